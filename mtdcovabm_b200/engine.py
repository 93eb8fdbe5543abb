"""Tensor-native day engine: the per-day sequence of the hot path over device-resident (or pinned-host) arrays.

One engine = one rank = one GPU = one `capi.Handle` on one torch stream.  The order of a day follows the reference's
day loop (sim.py:1392-2125): SEIR advance (seirstateutil.update_agent_state runs inside the itinerary stage at the
start of the day, itinerary.py:399) -> contact network (contactnetwork.py:65) -> transmission (epidemiology.py:88).
With `world > 1` cells are sharded over the ranks and the per-agent exposure deltas are exchanged once per day
(the reference's stateful-actor strategy, actor_dist.py:491-518): K3 writes this rank's records
straight into its slice of the exchange buffer and PUSHES exactly those records into every peer's buffer over NVLink
peer memory (CUDA IPC mappings, arrival flags, no host synchronisation; csrc/k3_transmission.cu); every rank applies
all records to its replica of the per-agent state on a side stream while the next day's K0/K1 already run.  If peer
mappings cannot be established (or MTD_EXCHANGE=nccl) the records go through one NCCL all-gather per day instead.

torch is used for device memory, streams and torch.distributed only; every computation is a kernel of libmtd_b200.so.
"""
from __future__ import annotations

import os

import torch

from . import capi


def exchange_deltas(gather: torch.Tensor, rank: int, group=None):
    """one all-gather per day of the fixed-capacity delta slices: gather[r] <- rank r's records (in place: every
    rank contributes gather[rank]).  Works on CUDA tensors over NCCL and on CPU tensors over gloo (tests)."""
    torch.distributed.all_gather_into_tensor(gather.view(-1), gather[rank].reshape(-1).clone(), group=group)


def setup_peer_exchange(h, rank, world, delta_cap, group=None):
    """map every peer's exchange buffer (CUDA IPC) -> "p2p"; "nccl" if that is not possible on this box"""
    import sys
    dist = torch.distributed
    ok, mine, why = 1, b"", ""
    if os.environ.get("MTD_EXCHANGE", "p2p") == "nccl":
        ok, why = 0, "MTD_EXCHANGE=nccl"
    else:
        try:
            mine = h.exchange_init(rank, world, delta_cap)
        except RuntimeError as e:
            ok, why = 0, str(e)
    handles = [None] * world
    dist.all_gather_object(handles, (ok, mine, why), group=group)
    if all(o for o, _, _ in handles):
        try:
            for r, (_, hb, _) in enumerate(handles):
                if r != rank:
                    h.exchange_open(r, hb)
            h.exchange_ready()
        except RuntimeError as e:
            ok, why = 0, str(e)
    else:
        ok = 0
        why = why or next(w for o, _, w in handles if not o)
    flags = [None] * world
    dist.all_gather_object(flags, ok, group=group)        # also the barrier after the mappings
    if all(flags):
        return "p2p"
    if rank == 0:
        sys.stderr.write("mtdcovabm_b200: peer exchange unavailable (%s); using the NCCL all-gather\n" % why)
    return "nccl"


class DayEngine:
    def __init__(self, n_agents, n_cells, max_intervals, params, device=0, rank=0, world=1, delta_cap=0,
                 max_rows=0, mode=capi.MODE_PHILOX, seed=0, group=None):
        if not torch.cuda.is_available():
            raise RuntimeError("mtdcovabm_b200 needs a CUDA device: there is no CPU fallback for this path")
        self.device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        self.rank, self.world, self.group = rank, world, group
        self.mode, self.seed = mode, seed
        # high priority: in the day pipeline the short K2/K3 kernels of the current day (this stream) get the SM slots that
        # free up before the next persistent K1 kernel of the coming day does (MTD_MAIN_PRIO=0: default priority)
        self.stream = torch.cuda.Stream(self.device, priority=0 if os.environ.get("MTD_MAIN_PRIO") == "0" else -1)
        # exposure records of one day (80 B each): sized for an epidemic peak of ~6 % of all ids in one day; only the
        # records actually produced are exchanged, so a generous capacity costs memory, not bandwidth
        self.delta_cap = int(delta_cap) if delta_cap else max(4096, n_agents // 16)
        self.h = capi.Handle(n_agents, n_cells, max_intervals, device=self.device.index or 0,
                             stream=self.stream.cuda_stream, max_rows=max_rows, max_exposures=self.delta_cap)
        self.h.set_params(params)
        self.gather = None
        self.exchange = "none"
        if world > 1:
            self.exchange = self._setup_peer_exchange()
            if self.exchange == "nccl":
                with torch.cuda.device(self.device):
                    self.gather = torch.zeros((world, self.delta_cap + 1, capi.DELTA_WORDS), dtype=torch.int32,
                                             device=self.device)
        self.n_agents, self.n_cells = n_agents, n_cells
        self._vaccination = False

    def _setup_peer_exchange(self):
        return setup_peer_exchange(self.h, self.rank, self.world, self.delta_cap, self.group)

    def _transmission(self, d):
        """K3 of the day + the exchange and application of the exposure records"""
        h = self.h
        if self.world == 1:
            h.simulate_direct_contacts(d)
        elif self.exchange == "p2p":
            h.simulate_direct_contacts(d, apply=capi.APPLY_EXCHANGE)
        else:       # this rank's cells -> delta records -> one all-gather -> every rank applies all records
            h.simulate_direct_contacts(d, delta_buf=self.gather[self.rank], cap_records=self.delta_cap, apply=False)
            exchange_deltas(self.gather, self.rank, self.group)
            h.apply_deltas(self.gather, self.world, self.delta_cap)

    # ---- population
    def set_population(self, pop):
        """pop: object with device tensors cell_type, cell_indid, age_bracket, epi_age_bracket, soc_rate, res_cellid,
        seir, inf_type, inf_sev, sched, sched_cnt (e.g. synth.SynthPopulation)"""
        D = capi.MEM_DEVICE
        with torch.cuda.stream(self.stream):
            self.h.set_cells(pop.cell_type, pop.cell_indid, mem=D)
            self.h.set_static(pop.age_bracket, pop.epi_age_bracket, pop.soc_rate, pop.res_cellid, mem=D)
            self.h.set_state(seir=pop.seir, inf_type=pop.inf_type, inf_sev=pop.inf_sev, sched=pop.sched,
                             sched_cnt=pop.sched_cnt, mem=D)
        self.h.sync()

    def set_inputs_ready(self, on=True):
        """Promise that the device interval arrays given to step() are complete when step() is called (pre-built day
        streams, nothing pending on any torch stream writes them): the library then runs K0 and K1 of the coming days
        beside K2/K3 of the current one (mtd_set_k0_overlap).  step_host() needs no promise: the library owns its uploads."""
        self.h.set_k0_overlap(on)

    # ---- one day
    def step(self, day, s, mhd=0.0, testing=False, quarantine=False, mem=capi.MEM_DEVICE, n=None):
        """s: dict(iv_agent, iv_cell, iv_start, iv_end, iv_order) -- device tensors (mem=MEM_DEVICE) or pinned host
        tensors / numpy arrays (mem=MEM_HOST).  Enqueues the whole day on the engine's stream; returns nothing."""
        h = self.h
        d = h.make_day(self.mode, day, mhd, testing, quarantine, seed=self.seed)
        with torch.cuda.stream(self.stream):
            h.advance_seir(day, None)
            h.simulate_contact_network(d, s["iv_agent"], s["iv_cell"], s["iv_start"], s["iv_end"],
                                       iv_order=s.get("iv_order"), n=n if n is not None else len(s["iv_agent"]),
                                       mem=mem)
            self._transmission(d)

    # ---- host boundary, pipelined: pinned host buffers in, pinned host buffers out, copies beside the kernels
    def enable_host_pipeline(self, cap_rows, cap_exposed=None, export_rows=True):
        """two sets of pinned result buffers (day d's results stay valid until day d+2 is submitted).  export_rows=False:
        the day's contact rows stay in HBM (a caller that runs contact tracing on the device, mtd_contact_tracing, never
        needs them on the host) and only the counts, the occupancy and the exposed ids are exported."""
        import numpy as np
        cap_exposed = int(cap_exposed or self.delta_cap)
        self.h.set_export(True)
        self._res = []
        self._export_rows = bool(export_rows)
        for _ in range(2):
            rows = torch.empty((int(cap_rows) if export_rows else 1, 3), dtype=torch.int32).pin_memory()
            kv = torch.empty((cap_exposed, 2), dtype=torch.int64).pin_memory()
            res = torch.zeros(capi.RESULT_WORDS, dtype=torch.int64).pin_memory()
            self._res.append((rows, kv, res))
        self._cap_rows, self._cap_exposed = int(cap_rows), cap_exposed
        self._k = 0
        self._staged = None

    def step_host(self, day, s, next_s=None, mhd=0.0, testing=False, quarantine=False):
        """One day from pinned HOST interval arrays `s`; `next_s` (the following day's arrays, if already known) is
        uploaded by the copy stream while this day computes.  Returns a HostDayResult whose buffers are filled
        asynchronously by the export stream: call .wait() before reading them."""
        h = self.h
        d = h.make_day(self.mode, day, mhd, testing, quarantine, seed=self.seed)
        n = len(s["iv_agent"])
        with torch.cuda.stream(self.stream):
            h.advance_seir(day, None)
            if self._staged is s:
                h.simulate_contact_network(d, None, None, None, None, n=n, mem=capi.MEM_STAGED)
            else:
                h.simulate_contact_network(d, s["iv_agent"], s["iv_cell"], s["iv_start"], s["iv_end"],
                                           iv_order=s.get("iv_order"), n=n, mem=capi.MEM_HOST)
            self._staged = None
            if next_s is not None:
                h.prefetch_intervals(next_s["iv_agent"], next_s["iv_cell"], next_s["iv_start"], next_s["iv_end"],
                                     next_s.get("iv_order"))
                self._staged = next_s
            self._transmission(d)
            rows, kv, res = self._res[self._k]
            self._k ^= 1
            ticket = h.export_day_async(rows if self._export_rows else None, self._cap_rows if self._export_rows else 0, kv,
                                        self._cap_exposed, res, mem=capi.MEM_HOST)
        return HostDayResult(h, ticket, rows if self._export_rows else None, kv, res)

    # ---- checkpoint / resume (the reference has none; the whole dynamic state of the path is a
    # handful of dense per-agent arrays, so a checkpoint is their copy plus the day and the Philox seed)
    def save_checkpoint(self, path, day):
        """state after `day` has been stepped -> one .npz (every rank holds the same replicated state: rank 0 writes)"""
        import numpy as np
        self.h.sync()
        st = self.h.get_state()
        last, rec = self.h.get_vaccination(with_records=self._vaccination)
        if self.rank == 0:
            extra = {"vacc_day": rec} if rec is not None else {}
            np.savez_compressed(path, day=np.int64(day), seed=np.int64(self.seed), mode=np.int64(self.mode),
                                n_agents=np.int64(self.n_agents), last_dose_day=last, **st, **extra)

    def load_checkpoint(self, path):
        """-> the day the checkpoint was taken after; static attributes / cells / params are set as for a new run"""
        import numpy as np
        z = np.load(path)
        if int(z["n_agents"]) != self.n_agents:
            raise RuntimeError("checkpoint is for %d agents, engine has %d" % (int(z["n_agents"]), self.n_agents))
        self.seed, self.mode = int(z["seed"]), int(z["mode"])
        self.h.set_state(seir=z["seir"], inf_type=z["inf_type"], inf_sev=z["inf_sev"], last_dose_day=z["last_dose_day"],
                         sched=z["sched"], sched_cnt=z["sched_cnt"], test_day=z["test_day"],
                         test_result_day=z["test_result_day"], quarantine=z["quarantine"], hosp=z["hosp"])
        if "vacc_day" in z.files:
            if not self._vaccination:
                raise RuntimeError("checkpoint carries vaccination records: enable_vaccination first")
            self.h.set_vaccination(z["last_dose_day"], z["vacc_day"])
        self.h.sync()
        return int(z["day"])

    def enable_vaccination(self, n_locals):
        self.h.enable_vaccination(n_locals)
        self._vaccination = True

    def sync(self):
        self.h.sync()

    def close(self):
        self.h.close()


class HostDayResult:
    """Results of one day in pinned host memory (written by the device over PCIe while the next day computes)."""

    def __init__(self, h, ticket, rows, kv, res):
        self._h, self.ticket, self._rows, self._kv, self._res = h, ticket, rows, kv, res
        self._done = False

    def wait(self):
        if not self._done:
            self._h.export_wait(self.ticket)
            self._done = True
            if int(self._res[11]) != self.ticket:
                raise RuntimeError("export record carries ticket %d, expected %d" % (int(self._res[11]), self.ticket))
            if int(self._res[10]) != 0:
                raise RuntimeError("device error %d during the day step" % int(self._res[10]))
        return self

    @property
    def n_rows(self):
        return int(self.wait()._res[1])

    @property
    def n_exposed(self):
        return int(self.wait()._res[2])

    @property
    def occupancy(self):
        return self.wait()._res[3:9].numpy()

    @property
    def pairs_evaluated(self):
        return int(self.wait()._res[9])

    @property
    def rows(self):
        """structured view (capi.ROW12) of the contact rows, in the reference's row order"""
        n = self.n_rows
        if self._rows is None:
            raise RuntimeError("this host pipeline keeps the contact rows on the device (export_rows=False)")
        return self._rows.numpy()[:n].view(capi.ROW12).reshape(n)

    def exposed(self, reference_order=True):
        """exposed agent ids; reference_order sorts by the (cell order, row order) key"""
        import numpy as np
        kv = self._kv.numpy()[:self.n_exposed]
        if reference_order:
            kv = kv[np.argsort(kv[:, 0].view(np.uint64), kind="stable")]
        return kv[:, 1].astype(np.int32)

    @property
    def nbytes(self):
        return (12 * self.n_rows if self._rows is not None else 0) + 16 * self.n_exposed + 8 * capi.RESULT_WORDS
