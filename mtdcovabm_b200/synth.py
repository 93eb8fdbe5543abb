"""Tiled synthetic populations for the throughput configurations (BASELINE.json configs 2-4).

The reference's Python itinerary stage cannot feed 520k / 5M agents (about 0.74 s per 1k agents per day), and the
10k population files are missing from the reference checkout, so the large configurations replicate the recorded
interval streams of the reference's own 1k-agent run (`data/itinerary_bank_1k.npz`, written by
tests/golden/make_bank.py) `reps` times: replica r gets agent ids `+ r*N1` and cell ids `+ r*C1`, which keeps the
measured cell-size distribution exactly (p50 6, p99 117, max ~280 intervals per cell).  Agent ids are then
shuffled over the whole id space (colleagues do not have neighbouring ids in a real population) and the stream is put
in agent-major emission order, the order in which the itinerary stage emits it (itinerary.py:1403-1452).

Everything here is host/torch plumbing that prepares INPUTS; no part of the hot path lives in this file.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from .params import DATA_DIR, NONE

T_ASYMP_REC = 5


def load_bank(path=None):
    z = np.load(path or os.path.join(DATA_DIR, "itinerary_bank_1k.npz"))
    bank = {k: z[k] for k in z.files}
    bank["n_days"] = len(bank["days"])
    return bank


def cell_owner(cell: torch.Tensor, world: int) -> torch.Tensor:
    """static owner map cell -> rank (multiplicative hash: tail cells spread evenly)"""
    if world == 1:
        return torch.zeros_like(cell)
    h = (cell.to(torch.int64) * 2654435761) & 0xFFFFFFFF
    return ((h >> 7) % world).to(cell.dtype)


def balanced_cell_owner(weight: np.ndarray, world: int, exact: int = 200_000) -> np.ndarray:
    """Static owner map cell -> rank that balances the expected K1 cost of the ranks (the reference
    balances its partitions the same way, greedily on a min-heap, with the member count as the weight, util.py:675-707).
    weight[c] = expected cost of cell c (sum over the day streams of m^2 + 16 m, m = intervals of the cell that day).
    Longest-processing-time-first: the `exact` heaviest cells go, heaviest first, to the currently lightest rank (heap);
    the long tail of light cells is dealt out in snake order of weight.  Deterministic; every rank computes the same map."""
    import heapq
    C = len(weight)
    owner = np.zeros(C, np.int32)
    if world == 1:
        return owner
    order = np.argsort(-weight, kind="stable")
    k = int(min(exact, C))
    heap = [(0.0, r) for r in range(world)]
    w = weight.astype(np.float64)
    for c in order[:k]:
        load, r = heapq.heappop(heap)
        owner[c] = r
        heapq.heappush(heap, (load + w[c], r))
    if k < C:
        rest = order[k:]
        pos = np.arange(len(rest)) % (2 * world)
        snake = np.where(pos < world, pos, 2 * world - 1 - pos)
        by_load = np.array([r for _, r in sorted(heap)], np.int32)      # lightest rank first
        owner[rest] = by_load[snake]
    return owner


class SynthPopulation:
    """`reps` replicas of the bank population on `device` (all ranks build the identical population)."""

    def __init__(self, bank, reps, device, seed=0, shuffle_ids=True, infectious_frac=0.01, start_day=1):
        self.bank, self.reps, self.device = bank, int(reps), torch.device(device)
        self.N1 = int(bank["n_agents"][0]); self.C1 = len(bank["cell_type"])
        self.n_agents = self.N1 * self.reps; self.n_cells = self.C1 * self.reps
        self.n_days = int(bank["n_days"])
        gen = torch.Generator().manual_seed(1234 + seed)
        N, dev = self.n_agents, self.device
        self.perm = torch.randperm(N, generator=gen).to(dev) if shuffle_ids else None

        def tile(a, dtype):
            t = torch.from_numpy(np.ascontiguousarray(a)).to(dev).to(dtype).repeat(self.reps)
            if self.perm is None:
                return t
            out = torch.empty_like(t)
            out[self.perm] = t
            return out

        self.age_bracket = tile(bank["age_bracket"], torch.int8)
        self.epi_age_bracket = tile(bank["epi_age_bracket"], torch.int8)
        self.soc_rate = tile(bank["soc_rate"], torch.float64)
        rc = torch.from_numpy(bank["res_cellid"].astype(np.int64)).to(dev).repeat(self.reps)
        rc = torch.where(rc >= 0, rc + torch.arange(self.reps, device=dev).repeat_interleave(self.N1) * self.C1, rc)
        if self.perm is not None:
            t = torch.empty_like(rc); t[self.perm] = rc; rc = t
        self.res_cellid = rc.to(torch.int32)
        self.cell_type = torch.from_numpy(bank["cell_type"]).to(dev).repeat(self.reps)
        self.cell_indid = torch.from_numpy(bank["cell_indid"]).to(dev).repeat(self.reps)
        present = tile(bank["present"], torch.bool)
        self.n_present = int(present.sum())
        u = torch.rand(N, generator=gen).to(dev)
        inf = present & (u < infectious_frac)
        self.seir = torch.where(inf, 3, torch.where(present, 1, 0)).to(torch.int8)
        self.inf_type = torch.where(inf, 3, 0).to(torch.int8)          # Asymptomatic
        self.inf_sev = torch.where(inf, 1, 0).to(torch.int8)           # Mild
        # the initially infectious recover 2..9 days in: one AsymptomaticToRecovered entry each
        rec_day = start_day + 2 + (torch.rand(N, generator=gen).to(dev) * 8).to(torch.int32)
        self.sched = torch.zeros((N, 12), dtype=torch.int32, device=dev)
        self.sched[:, 0] = torch.where(inf, (rec_day << 16) | (T_ASYMP_REC << 8) | 60, 0)
        self.sched_cnt = inf.to(torch.uint8)

    def owner_map(self, world):
        """cell -> rank, balanced on the cells' expected cost over the bank's day streams (balanced_cell_owner);
        MTD_OWNER=hash switches back to the multiplicative hash"""
        key = ("owner", world)
        if key not in self.__dict__:
            if os.environ.get("MTD_OWNER", "balanced") == "hash":
                self.__dict__[key] = cell_owner(torch.arange(self.n_cells, device=self.device, dtype=torch.int32), world)
            else:
                b = self.bank
                w1 = np.zeros(self.C1, np.float64)
                for k in range(self.n_days):
                    m = np.bincount(b["d%d_iv_cell" % k].astype(np.int64), minlength=self.C1).astype(np.float64)
                    w1 += m * m + 16.0 * m
                owner = balanced_cell_owner(np.tile(w1, self.reps), world)
                self.__dict__[key] = torch.from_numpy(owner).to(self.device)
        return self.__dict__[key]

    def stream(self, k, rank=0, world=1):
        """bank day k tiled -> dict(iv_agent, iv_cell, iv_start, iv_end, iv_order, n_total) of this rank's shard"""
        b, dev = self.bank, self.device
        k = k % self.n_days
        a1 = torch.from_numpy(b["d%d_iv_agent" % k].astype(np.int64)).to(dev)
        c1 = torch.from_numpy(b["d%d_iv_cell" % k].astype(np.int64)).to(dev)
        s1 = torch.from_numpy(b["d%d_iv_start" % k]).to(dev); e1 = torch.from_numpy(b["d%d_iv_end" % k]).to(dev)
        M1 = len(a1)
        r = torch.arange(self.reps, device=dev).repeat_interleave(M1)
        agent = a1.repeat(self.reps) + r * self.N1
        cell = (c1.repeat(self.reps) + r * self.C1).to(torch.int32)
        start, end = s1.repeat(self.reps), e1.repeat(self.reps)
        del r
        if self.perm is not None:
            agent = self.perm[agent]
            agent, o = torch.sort(agent, stable=True)          # agent-major emission order, per-agent order kept
            cell, start, end = cell[o], start[o], end[o]
            del o
        agent = agent.to(torch.int32)
        M = agent.numel()
        order = None
        if world > 1:
            sel = self.owner_map(world)[cell.long()] == rank
            order = torch.nonzero(sel).flatten().to(torch.int32)   # read as uint32 by the C ABI
            agent, cell, start, end = agent[sel], cell[sel], start[sel], end[sel]
        return dict(iv_agent=agent.contiguous(), iv_cell=cell.contiguous(), iv_start=start.contiguous(),
                    iv_end=end.contiguous(), iv_order=order, n_total=M)
