#!/usr/bin/env python
"""Times the UNMODIFIED reference's own CPU path for this hot path -- contactnetwork_mp.contactnetwork_parallel
(contactnetwork_mp.py:16; called by sim.py:1759 with the pool that sim.py:937-943 builds) -- inside a normal
sim.main() run of config 1 (population/1kagents2ktourists2019, data/epidemiology.json) or of the tiled 10k population,
with `numprocesses` = 1 and = os.cpu_count().  time.perf_counter() around every call; the pairs a day holds are counted
from vars_util.cells_agents_timesteps before the call (sum over cells of (m^2 - sum_a cnt_a^2) / 2: the metric's unit).

Build container only (/root/reference).  Writes / updates profiles/r02_reference_python_timing.json, which bench.py
carries in its JSON line as cpu_baseline.reference_python (the reference's sources do not travel to the GPU box).

    python tools/time_reference.py [--pop 1k|10k] [--days 12] [--procs 1,8]
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))


def one(pop, days, procs, out_path):
    import ref_harness as rh
    rh._install_stubs()
    if pop == "10k":
        import make_pop10k
        if not os.path.exists("/tmp/pop10k/agents.json"):
            make_pop10k.tile("/tmp/pop10k", 10)
        wd = rh.prepare_workdir("/tmp/reftime_%s_%d" % (pop, procs), popsubfolder="tiled10k", pop_src="/tmp/pop10k")
        sub = "tiled10k"
    else:
        wd = rh.prepare_workdir("/tmp/reftime_%s_%d" % (pop, procs))
        sub = "1kagents2ktourists2019"
    simdir = os.path.join(wd, "simulator")
    sys.path.insert(0, simdir)
    os.chdir(wd)
    import random
    import numpy as np
    random.seed(6); np.random.seed(6)
    import sim
    import contactnetwork_mp as cmp
    sim.params.update({"popsubfolder": sub, "simulationdays": days, "numprocesses": procs, "use_mp": True,
                       "logmemoryinfo": False, "logfilename": "timing_%d.txt" % procs})
    rows = []
    orig = cmp.contactnetwork_parallel

    def timed(manager, pool, day, weekday, n_locals, n_tourists, ratio, agents_epi, vars_util, *a, **k):
        pairs = m_total = 0
        agents = set()
        for lst in vars_util.cells_agents_timesteps.values():
            m = len(lst); m_total += m
            cnt = {}
            for (ag, s, e) in lst:
                cnt[ag] = cnt.get(ag, 0) + 1
                agents.add(ag)
            pairs += (m * m - sum(v * v for v in cnt.values())) // 2
        t0 = time.perf_counter()
        out = orig(manager, pool, day, weekday, n_locals, n_tourists, ratio, agents_epi, vars_util, *a, **k)
        dt = time.perf_counter() - t0
        rows.append({"day": int(day), "seconds": dt, "pairs": int(pairs), "intervals": int(m_total), "active_agents": len(agents)})
        return out

    cmp.contactnetwork_parallel = timed
    sim.contactnetwork_mp.contactnetwork_parallel = timed
    stdout = sys.stdout
    try:
        sim.main()
    finally:
        sys.stdout = stdout
    json.dump(rows, open(out_path, "w"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pop", default="1k")
    ap.add_argument("--days", type=int, default=12)
    ap.add_argument("--procs", default="1,%d" % (os.cpu_count() or 1))
    ap.add_argument("--one", type=int, default=0)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    if a.one:
        one(a.pop, a.days, a.one, a.out)
        return
    path = os.path.join(ROOT, "profiles", "r02_reference_python_timing.json")
    res = json.load(open(path)) if os.path.exists(path) else {}
    for p in [int(x) for x in a.procs.split(",")]:
        tmp = "/tmp/reftime_%s_%d.json" % (a.pop, p)
        subprocess.check_call([sys.executable, os.path.abspath(__file__), "--pop", a.pop, "--days", str(a.days), "--one", str(p),
                               "--out", tmp], stdout=subprocess.DEVNULL)
        rows = json.load(open(tmp))[1:]                 # the first day carries one-off set-up
        sec = sum(r["seconds"] for r in rows); pairs = sum(r["pairs"] for r in rows)
        res["%s_procs%d" % (a.pop, p)] = {
            "population": "1k agents + 2k tourists (config 1)" if a.pop == "1k" else "10k agents + 20k tourists (1k population tiled x10)",
            "call": "contactnetwork_mp.contactnetwork_parallel (contactnetwork_mp.py:16) inside sim.main(), numprocesses=%d" % p,
            "processes": p, "host_cores": os.cpu_count(), "days_timed": len(rows), "seconds": sec, "pairs": pairs,
            "pairs_per_s": pairs / sec, "agent_days_per_s": sum(r["active_agents"] for r in rows) / sec,
            "where": "build container (the reference's sources do not travel to the GPU box)",
            "python": sys.version.split()[0]}
        print(p, res["%s_procs%d" % (a.pop, p)], flush=True)
        json.dump(res, open(path, "w"), indent=1)


if __name__ == "__main__":
    main()
