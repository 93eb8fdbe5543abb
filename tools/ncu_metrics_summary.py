#!/usr/bin/env python
"""Per-kernel means of every metric in an `ncu --metrics ... --csv` log, and totals per day:
    python tools/ncu_metrics_summary.py gpurun_out/q3_traffic.csv [--day-marker 'k1_block_kernel<1024, 16'] [--json out.json --intervals N]"""
import collections
import csv
import json
import sys


def load(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    ki, mi, vi, ii = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("ID")
    per = collections.OrderedDict()
    for r in rows[1:]:
        per.setdefault(int(r[ii]), {"kernel": r[ki]})[r[mi]] = float(r[vi].replace(",", ""))
    return list(per.values())


def short(k):
    k = k.replace("void ", "").replace("<unnamed>::", "")
    return k.split("(")[0]


def main():
    L = load(sys.argv[1])
    marker = sys.argv[sys.argv.index("--day-marker") + 1] if "--day-marker" in sys.argv else "k1_block_kernel<1024, 16"
    starts = [i for i, x in enumerate(L) if marker in x["kernel"]]
    n_days = len(starts) - 1                      # whole days only: from the first marker launch up to the last one
    sel = L[starts[0]:starts[-1]]
    per_day = len(sel) / n_days
    by_k = collections.OrderedDict()
    for x in sel:
        d = by_k.setdefault(short(x["kernel"]), collections.defaultdict(float))
        d["n"] += 1
        for m, v in x.items():
            if m != "kernel":
                d[m] += v
    tot = collections.defaultdict(float)
    print("mean per day over %d days (%.1f profiled launches each)" % (n_days, per_day))
    print("  %-44s %5s %8s %8s %8s %10s %6s %6s %6s %6s" % ("kernel", "n/day", "us", "rd MB", "wr MB", "warp-inst", "issue%", "warps%", "thr/in", "L2hit"))
    for k, d in by_k.items():
        n = d["n"]
        hit = d.get("lts__t_sectors_srcunit_tex_op_read_lookup_hit.sum", 0) / max(d.get("lts__t_sectors_srcunit_tex_op_read.sum", 1), 1)
        print("  %-44s %5.1f %8.1f %8.1f %8.1f %10.3e %6.1f %6.1f %6.1f %6.2f" % (
            k[:44], n / n_days, d["gpu__time_duration.sum"] / 1e3 / n_days, d["dram__bytes_read.sum"] / 1e6 / n_days,
            d["dram__bytes_write.sum"] / 1e6 / n_days, d["smsp__inst_executed.sum"] / n_days,
            d["smsp__issue_active.avg.pct_of_peak_sustained_active"] / n, d["sm__warps_active.avg.pct_of_peak_sustained_active"] / n,
            d["smsp__thread_inst_executed_per_inst_executed.ratio"] / n, hit))
        for m in ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum"):
            if k.startswith("k1_"):
                tot[m] += d[m] / n_days
    print("  K1 total per day: read %.1f MB + write %.1f MB = %.1f MB, warp-instr %.3e, serialised time %.1f us" % (
        tot["dram__bytes_read.sum"] / 1e6, tot["dram__bytes_write.sum"] / 1e6,
        (tot["dram__bytes_read.sum"] + tot["dram__bytes_write.sum"]) / 1e6, tot["smsp__inst_executed.sum"],
        tot["gpu__time_duration.sum"] / 1e3))
    if "--json" in sys.argv:
        out = sys.argv[sys.argv.index("--json") + 1]
        iv = float(sys.argv[sys.argv.index("--intervals") + 1])
        json.dump({"source": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,... -k regex:k1_|k2_ "
                             "(the K1 launches of whole days of the bench; mean per day), " + sys.argv[1],
                   "kernel": "k1_contacts (%.1f profiled launches per day incl. K2 and the state kernels; K1 = the k1_* kernels)" % per_day, "days": n_days,
                   "dram_bytes_read": tot["dram__bytes_read.sum"], "dram_bytes_write": tot["dram__bytes_write.sum"],
                   "dram_bytes_total": tot["dram__bytes_read.sum"] + tot["dram__bytes_write.sum"],
                   "warp_instructions": tot["smsp__inst_executed.sum"], "intervals_of_day": iv}, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()
