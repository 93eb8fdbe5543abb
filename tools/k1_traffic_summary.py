#!/usr/bin/env python
"""Per-kernel and per-day DRAM traffic of the K1 launches from an `ncu --metrics dram__bytes_read.sum,... --csv` log:
    python tools/k1_traffic_summary.py gpurun_out/o_k1_new.csv [--json out.json --intervals N]"""
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


def main():
    L = load(sys.argv[1])
    # a day's K1 stage opens with the <1024> launch; keep whole days only
    starts = [i for i, x in enumerate(L) if "k1_large_kernel<1024" in x["kernel"]]
    n_days = len(starts) - 1 if len(L) - starts[-1] < starts[1] - starts[0] else len(starts)
    per_day = starts[1] - starts[0]
    by_k = collections.defaultdict(lambda: collections.defaultdict(float))
    for x in L[starts[0]:starts[0] + n_days * per_day]:
        k = x["kernel"].split("::")[-1].split("(")[0]
        for m, v in x.items():
            if m != "kernel":
                by_k[k][m] += v / n_days
    tot = collections.defaultdict(float)
    print("mean per day over %d days (%d K1 launches each)" % (n_days, per_day))
    for k, d in by_k.items():
        hit = d.get("lts__t_sectors_srcunit_tex_op_read_lookup_hit.sum", 0) / max(d.get("lts__t_sectors_srcunit_tex_op_read.sum", 1), 1)
        print("  %-28s read %7.1f MB  write %7.1f MB  warp-instr %.3e  time %7.1f us  L2 read hit %.2f" % (
            k, d["dram__bytes_read.sum"] / 1e6, d["dram__bytes_write.sum"] / 1e6, d["smsp__inst_executed.sum"],
            d["gpu__time_duration.sum"] / 1e3, hit))
        for m, v in d.items():
            tot[m] += v
    print("  total: read %.1f MB write %.1f MB = %.1f MB/day, warp-instr %.3e, serialised time %.1f us" % (
        tot["dram__bytes_read.sum"] / 1e6, tot["dram__bytes_write.sum"] / 1e6,
        (tot["dram__bytes_read.sum"] + tot["dram__bytes_write.sum"]) / 1e6, tot["smsp__inst_executed.sum"],
        tot["gpu__time_duration.sum"] / 1e3))
    if "--json" in sys.argv:
        out = sys.argv[sys.argv.index("--json") + 1]
        iv = float(sys.argv[sys.argv.index("--intervals") + 1])
        json.dump({"source": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum -k regex:k1_ "
                             "(the K1 launches of whole days of the bench; mean per day), " + sys.argv[1],
                   "kernel": "k1_contacts (%d launches per day: 6 size-class kernels + ordering helpers)" % per_day, "days": n_days,
                   "dram_bytes_read": tot["dram__bytes_read.sum"], "dram_bytes_write": tot["dram__bytes_write.sum"],
                   "dram_bytes_total": tot["dram__bytes_read.sum"] + tot["dram__bytes_write.sum"],
                   "warp_instructions": tot["smsp__inst_executed.sum"], "intervals_of_day": iv}, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()
