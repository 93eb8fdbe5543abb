#!/usr/bin/env python
"""Aggregate the per-instruction stall samples of an ncu report by CUDA source line (no GUI needed).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [--top 30] [--kernel regex]

ncu's `--page source --csv` only prints the SASS view with metrics; the line mapping comes from nvdisasm -g on the
cubins of mtdcovabm_b200/libmtd_b200.so (compiled with -lineinfo).  Instruction order is identical in both listings.
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "mtdcovabm_b200", "libmtd_b200.so")


def sass_lines():
    """-> {mangled function name: [(offset, file, line, inlined_chain)]} in instruction order"""
    out = {}
    with tempfile.TemporaryDirectory() as td:
        subprocess.run(["cuobjdump", "-xelf", "all", SO], cwd=td, check=True, stdout=subprocess.DEVNULL)
        for f in sorted(os.listdir(td)):
            txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(td, f)], capture_output=True, text=True).stdout
            fn, cur = None, None
            for ln in txt.splitlines():
                m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
                if m:
                    fn = m.group(1); out[fn] = []; cur = None
                    continue
                m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
                if m:
                    cur = (os.path.basename(m.group(1)), int(m.group(2)))
                    continue
                m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);", ln)
                if m and fn:
                    out[fn].append((int(m.group(1), 16), cur, m.group(2)))
    return out


def main():
    rep = sys.argv[1]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 30
    kre = re.compile(sys.argv[sys.argv.index("--kernel") + 1]) if "--kernel" in sys.argv else None
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    sl = sass_lines()
    starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
    seen = set()
    for n, ki in enumerate(starts):
        name = rows[ki][1]
        if kre and not kre.search(name):
            continue
        if name in seen:
            continue
        seen.add(name)
        end = starts[n + 1] if n + 1 < len(starts) else len(rows)
        hdr = rows[ki + 1]
        si, ii = hdr.index("# Samples"), hdr.index("Instructions Executed")
        data = [r for r in rows[ki + 2:end] if len(r) > ii]
        # match the mangled function: same base name and same template integers, same instruction count
        base = re.match(r"(?:void )?(?:<unnamed>::)?(\w+)", name).group(1)
        ints = re.findall(r"\(int\)(\d+)", name)
        cands = [f for f in sl if base in f and all(("Li%sE" % x) in f for x in ints) and len(sl[f]) == len(data)]
        if not cands:
            cands = [f for f in sl if base in f and len(sl[f]) == len(data)]
        if not cands:
            print("## %s: no SASS match (%d instrs)" % (name, len(data)))
            continue
        ins = sl[cands[0]]
        agg = {}
        tot_s = tot_i = 0.0
        stall_cols = [j for j, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        for r, (off, loc, op) in zip(data, ins):
            s = float(r[si] or 0); i = float(r[ii] or 0)
            tot_s += s; tot_i += i
            a = agg.setdefault(loc, [0.0, 0.0, {}])
            a[0] += s; a[1] += i
            for j in stall_cols:
                v = float(r[j] or 0)
                if v:
                    a[2][hdr[j]] = a[2].get(hdr[j], 0) + v
        print("## %s   samples=%d  warp-instructions=%.3g" % (name, tot_s, tot_i))
        src_cache = {}
        for loc, (s, i, st) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
            text = ""
            if loc:
                p = os.path.join(ROOT, "mtdcovabm_b200", "csrc", loc[0])
                if p not in src_cache and os.path.exists(p):
                    src_cache[p] = open(p).read().splitlines()
                if p in src_cache and loc[1] - 1 < len(src_cache[p]):
                    text = src_cache[p][loc[1] - 1].strip()[:90]
            tops = ",".join("%s:%d" % (k.replace("stall_", ""), 100 * v / max(s, 1)) for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
            print("  %5.1f%% smp %5.1f%% ins  %-22s %-32s | %s" % (100 * s / max(tot_s, 1), 100 * i / max(tot_i, 1),
                                                              "%s:%d" % loc if loc else "?", tops, text))


if __name__ == "__main__":
    main()
