#!/usr/bin/env python
"""SASS census of libmtd_b200.so per kernel (cuobjdump -sass): instruction count and the mnemonics that show what the
kernel is made of -- bulk-async copies (UBLKCP = cp.async.bulk), mbarrier (SYNCS), warp votes / matches / shuffles /
reductions, shared and global atomics, fp64 arithmetic, MUFU -- plus registers and shared memory from cuobjdump -res-usage.
    python tools/sass_summary.py > profiles/r02_sass_summary.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "mtdcovabm_b200", "libmtd_b200.so")
GROUPS = collections.OrderedDict([
    ("bulk-async copy (UBLKCP)", r"^UBLKCP"), ("mbarrier (SYNCS)", r"^SYNCS"), ("tcgen05 / UTC*", r"^UTC|^TCGEN"),
    ("vote / match (VOTE, MATCH)", r"^VOTE|^MATCH"), ("shuffle (SHFL)", r"^SHFL"), ("warp reduce (REDUX)", r"^REDUX"),
    ("popc / flo / brev", r"^POPC|^FLO|^BREV"), ("shared atomics (ATOMS)", r"^ATOMS"), ("global atomics (ATOMG, RED)", r"^ATOMG|^RED\b|^ATOM\b"),
    ("fp64 (DADD, DMUL, DFMA, DSETP, ...)", r"^D(ADD|MUL|FMA|SETP|MNMX)"), ("MUFU", r"^MUFU"), ("LDG", r"^LDG"), ("STG", r"^STG"),
    ("LDS", r"^LDS"), ("STS", r"^STS"), ("BAR", r"^BAR"), ("IMAD / integer mul", r"^IMAD|^IMUL"),
])


def demangle(n):
    try:
        return subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
    except Exception:
        return n


def main():
    res = subprocess.run(["cuobjdump", "-res-usage", SO], capture_output=True, text=True).stdout
    usage = {}
    cur = None
    for ln in res.splitlines():
        m = re.match(r"\s*Function (\S+):", ln)
        if m:
            cur = m.group(1)
            continue
        m = re.search(r"REG:(\d+).*?SHARED:(\d+)", ln)
        if m and cur:
            usage[cur] = (int(m.group(1)), int(m.group(2)))
    sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for ln in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = m.group(1); kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", ln)
        if m and cur:
            kernels[cur][m.group(1).split(".")[0]] += 1
    print("SASS census of mtdcovabm_b200/libmtd_b200.so (sm_100a), one line per kernel: static instruction counts")
    print("no kernel of this path holds a contraction, so there is no tcgen05 / TMEM instruction; the Blackwell-specific part is the")
    print("bulk-async (TMA engine) staging of the CSR segments in the block-per-cell K1 kernels (UBLKCP + SYNCS = cp.async.bulk + mbarrier)\n")
    for k, cnt in kernels.items():
        name = re.sub(r"\(anonymous namespace\)::", "", demangle(k)).split("(")[0].replace("void ", "")
        tot = sum(cnt.values())
        parts = []
        for g, rx in GROUPS.items():
            n = sum(v for op, v in cnt.items() if re.match(rx, op))
            if n:
                parts.append("%s %d" % (g.split(" (")[0], n))
        reg, sh = usage.get(k, (0, 0))
        print("%-58s %6d instr  %3d regs %6d B smem | %s" % (name[:58], tot, reg, sh, ", ".join(parts)))


if __name__ == "__main__":
    main()
