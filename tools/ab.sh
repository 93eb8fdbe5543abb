#!/bin/bash
# A/B of environment switches on the device-resident bench (run on the box):  bash tools/ab.sh <tag> <agents> "ENV1=a ENV2=b" "ENV1=c" ...
# prints ms/day and the pipelined / serial stage times per variant; logs under gpurun_out/<tag>_<i>.json
T=$1; A=$2; shift 2; O=gpurun_out; mkdir -p $O; i=0
for v in "$@"; do
  env $v timeout 300 python bench.py ${BENCH_ARGS:-} --agents $A --steps ${STEPS:-28} --warmup 7 --no-cpu-baseline --no-e2e > $O/${T}_$i.json 2> $O/${T}_$i.err || tail -n 3 $O/${T}_$i.err
  python - "$O/${T}_$i.json" "$v" <<'PY'
import json, sys
j = json.load(open(sys.argv[1])); r = j["roofline"]
f = lambda d: " ".join("%s=%.3f" % (k.split("_")[0] if k != "k3_apply" else "app", v) for k, v in d.items() if k != "h2d")
print("%-44s %.3f ms/day | pipe %s | serial %s" % (sys.argv[2], j["ms_per_step"], f(r["stage_ms"]), f(r["stage_ms_serial"])))
PY
  i=$((i+1))
done
