#!/bin/bash
# A/B of environment switches on the device-resident bench at N GPUs (run on the box): bash tools/ab_n.sh <tag> <N> "ENV=.." ...
T=$1; G=$2; shift 2; O=gpurun_out; mkdir -p $O; i=0
for v in "$@"; do
  env $v timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port $((29520 + i)) bench.py --gpus $G --steps ${STEPS:-50} --warmup 7 --no-cpu-baseline --no-e2e > $O/${T}_$i.json 2> $O/${T}_$i.err || tail -n 3 $O/${T}_$i.err
  python - "$O/${T}_$i.json" "$v" <<'PY'
import json, sys
j = json.load(open(sys.argv[1])); r = j["roofline"]
f = lambda d: " ".join("%s=%.3f" % (k.split("_")[0] if k != "k3_apply" else "app", v) for k, v in d.items() if k != "h2d")
pr = j.get("per_rank_stage_ms") or {}
spread = " k1 per rank %.3f..%.3f" % (min(pr["k1_contacts"]), max(pr["k1_contacts"])) if pr else ""
print("%-44s %.3f ms/day host %.3f | pipe %s | serial %s |%s" % (sys.argv[2], j["ms_per_step"], j.get("host_submit_ms_per_step", 0), f(r["stage_ms"]), f(r["stage_ms_serial"]), spread))
PY
  i=$((i+1))
done
