# 8-GPU bench of the K1 implementations given in $IMPLS (device-resident loop only)
set -u
O=gpurun_out; mkdir -p $O
G=${GPUS:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1"
for impl in ${IMPLS:-block}; do
  MTD_K1_IMPL=$impl timeout 600 $TR --master-port 29511 bench.py --gpus $G --no-e2e > $O/n${G}_bench_$impl.json 2> $O/n${G}_bench_$impl.err; echo "bench[$impl]=$?"
  python -c "
import json; j=json.load(open('$O/n${G}_bench_$impl.json')); r=j['roofline']; print('$impl', j['ms_per_step'], j['value'], r['stage_ms'], r['stage_ms_serial'])"
done
