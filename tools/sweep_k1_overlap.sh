for cfg in "0 64 64" "1 64 64" "1 2 64" "1 2 3" "1 1 2" "1 3 4"; do
set -- $cfg
MTD_K1_OVERLAP=$1 MTD_K1_CAP320=$2 MTD_K1_CAP128=$3 python bench.py --steps 28 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('$cfg', round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['roofline']['stage_ms'].items()})"
done
