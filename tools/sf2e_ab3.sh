# fully unrolled dual contraction loops: weight rows from shared memory (WBIT 2) or the constant bank (WBIT 1); packed element-wise ops again
for cfg in "15 2 0" "15 1 0" "15 2 1" "15 1 0"; do
  set -- $cfg
  GPFLOW_NVCC_EXTRA="-DGPF_SF2E_UNROLL=$1 -DGPF_SF2E_WBIT=$2 -DGPF_SF2E_PACKED=$3" python -m gpflow_b200.build --force > /dev/null 2>&1
  python bench.py --rows c4_2d_euler_5e-4 --no-cpu-baseline --no-e2e --steps 5 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1])
for r in j['rows']: print('unroll $1 wbit $2 packed $3', r['value'], r['ms_per_step'])"
done
