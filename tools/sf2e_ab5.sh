# weight rows of the dual reverse contraction from shared memory (2) or the constant bank (1), C4 Euler row
for r in 2 1 2 1; do
  GPFLOW_NVCC_EXTRA="-DGPF_SF2E_RBIT=$r" python -m gpflow_b200.build --force > /dev/null 2>&1
  python bench.py --rows c4_2d_euler_5e-4 --no-cpu-baseline --no-e2e --steps 5 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1])
for r in j['rows']: print('rbit $r', r['value'], r['ms_per_step'])"
done
