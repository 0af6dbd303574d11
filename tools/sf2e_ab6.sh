# layer-1 weight rows of the dual reverse sweep from shared memory (4) or the constant bank (1), C4 Euler row
for r in 4 1 4 1; do
  GPFLOW_NVCC_EXTRA="-DGPF_SF2E_W1BIT=$r" python -m gpflow_b200.build --force > /dev/null 2>&1
  python bench.py --rows c4_2d_euler_5e-4 --no-cpu-baseline --no-e2e --steps 5 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1])
for r in j['rows']: print('w1bit $r', r['value'], r['ms_per_step'])"
done
