# unroll factor of the pair-accumulator rank update of the dual reverse sweep (GPF_SF2E_RANK_UNROLL), C4 Euler row
for u in 2 4 16 2; do
  GPFLOW_NVCC_EXTRA="-DGPF_SF2E_RANK_UNROLL=$u" python -m gpflow_b200.build --force > /dev/null 2>&1
  python bench.py --rows c4_2d_euler_5e-4 --no-cpu-baseline --no-e2e --steps 5 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1])
for r in j['rows']: print('rank unroll $u', r['value'], r['ms_per_step'])"
done
