# C4 Euler row: staggered warp start of gpflow_sf2e_kernel (cycles spread over the 8 warps of an SM), and the row vs batch size
for s in 0 40000 160000 640000 2500000 0; do
  GPFLOW_SF2E_STAGGER=$s python bench.py --rows c4_2d_euler_5e-4 --no-cpu-baseline --no-e2e --steps 5 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1])
for r in j['rows']: print('stagger $s', r['value'], r['ms_per_step'])"
done
for b in 18 19 20 21; do
  python bench.py --euler 5e-4 --rows none --no-cpu-baseline --no-e2e --steps 10 --warmup 3 --log2-batch $b 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('log2B $b', j['value'], j['ms_per_step'])"
done
