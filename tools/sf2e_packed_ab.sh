# A/B of GPF_SF2E_PACKED (element-wise dual operations as packed fp32x2 instructions) on the C4 Euler row, same box
for v in 1 0 1 0; do
  GPFLOW_NVCC_EXTRA="-DGPF_SF2E_PACKED=$v" python -m gpflow_b200.build --force > /dev/null 2>&1
  python bench.py --rows c4_2d_euler_5e-4 --no-cpu-baseline --no-e2e --steps 5 > gpurun_out/e_ab.json 2>gpurun_out/e_ab.err
  python - <<PY
import json
j=json.loads(open("gpurun_out/e_ab.json").read().strip().splitlines()[-1])
for r in j["rows"]: print("packed $v", r["value"], r["ms_per_step"], r["reg"])
PY
done
