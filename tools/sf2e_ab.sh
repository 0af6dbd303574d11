for cfg in "1 1" "1 2" "3 1" "3 2" "5 2"; do
  set -- $cfg
  GPFLOW_NVCC_EXTRA="-DGPF_SF2E_UNROLL=$1 -DGPF_SF2E_WBIT=$2" python -m gpflow_b200.build --force > /dev/null 2>&1
  python bench.py --rows c4_2d_euler_5e-4 --no-cpu-baseline --no-e2e --steps 3 > gpurun_out/e_ab.json 2>gpurun_out/e_ab.err
  python - <<PY
import json
j=json.loads(open("gpurun_out/e_ab.json").read().strip().splitlines()[-1])
for r in j["rows"]: print("unroll $1 wbit $2", r["value"], r["ms_per_step"], r["reg"])
PY
done
