for m in 7 3 1; do
  GPFLOW_NVCC_EXTRA="-DGPF_SF2R_CONST=$m" python -m gpflow_b200.build --force > /dev/null 2>&1
  python bench.py --rows none --no-cpu-baseline --no-e2e --steps 5 > gpurun_out/ab_$m.json 2>gpurun_out/ab_$m.err
  python - <<PY
import json
j=json.loads(open("gpurun_out/ab_$m.json").read().strip().splitlines()[-1])
print("mode $m", j["value"], j["ms_per_step"], j["roofline"]["frac"], j["clocks"]["sm_mhz"], j["clocks"]["reasons"])
PY
done
GPFLOW_NVCC_EXTRA="-DGPF_SF2R_CONST=7" python -m gpflow_b200.build --force > /dev/null 2>&1
python -m pytest tests/test_sf2_parity.py -q -m gpu 2>&1 | grep -E "^(FAILED|ERROR)|passed|failed|Error|assert" | head -20
