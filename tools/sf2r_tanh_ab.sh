for t in 1 0; do
  GPFLOW_NVCC_EXTRA="-DGPF_SF2R_TANH=$t" python -m gpflow_b200.build --force > /dev/null 2>&1
  echo "== GPF_SF2R_TANH=$t"
  python tools/sf2_accuracy.py 2>&1 | grep -v v1
  python bench.py --rows none --no-cpu-baseline --no-e2e --steps 5 > gpurun_out/tanh_$t.json 2>gpurun_out/tanh_$t.err
  python - <<PY
import json
j=json.loads(open("gpurun_out/tanh_$t.json").read().strip().splitlines()[-1])
print("tanh $t", j["value"], j["ms_per_step"], j["roofline"]["frac"], j["clocks"]["sm_mhz"], j["clocks"]["reasons"])
PY
done
GPFLOW_SF2_FORM=0 python tools/sf2_accuracy.py 2>&1
python -m gpflow_b200.build --force > /dev/null 2>&1
python -m pytest tests/test_sf2_parity.py tests/test_train_trajectory.py tests/test_cuda_parity.py tests/test_module_surface.py -q -m gpu 2>&1 | tail -3
