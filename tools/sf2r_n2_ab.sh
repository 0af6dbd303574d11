run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 --rows none --no-cpu-baseline --no-e2e --steps 10 --warmup 3 --log2-batch 22 > gpurun_out/n2x.json 2>/dev/null; python - <<PY
import json
j=json.loads(open("gpurun_out/n2x.json").read().strip().splitlines()[-1])
print("$2", "N=2 2^21/GPU", j["ms_per_step"], j["roofline"]["kernel_ms"])
PY
python bench.py --rows none --no-cpu-baseline --no-e2e --steps 10 --warmup 3 --log2-batch 21 > gpurun_out/n1x.json 2>/dev/null; python - <<PY
import json
j=json.loads(open("gpurun_out/n1x.json").read().strip().splitlines()[-1])
print("$2", "N=1 2^21", j["ms_per_step"], j["roofline"]["kernel_ms"])
PY
}
run 29521 "default(const fwd)"
GPFLOW_SF2_FORM=0 run 29522 "old sf2 kernel"
GPFLOW_NVCC_EXTRA="-DGPF_SF2R_CONST=0" python -m gpflow_b200.build --force > /dev/null 2>&1
run 29523 "sf2r smem weights (no const staging)"
