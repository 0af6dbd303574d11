for m in 2 1 0; do
  GPFLOW_NVCC_EXTRA="-DGPF_SF2R_MMA=$m" python -m gpflow_b200.build --force > /dev/null 2>&1
  python bench.py --rows none --no-cpu-baseline --no-e2e --steps 8 > gpurun_out/mma_$m.json 2>/dev/null
  python - <<PY
import json
j=json.loads(open("gpurun_out/mma_$m.json").read().strip().splitlines()[-1])
print("GPF_SF2R_MMA=$m", j["roofline"]["kernel_ms"], j["value"], j["roofline"]["frac"])
PY
done
GPFLOW_NVCC_EXTRA="-DGPF_SF2R_MMA=2" python -m gpflow_b200.build --force > /dev/null 2>&1
python tools/sf2_accuracy.py 2>&1 | grep -v v1
ncu --metrics sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio --clock-control none -k regex:gpflow_sf2r_kernel -c 1 python bench.py --rows none --no-cpu-baseline --no-e2e --log2-batch 21 --steps 2 --warmup 1 2>&1 | grep -E "pct|ratio|inst_executed" 
