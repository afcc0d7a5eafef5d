mkdir -p gpurun_out
timeout 600 python tools/reactor_diag.py > gpurun_out/diag.log 2>&1; echo "rc=$?" >> gpurun_out/diag.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_wdot.log 2>&1
python - > gpurun_out/pcie.log 2>&1 <<'PY'
import torch, time
a = torch.empty(1<<28, dtype=torch.float64).pin_memory(); d = torch.empty_like(a, device='cuda')
for _ in range(2):
    torch.cuda.synchronize(); t=time.perf_counter(); d.copy_(a, non_blocking=True); torch.cuda.synchronize(); print('H2D GB/s', a.numel()*8/(time.perf_counter()-t)/1e9)
    torch.cuda.synchronize(); t=time.perf_counter(); a.copy_(d, non_blocking=True); torch.cuda.synchronize(); print('D2H GB/s', a.numel()*8/(time.perf_counter()-t)/1e9)
PY
cat gpurun_out/diag.log; tail -30 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_wdot.log gpurun_out/pcie.log
