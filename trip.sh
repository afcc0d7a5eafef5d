mkdir -p gpurun_out
timeout 600 python tools/reactor_diag.py > gpurun_out/diag.log 2>&1; echo "rc=$?" >> gpurun_out/diag.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload ignition --steps 2 --warmup 1 --reactors 32768 > gpurun_out/bench_ign.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:advanceKernel -c 1 -o gpurun_out/advance_r01 python bench.py --workload ignition --steps 1 --warmup 0 --reactors 2048 --no-cpu-baseline > gpurun_out/ncu3.log 2>&1
cat gpurun_out/diag.log; tail -15 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_ign.log; tail -3 gpurun_out/ncu3.log
