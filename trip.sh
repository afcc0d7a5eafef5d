mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_wdot_u4.log 2>&1
CTB_RATES_U=2 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_wdot_u2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ratesThreadKernel -c 1 -o gpurun_out/rates_thread_r01 python bench.py --steps 1 --warmup 0 --states 1048576 --no-cpu-baseline > gpurun_out/ncu_rt.log 2>&1
tail -5 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_wdot_u4.log gpurun_out/bench_wdot_u2.log; tail -3 gpurun_out/ncu_rt.log
