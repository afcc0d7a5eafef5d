mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 --no-ignition --no-cpu-baseline > gpurun_out/bench_wdot.log 2>&1
cut -c1-120 gpurun_out/bench_wdot.log
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
CTB_PROFILE=1 timeout 600 python bench.py --workload ignition --steps 1 --warmup 0 --reactors 65536 --no-cpu-baseline > gpurun_out/bench_ign64k.log 2> gpurun_out/bench_ign64k.err
cut -c1-120 gpurun_out/bench_ign64k.log; grep "kernel time" gpurun_out/bench_ign64k.err | tail -1
