mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
CTB_PROFILE=1 timeout 600 python bench.py --workload ignition --steps 1 --warmup 0 --reactors 65536 --no-cpu-baseline > gpurun_out/bench_ign64k.log 2> gpurun_out/bench_ign64k.err
tail -4 gpurun_out/pytest_gpu.log; cut -c1-250 gpurun_out/bench_ign64k.log; tail -4 gpurun_out/bench_ign64k.err
