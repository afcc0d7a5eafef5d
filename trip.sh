mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
CTB_PROFILE=1 timeout 600 python bench.py --workload ignition --steps 1 --warmup 0 --reactors 65536 --no-cpu-baseline > gpurun_out/bench_ign64k.log 2> gpurun_out/bench_ign64k.err
CTB_BDF_REUSE_JAC=1 CTB_PROFILE=1 timeout 600 python bench.py --workload ignition --steps 1 --warmup 0 --reactors 65536 --no-cpu-baseline > gpurun_out/bench_ign64k_reuse.log 2> gpurun_out/bench_ign64k_reuse.err
timeout 900 python bench.py --steps 5 --warmup 3 --no-ignition > gpurun_out/bench_wdot.log 2>&1
tail -4 gpurun_out/pytest_gpu.log; cut -c1-250 gpurun_out/bench_ign64k.log; tail -4 gpurun_out/bench_ign64k.err; cut -c1-250 gpurun_out/bench_ign64k_reuse.log; tail -4 gpurun_out/bench_ign64k_reuse.err; cut -c1-900 gpurun_out/bench_wdot.log
