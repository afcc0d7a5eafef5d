mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
CTB_PROFILE=1 timeout 600 python bench.py --workload ignition --steps 1 --warmup 1 --reactors 65536 --no-cpu-baseline > gpurun_out/bench_ign64k.log 2> gpurun_out/bench_ign64k.err
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_wdot_group.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/launches_ign.csv python bench.py --workload ignition --steps 1 --warmup 0 --reactors 65536 --no-cpu-baseline > gpurun_out/ncu_ign.log 2>&1
tail -5 gpurun_out/pytest_gpu.log; cut -c1-220 gpurun_out/bench_ign64k.log; tail -2 gpurun_out/bench_ign64k.err;  cut -c1-220 gpurun_out/bench_wdot_group.log
