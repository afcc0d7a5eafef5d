mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload ignition --steps 1 --warmup 1 --reactors 16384 --no-cpu-baseline > gpurun_out/bench_ign16k.log 2>&1
timeout 900 python bench.py --workload ignition --steps 1 --warmup 1 --reactors 131072 --no-cpu-baseline > gpurun_out/bench_ign128k.log 2>&1
CTB_RATES_KERNEL=thread timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_wdot_thread.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_ign.csv python bench.py --workload ignition --steps 1 --warmup 0 --reactors 65536 --no-cpu-baseline > gpurun_out/ncu_ign.log 2>&1
tail -15 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_ign16k.log gpurun_out/bench_ign128k.log gpurun_out/bench_wdot_thread.log; tail -3 gpurun_out/ncu_ign.log
