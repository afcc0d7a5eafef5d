mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_default.log 2> gpurun_out/bench_default.err; echo "rc=$?" >> gpurun_out/bench_default.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.log 2> gpurun_out/bench_reference.err
timeout 600 python bench.py --workload ignition --steps 2 --warmup 1 --reactors 262144 > gpurun_out/bench_ign256k.log 2> gpurun_out/bench_ign256k.err
timeout 600 python bench.py --impl reference --workload ignition --steps 2 --warmup 1 > gpurun_out/bench_ign_reference.log 2> gpurun_out/bench_ign_reference.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_default.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
cap() { # name kernel-regex skip rest...
  name=$1; k=$2; s=$3; shift 3
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -s $s -c 1 -f -o /tmp/$name "$@" > gpurun_out/ncu_$name.log 2>&1
  ncu -i /tmp/$name.ncu-rep --page raw --csv > gpurun_out/${name}_raw.csv 2>/dev/null
  ncu -i /tmp/$name.ncu-rep --page details > gpurun_out/${name}_details.txt 2>/dev/null
}
cap r01_wdot_full ratesKernel 2 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-ignition
cap r01_rhs_full batchRhsKernel 300 python bench.py --workload ignition --steps 1 --warmup 0 --reactors 65536 --no-cpu-baseline
cap r01_setup_full batchSetupKernel 100 python bench.py --workload ignition --steps 1 --warmup 0 --reactors 65536 --no-cpu-baseline
tail -3 gpurun_out/pytest_gpu.log; cut -c1-300 gpurun_out/bench_default.log; tail -2 gpurun_out/bench_default.err; cut -c1-200 gpurun_out/bench_reference.log; cut -c1-200 gpurun_out/bench_ign256k.log; cut -c1-250 gpurun_out/bench_ign_reference.log
