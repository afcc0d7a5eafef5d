timeout 600 ncu --set full --clock-control none --import-source on --launch-skip 2000 --launch-count 4 -o gpurun_out/r02n_round python bench.py --workload ignition --reactors 75776 --steps 1 --warmup 0 --no-cpu-baseline --no-parity > gpurun_out/r02n_ncu.log 2>&1
tail -3 gpurun_out/r02n_ncu.log; ls -la gpurun_out/*.ncu-rep
