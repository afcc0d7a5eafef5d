set -x
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ratesTileKernel -c 1 -o gpurun_out/r02_wdot_tile python bench.py --no-cpu-baseline --no-sub --no-parity --steps 1 --warmup 0 --states 1048576 > gpurun_out/ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:jacTileKernel -c 1 -o gpurun_out/r02_ddx_tile python bench.py --workload ddx --no-cpu-baseline --no-parity --steps 1 --warmup 0 --states 524288 > gpurun_out/ncu2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --launch-skip 1500 --launch-count 5 -o gpurun_out/r02_round python bench.py --workload ignition --reactors 32768 --steps 1 --warmup 0 --no-cpu-baseline --no-parity > gpurun_out/ncu3.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_default_bench.csv python bench.py --no-cpu-baseline --no-parity --steps 2 --warmup 1 > gpurun_out/ncu4.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/r02_launches_default_bench.csv
