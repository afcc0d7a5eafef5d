timeout 300 python -m pytest tests/test_gpu_kinetics.py -q -k "ddx" > gpurun_out/r02k_tests.log 2>&1; tail -12 gpurun_out/r02k_tests.log | cut -c1-200
timeout 300 python bench.py --workload ddx --steps 2 --warmup 1 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, d['roofline'], d['parity'], d['e2e']['value']/1e6)"
