timeout 600 python -m pytest tests/test_gpu_kinetics.py tests/test_gpu_reactor.py -q -x > gpurun_out/r02t_tests.log 2>&1; tail -4 gpurun_out/r02t_tests.log | cut -c1-200
timeout 200 python bench.py --no-cpu-baseline --no-sub --no-parity --steps 5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, d['roofline']['frac'], d['e2e']['value']/1e6)"
CTB_PROFILE=1 timeout 200 python bench.py --workload ignition --reactors 131072 --steps 1 --warmup 0 --no-cpu-baseline --no-parity > gpurun_out/r02t_ign.json 2> gpurun_out/r02t_ign.err; grep "ctb profile" gpurun_out/r02t_ign.err | head -1; python -c "
import json; d=json.load(open('gpurun_out/r02t_ign.json')); print(d['value'], d['roofline']['frac'])"
