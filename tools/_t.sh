timeout 300 python -m pytest tests/test_gpu_reactor.py -q > gpurun_out/r02l_tests.log 2>&1; tail -8 gpurun_out/r02l_tests.log | cut -c1-200
CTB_PROFILE=1 timeout 200 python bench.py --workload ignition --reactors 131072 --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/r02l_ign.json 2> gpurun_out/r02l_ign.err; grep "ctb profile" gpurun_out/r02l_ign.err | head -3; python -c "
import json; d=json.load(open('gpurun_out/r02l_ign.json')); print(d['value'], d['roofline']['frac'], d['roofline']['per_reactor'], d['parity'])"
