timeout 250 python tools/_cmp.py 2>&1 | tail -12 | cut -c1-150
timeout 300 python -m pytest tests/test_gpu_reactor.py -q > gpurun_out/r02j_tests.log 2>&1; tail -8 gpurun_out/r02j_tests.log
CTB_PROFILE=1 timeout 200 python bench.py --workload ignition --reactors 131072 --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/r02j_ign.json 2> gpurun_out/r02j_ign.err; grep "ctb profile" gpurun_out/r02j_ign.err | tail -3; python -c "
import json; d=json.load(open('gpurun_out/r02j_ign.json')); print(d['value'], d['roofline']['frac'], d['parity'])"
