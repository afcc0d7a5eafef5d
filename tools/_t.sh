timeout 600 python -m pytest tests/test_gpu_reactor.py -q > gpurun_out/r02q_tests.log 2>&1; tail -25 gpurun_out/r02q_tests.log | cut -c1-220
