timeout 300 python -m pytest tests/test_gpu_reactor.py -q -k "jacobian or rhs" > gpurun_out/r02o_tests.log 2>&1; tail -25 gpurun_out/r02o_tests.log | cut -c1-220
