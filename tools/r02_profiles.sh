#!/bin/bash
# Regenerates the round-2 single-GPU artefacts under gpurun_out/ (copied to profiles/ by hand after
# reading them). Run on the GPU box from the repository root:  bash tools/r02_profiles.sh [quick]
mkdir -p gpurun_out
O=gpurun_out
# the other configs as stand-alone bench lines (BASELINE.json configs 1, 4, 5)
timeout 300 python bench.py --workload h2o2 > $O/r02_bench_h2o2.json 2> $O/h2o2.err
timeout 900 python bench.py --workload ndodecane --steps 1 --warmup 0 > $O/r02_bench_ndodecane.json 2> $O/ndodecane.err
timeout 900 python bench.py --workload substep --steps 1 --warmup 0 > $O/r02_bench_substep.json 2> $O/substep.err
# per-kernel split of one ignition batch (CUDA events around every kernel of a round)
CTB_PROFILE=1 timeout 300 python bench.py --workload ignition --reactors 131072 --steps 1 --warmup 0 \
    --no-cpu-baseline --no-parity > $O/r02_ign_131072.json 2> $O/r02_batch_integrator_profile.txt
# the reference arm (CPU only)
timeout 900 python bench.py --impl reference > $O/r02_bench_reference.json 2> $O/reference.err
if [ "$1" != "quick" ]; then
    # launch list of the default bench command (after it has exited 0 without ncu)
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
        --log-file $O/r02_launches_default_bench.csv python bench.py --no-cpu-baseline --no-parity --steps 2 --warmup 1 \
        > $O/ncu_launches.log 2>&1
fi
ls -la $O
