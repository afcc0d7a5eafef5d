mkdir -p gpurun_out
for n in 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_${n}gpu.log 2> gpurun_out/bench_${n}gpu.err; echo "rc=$?" >> gpurun_out/bench_${n}gpu.err
cut -c1-130 gpurun_out/bench_${n}gpu.log; python -c "
import json;d=json.loads(open('gpurun_out/bench_${n}gpu.log').read().strip().splitlines()[-1]);print('e2e',d['e2e']['value'],'launches',d['gpu_launches'])"; tail -1 gpurun_out/bench_${n}gpu.err
done
