# usage: bash tools/_mg.sh N   (inside gpurun --gpus N)
N=$1
mkdir -p gpurun_out
if [ "$N" = "8" ]; then
  for R in 1 2 4 8; do
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $R --master-addr 127.0.0.1 --master-port 29511 tools/pcie_probe.py > gpurun_out/pcie_$R.txt 2>&1
  done
fi
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --no-cpu-baseline > gpurun_out/mg_$N.out 2> gpurun_out/mg_$N.err
grep '^{' gpurun_out/mg_$N.out | tail -1 > gpurun_out/r02_bench_${N}gpu.json
python -c "
import json
d=json.load(open('gpurun_out/r02_bench_${N}gpu.json'))
print('wdot', d['value']/1e6, 'e2e', d['e2e']['value']/1e6, 'ign', d['ignition']['value'], d['ignition']['parity'].get('ok'), 'ddx', d['ddx']['value']/1e6)
"
tail -3 gpurun_out/pcie_8.txt 2>/dev/null
