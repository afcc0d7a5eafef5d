mkdir -p gpurun_out
for sp in 512 448 384 320; do
CTB_SLOTS_PER_SM=$sp CTB_PROFILE=1 timeout 300 python bench.py --workload ignition --reactors 524288 --steps 1 --warmup 0 --no-cpu-baseline --no-parity > gpurun_out/t_ign_$sp.json 2> gpurun_out/t_ign_$sp.err; echo "slots/SM $sp"; grep "ctb profile" gpurun_out/t_ign_$sp.err | head -1; python -c "
import json; d=json.load(open('gpurun_out/t_ign_$sp.json')); print('ign', d['value'], d['roofline']['frac'], d['roofline']['rounds'])"
done
