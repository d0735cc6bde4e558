show() { python -c "
import json;d=json.load(open('$1'));r=d['roofline'];print('$1',round(d['value'],1),'z us/tree-sweep',round(1e3*r['avg_launch_ms']/r['tree_sweeps_per_launch'],3),'y',round(1e3*r['qy_sweep']['avg_launch_ms']/r['tree_sweeps_per_launch'],3), 'ts/launch', round(r['tree_sweeps_per_launch'],1))"; }
for v in "-DARB_EXP_HALFSTREAM" "-DARB_EXP_HALFSTREAM -DARB_EXP_NOSTORE"; do
ARB_NVCC_EXTRA="$v" python -m arborist_b200._build --force > /dev/null
n=$(echo "$v" | tr -d ' -')
timeout 300 python bench.py --no-e2e --no-cpu-baseline --steps 2 > gpurun_out/exp_$n.json 2>gpurun_out/exp_$n.err; show gpurun_out/exp_$n.json
done
