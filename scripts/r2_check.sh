# round-2 GPU check: parity tests, then bench lines (gpurun -- 'bash scripts/r2_check.sh TAG [configs...]')
TAG=${1:-r2a}; shift
CONFIGS=${@:-config3 config5 config2}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?"; grep -a "MAP rows\|passed\|failed\|Error\|error" gpurun_out/${TAG}_pytest_gpu.log | tail -8
show() { python -c "
import json,sys
d=json.load(open('$1'));r=d['roofline']
e=d.get('e2e') or {}
print('$1',round(d['value'],1),'trees/s', round(d['ms_per_step'],2),'ms', {k:round(v*d['ms_per_step'],2) for k,v in r['kernel_share_of_step'].items()}, 'frac z/y/step', round(r['frac'],3), round(r['qy_sweep']['frac'],3), round(r['whole_step']['frac'],3), 'launches', d['gpu_launches'], 'e2e', round(e.get('value',0),1), e.get('host_stages_ms_steps_rank0',[None])[-1])"; }
for c in $CONFIGS; do
timeout 900 python bench.py --config $c --no-cpu-baseline > gpurun_out/${TAG}_bench_${c}.json 2> gpurun_out/${TAG}_bench_${c}.err && show gpurun_out/${TAG}_bench_${c}.json || tail -5 gpurun_out/${TAG}_bench_${c}.err
done
