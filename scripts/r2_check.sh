# round-2 GPU check: parity tests, then device-only bench lines of configs 3 / 5 / 2 (gpurun -- 'bash scripts/r2_check.sh TAG')
TAG=${1:-r2a}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/${TAG}_pytest_gpu.log
show() { python -c "
import json,sys
d=json.load(open('$1'));r=d['roofline']
print('$1',round(d['value'],1),'trees/s', round(d['ms_per_step'],2),'ms', {k:round(v*d['ms_per_step'],2) for k,v in r['kernel_share_of_step'].items()}, 'frac z/y/step', round(r['frac'],3), round(r['qy_sweep']['frac'],3), round(r['whole_step']['frac'],3), 'launches', d['gpu_launches'])"; }
for c in config3 config5 config2; do
timeout 600 python bench.py --config $c --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_bench_${c}_noe2e.json 2> gpurun_out/${TAG}_bench_${c}.err && show gpurun_out/${TAG}_bench_${c}_noe2e.json || tail -5 gpurun_out/${TAG}_bench_${c}.err
done
