show() { python -c "
import json;d=json.load(open('$1'));r=d['roofline'];print('$1',round(d['value'],1),{k:round(v*d['ms_per_step'],1) for k,v in r['kernel_share_of_step'].items()},d['clocks']['sm_mhz'])"; }
for v in 1 0; do
ARB_NVCC_EXTRA="-DARB_PIECE_LAYOUT=$v" python -m arborist_b200._build --force > /dev/null
timeout 300 python bench.py --no-e2e --no-cpu-baseline > gpurun_out/y256_$v.json 2>gpurun_out/y256_$v.err; show gpurun_out/y256_$v.json
done
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
