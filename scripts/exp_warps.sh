show() { python -c "
import json;d=json.load(open('$1'));r=d['roofline'];print('$1',round(d['value'],1),{k:round(v*d['ms_per_step'],1) for k,v in r['kernel_share_of_step'].items()})"; }
for v in 16 20 24 12; do
ARB_NVCC_EXTRA="-DARB_SWEEP_WARPS=$v" python -m arborist_b200._build --force > /dev/null
timeout 300 python bench.py --no-e2e --no-cpu-baseline > gpurun_out/warps_$v.json 2>gpurun_out/warps_$v.err; show gpurun_out/warps_$v.json
done
