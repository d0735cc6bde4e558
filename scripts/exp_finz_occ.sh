# q(z) finishing kernel compiled for 4 CTAs/SM (64 registers, 116 bytes of spills) or 3 (85 registers, none)
show() { python -c "
import json;d=json.load(open('$1'));r=d['roofline'];print('$2', round(d['value'],1), {k:round(v*d['ms_per_step'],2) for k,v in r['kernel_share_of_step'].items()})"; }
for occ in 4 3; do
ARB_NVCC_EXTRA="-DARB_FINZ_OCC=$occ" python -m arborist_b200._build --force > /dev/null
for c in config3 config4; do
T=""; [ $c = config4 ] && T="--trees 32 --steps 2"
timeout 600 python bench.py --config $c $T --no-e2e --no-cpu-baseline > gpurun_out/finz_${occ}_$c.json 2>/dev/null; show gpurun_out/finz_${occ}_$c.json "occ=$occ $c"
done; done
