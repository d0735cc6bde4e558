show() { python -c "
import json;d=json.load(open('$1'));r=d['roofline'];print('$1',round(d['value'],1),{k:round(v*d['ms_per_step'],1) for k,v in r['kernel_share_of_step'].items()})"; }
for v in "-DARB_FIN_MINB=1" "-DARB_FIN_MINB=4" "-DARB_FIN_MINB=3"; do
ARB_NVCC_EXTRA="$v" python -m arborist_b200._build --force > /dev/null
n=$(echo "$v" | tr -d ' -=')
timeout 300 python bench.py --no-e2e --no-cpu-baseline > gpurun_out/fin_$n.json 2>gpurun_out/fin_$n.err; show gpurun_out/fin_$n.json
done
