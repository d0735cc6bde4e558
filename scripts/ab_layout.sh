show() { python -c "
import json,sys;d=json.load(open('$1'));r=d['roofline'];print('$1',round(d['value'],1),{k:round(v*d['ms_per_step'],1) for k,v in r['kernel_share_of_step'].items()})"; }
timeout 300 python bench.py --no-e2e --no-cpu-baseline > gpurun_out/r1k_l0.json 2>gpurun_out/r1k_l0.err; show gpurun_out/r1k_l0.json
timeout 300 python bench.py --no-e2e --no-cpu-baseline > gpurun_out/r1k_l0b.json 2>gpurun_out/r1k_l0b.err; show gpurun_out/r1k_l0b.json
ARB_NVCC_EXTRA="-DARB_PIECE_LAYOUT=1" python -m arborist_b200._build --force
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 300 python bench.py --no-e2e --no-cpu-baseline > gpurun_out/r1k_l1.json 2>gpurun_out/r1k_l1.err; show gpurun_out/r1k_l1.json
timeout 300 python bench.py --no-e2e --no-cpu-baseline --groups 2 > gpurun_out/r1k_l1g2.json 2>gpurun_out/r1k_l1g2.err; show gpurun_out/r1k_l1g2.json
