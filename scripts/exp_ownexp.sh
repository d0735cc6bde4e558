#!/bin/bash
# A/B: libdevice exp vs arb_exp (coefficients in the constant bank) in the finishing kernels.
# For each build: GPU tests, ELBOs of the `small` golden printed with repr (must be identical between builds), bench.
set -u
for v in 0 1; do
  ARB_NVCC_EXTRA="-DARB_OWN_EXP=$v" python -c "from arborist_b200 import _build; _build.build_library(force=True)" > gpurun_out/ownexp_build_$v.log 2>&1 || { echo "build $v failed"; tail -5 gpurun_out/ownexp_build_$v.log; continue; }
  timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -1
  python - <<PY > gpurun_out/ownexp_elbo_$v.txt
import sys; sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from conftest import load_golden, golden_frames
from arborist_b200 import arborist, synth
for name in ("small", "tiny"):
    trees, rc, cl = golden_frames(load_golden(name))
    lik, best, fits = arborist(trees, rc, cl, max_iter=25, add_normal=True)
    print(name, [repr(lik[t]) for t in range(len(trees))], repr(float(best.q_z.sum())), repr(float(best.q_y.max())))
d = synth.generate(synth.CONFIGS["config2"])
lik, best, fits = arborist(d.trees, d.read_counts_frame(), d.clusters_frame(), max_iter=25, add_normal=True)
print("config2", repr(sum(lik.values())), repr(float(best.q_z[:, 1].sum())))
PY
  timeout 300 python bench.py --no-cpu-baseline --no-e2e > gpurun_out/ownexp_$v.json 2> gpurun_out/ownexp_$v.err
  python -c "
import json
j=json.loads(open('gpurun_out/ownexp_$v.json').read().strip().splitlines()[-1])
print('ARB_OWN_EXP=$v', round(j['value'],1), {k: round(x,4) for k,x in j['roofline']['kernel_share_of_step'].items()}, round(j['ms_per_step'],1))"
done
cmp gpurun_out/ownexp_elbo_0.txt gpurun_out/ownexp_elbo_1.txt && echo "ELBOs and posteriors: identical bits"
