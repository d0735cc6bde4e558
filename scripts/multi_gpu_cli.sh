# The console entry point on a multi-GPU box WITHOUT torchrun: one process, one context per GPU
# (arb_fit_trees_multi).  gpurun --gpus N -- 'bash scripts/multi_gpu_cli.sh'
set -e
python - <<'PY'
import os, time, threading, subprocess, sys
import numpy as np
sys.path.insert(0, os.getcwd())
from arborist_b200 import synth, engine
from arborist_b200.utils import write_trees
d = synth.generate(synth.CONFIGS["config3"])
os.makedirs("gpurun_out/cli", exist_ok=True)
d.read_counts_frame().to_csv("gpurun_out/cli/rc.csv", index=False)
d.clusters_frame().to_csv("gpurun_out/cli/clusters.csv", index=False, header=False)
rng = np.random.default_rng(7)
trees = list(d.trees)
while len(trees) < 4000:                      # 4,000 candidate trees: 2.6 s of one GPU
    trees.append(synth._edges(synth.random_tree(rng, d.config.n_clones - 1)))
write_trees(trees, "gpurun_out/cli/trees.txt")
print("devices visible:", engine.device_count())
PY
nvidia-smi --query-gpu=index,utilization.gpu,memory.used --format=csv,noheader -lms 50 > gpurun_out/cli/smi.csv &
SMI=$!
for dev in 1 all; do
T0=$(date +%s.%N)
ARBORIST_DEVICES=$dev python -m arborist.arborist -R gpurun_out/cli/rc.csv -Y gpurun_out/cli/clusters.csv \
  -T gpurun_out/cli/trees.txt --add-normal --max-iter 25 --ranking gpurun_out/cli/ranking_$dev.csv --cell-assign gpurun_out/cli/cells_$dev.csv 2>&1 | tail -4
echo "ARBORIST_DEVICES=$dev: wall $(python -c "import time;print(round($(date +%s.%N) - $T0, 2))") s (process start, CSV ingest, fit, outputs)"
done
kill $SMI
cmp gpurun_out/cli/ranking_1.csv gpurun_out/cli/ranking_all.csv && cmp gpurun_out/cli/cells_1.csv gpurun_out/cli/cells_all.csv && echo "1 GPU and all GPUs: identical ranking.csv and cell assignments"
python - <<'PY'
import collections
busy = collections.defaultdict(int)
for ln in open("gpurun_out/cli/smi.csv"):
    f = [x.strip() for x in ln.split(",")]
    if len(f) == 3 and f[1].split()[0].isdigit() and int(f[1].split()[0]) > 50:
        busy[int(f[0])] += 1
print("samples with > 50 % utilisation per GPU:", dict(sorted(busy.items())))
PY
