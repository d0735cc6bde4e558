"""Generate tests/golden/cli_tiny/ from the UNMODIFIED reference CLI (build container only).

    python oracle/gen_cli_golden.py

Writes the three input files the reference's ``main()`` reads (A.py:1133-1140: read counts
CSV with a header, SNV clustering CSV without one, tree file for ``read_trees``) from the
``tiny`` synthetic workload, runs ``arborist.arborist.main()`` of /root/reference on them
with every output flag (A.py:1154-1188), and keeps inputs + outputs as fixtures.  The GPU
test ``tests/test_gpu_parity.py::test_cli_outputs_match_reference`` replays the same command
line through ``arborist_b200.arborist.main`` and diffs the files.
"""
from __future__ import annotations

import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
OUT = os.path.join(ROOT, "tests", "golden", "cli_tiny")

ARGS = ["--add-normal", "--max-iter", "25", "--alpha", "0.001", "--prior", "0.7", "-j", "1"]
OUTPUTS = {"--ranking": "ranking.csv", "--cell-assign": "cell_assign.csv", "--snv-assign": "snv_assign.csv",
           "--q_z": "q_z.csv", "--q_y": "q_y.csv", "-t": "tree.txt"}


def write_inputs(out):
    from arborist_b200 import synth

    d = synth.generate(synth.CONFIGS["tiny"])
    os.makedirs(out, exist_ok=True)
    d.read_counts_frame().to_csv(os.path.join(out, "read_counts.csv"), index=False)
    d.clusters_frame().to_csv(os.path.join(out, "clusters.csv"), index=False, header=False)
    with open(os.path.join(out, "trees.txt"), "w") as f:
        for t, tree in enumerate(d.trees):
            f.write(f"# tree {t}\n")
            for u, v in tree:
                f.write(f"{u} {v}\n")


def command_line(indir, outdir):
    argv = ["-R", os.path.join(indir, "read_counts.csv"), "-Y", os.path.join(indir, "clusters.csv"),
            "-T", os.path.join(indir, "trees.txt"), *ARGS]
    for flag, name in OUTPUTS.items():
        argv += [flag, os.path.join(outdir, name)]
    return argv


def main():
    from gen_golden import load_reference

    write_inputs(OUT)
    A = load_reference()
    exp = os.path.join(OUT, "expected")
    os.makedirs(exp, exist_ok=True)
    old = sys.argv
    sys.argv = ["arborist", *command_line(OUT, exp)]
    try:
        A.main()
    finally:
        sys.argv = old
    print("wrote", sorted(os.listdir(exp)))


if __name__ == "__main__":
    main()
