"""Lane-steps and shared-memory wavefronts per nonzero of two tile shapes, replayed in numpy on a real workload
(no GPU): (a) the built layout -- tiles of 8 segments sorted by length INSIDE every panel; (b) "row-owned" tiles --
the 8 rows of a warp's tile are the same in every panel so that their running products can stay in registers
across panels (round-1 verdict, item 1).  A tile runs max(segment length) steps, rounded up to quads.

    python scripts/row_block_model.py [config3]       -> profiles/r2_row_block_model.md
"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from arborist_b200 import synth

name = sys.argv[1] if len(sys.argv) > 1 else "config3"
d = synth.generate(synth.CONFIGS[name])
alt_sum = np.bincount(d.snv, weights=d.alt, minlength=d.config.n_snvs)
keep = alt_sum[d.snv] > 0
_, snv = np.unique(d.snv[keep], return_inverse=True)
cell = d.cell[keep]
n, m = int(cell.max()) + 1, int(snv.max()) + 1
print(f"{name}: {n} cells x {m} SNVs after the zero-alt filter, nnz {len(snv)}\n")
print("| sweep | panels | mean segment | lane-steps / nonzero, tiles sorted per panel (built) | lane-steps / nonzero, row-owned tiles | "
      "wavefronts / nonzero, row-owned tiles with idle lane pairs predicated off |\n|---|---|---|---|---|---|")


def study(rows, cols, n_rows, n_cols, label):
    for panel_rows, kind in ((1792, "224 KB panels"), (896, "112 KB panels (double-buffered)")):
        n_p = int(np.ceil(n_cols / panel_rows))
        L = np.zeros((n_rows, n_p), dtype=np.int64)
        np.add.at(L, (rows, cols % n_p), 1)
        tot = L.sum()
        built = 0
        for p in range(n_p):
            l = np.sort(L[:, p])[::-1]
            l = np.concatenate([l, np.zeros((-len(l)) % 8, dtype=np.int64)]).reshape(-1, 8)
            built += (np.ceil(l.max(1) / 4) * 4).sum() * 8
        order = np.argsort(-L.sum(1), kind="stable")
        Ls = np.concatenate([L[order], np.zeros(((-n_rows) % 8, n_p), dtype=np.int64)]).reshape(-1, 8, n_p)
        owned = (np.ceil(Ls.max(1) / 4) * 4).sum() * 8
        pairs = Ls.reshape(-1, 4, 2, n_p).max(2).sum() * 2
        print(f"| {label}, {kind} | {n_p} | {tot / (n_rows * n_p):.1f} | {built / tot:.3f} | {owned / tot:.3f} | {pairs / tot:.3f} |")


study(cell, snv, n, m, "q(z)")
study(snv, cell, m, n, "q(y)")
