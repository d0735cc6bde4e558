"""Summarise ncu output into the markdown kept under profiles/.

    python profiles/summarize.py launches  <launches.csv>            -> per-kernel time shares
    python profiles/summarize.py full      <raw.csv from --page raw> -> selected counters per kernel

The launch list comes from
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file X.csv python bench.py ...
and the raw page from
    ncu -i X.ncu-rep --page raw --csv > raw.csv
"""
import csv
import sys
from collections import OrderedDict

FULL_METRICS = [
    "gpu__time_duration.sum",
    "launch__grid_size",
    "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum",
    "sm__cycles_elapsed.max",
    "l1tex__t_sector_hit_rate.pct",
]
STALL_PREFIX = "smsp__average_warps_issue_stalled_"
STALL_SUFFIX = "_per_issue_active.ratio"


def launches(path):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = OrderedDict()
    for r in rows[1:]:
        ns = float(r[iv].replace(",", ""))
        if r[iu] == "us":
            ns *= 1e3
        elif r[iu] == "ms":
            ns *= 1e6
        a = agg.setdefault(r[ik], [0, 0.0])
        a[0] += 1
        a[1] += ns
    total = sum(v[1] for v in agg.values())
    print("| kernel | launches | total ms | share |\n|---|---|---|---|")
    for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k[:70]}` | {n} | {ns / 1e6:.3f} | {100 * ns / total:.1f}% |")
    print(f"| total | | {total / 1e6:.3f} | |")


def full(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    names = [d[hdr.index("Kernel Name")].replace("void ", "").split("(")[0] for d in data]
    print("| metric | " + " | ".join(names) + " |\n|---|" + "---|" * len(names))
    for m in FULL_METRICS:
        if m in hdr:
            i = hdr.index(m)
            print(f"| {m} [{units[i]}] | " + " | ".join(d[i] for d in data) + " |")
    stalls = [(h, i) for i, h in enumerate(hdr) if h.startswith(STALL_PREFIX) and h.endswith(STALL_SUFFIX)]
    print("\nWarp stalls per issue-active cycle (top 6 per kernel):\n")
    for n, d in zip(names, data):
        top = sorted(((float(d[i]), h[len(STALL_PREFIX):-len(STALL_SUFFIX)]) for h, i in stalls), reverse=True)[:6]
        print(f"* `{n}`: " + ", ".join(f"{k} {v:.2f}" for v, k in top))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
