"""Read-count CSV of a synthetic workload: pandas.read_csv + arb_ingest_ids against the device reader (arb_csv_*)."""
import os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, pandas as pd
from arborist_b200 import engine, synth

name = sys.argv[1] if len(sys.argv) > 1 else "config3"
d = synth.generate(synth.CONFIGS[name])
rc = d.read_counts_frame()
path = os.path.join(tempfile.mkdtemp(), "rc.csv")
t0 = time.perf_counter()
try:                                        # pyarrow writes 10^8 rows in seconds, DataFrame.to_csv in minutes
    import pyarrow as pa, pyarrow.csv as pacsv
    pacsv.write_csv(pa.Table.from_pandas(rc, preserve_index=False), path,
                    pacsv.WriteOptions(quoting_style="none"))
except ImportError:
    rc.to_csv(path, index=False)
print(f"csv written in {time.perf_counter() - t0:.1f} s", flush=True)
print(f"{name}: {len(rc)} rows, {os.path.getsize(path) / 1e6:.1f} MB")
for rep in range(int(sys.argv[2]) if len(sys.argv) > 2 else 2):
    t0 = time.perf_counter()
    f = pd.read_csv(path)
    t1 = time.perf_counter()
    e = engine.Engine()
    cu, su = e.ingest_ids(f["cell"], f["snv"], f["alt"], f["total"], 0.001)
    t2 = time.perf_counter()
    a = e.export_sparse() if rep == 0 else None
    e.close()
    t3 = time.perf_counter()
    e = engine.Engine()
    cu2, su2, rows = e.ingest_csv(path, 0.001)
    t4 = time.perf_counter()
    if rep == 0:
        b = e.export_sparse()
        print("same matrix:", all(np.array_equal(x, y) for x, y in zip(a, b)), np.array_equal(cu, cu2), np.array_equal(su, su2))
    e.close()
    print(f"pandas.read_csv {1e3*(t1-t0):.0f} ms + ingest_ids {1e3*(t2-t1):.0f} ms   |   device CSV reader + ingest {1e3*(t4-t3):.0f} ms")
