# round-2 final measurements on one GPU: default bench line, reference arm, ncu launch list of the same command
mkdir -p gpurun_out
python bench.py > gpurun_out/r2final_bench_config3.json 2> gpurun_out/r2final_bench.err; echo "bench exit $?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2final_reference_arm.json 2> gpurun_out/r2final_ref.err; echo "ref exit $?"
python bench.py --config config5 --no-cpu-baseline > gpurun_out/r2final_bench_config5.json 2>/dev/null
python bench.py --config config2 --no-cpu-baseline --steps 5 > gpurun_out/r2final_bench_config2.json 2>/dev/null
python bench.py --config config4 --trees 64 --steps 2 --no-e2e --no-cpu-baseline > gpurun_out/r2final_bench_config4_64trees.json 2>/dev/null
CMD="python bench.py --steps 1 --warmup 1 --trees 148 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/r2final_plain148.json 2>/dev/null && ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r2final_launches_bench_148trees.csv $CMD > gpurun_out/r2final_ncu.log 2>&1
echo "ncu exit $?"
python - <<'PY'
import json
for f in ("config3","config5","config2","config4_64trees"):
    try:
        d=json.load(open(f"gpurun_out/r2final_bench_{f}.json")); r=d["roofline"]; e=d.get("e2e") or {}
        print(f, round(d["value"],1), "e2e", round(e.get("value",0),1), "frac", round(r["frac"],3), round(r["qy_sweep"]["frac"],3), round(r["whole_step"]["frac"],3), "floor", round(r["lsu_floor"]["z_frac_of_floor"],3), round(r["lsu_floor"]["y_frac_of_floor"],3), (d.get("cpu_baseline") or {}).get("value"))
    except Exception as ex: print(f, "failed", ex)
print(open("gpurun_out/r2final_reference_arm.json").read()[:300])
PY
