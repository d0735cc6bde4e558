# round-3 confirmation on one GPU: GPU tests, default bench line, reference arm (short), launch list, config-2 ncu capture
python -m pytest tests -x -q -m gpu > gpurun_out/r3d_pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/r3d_pytest_gpu.log; tail -3 gpurun_out/r3d_pytest_gpu.log
python bench.py > gpurun_out/r3d_bench_config3.json 2> gpurun_out/r3d_bench_config3.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r3d_reference_arm.json 2> gpurun_out/r3d_reference_arm.err; echo "ref rc=$?"
python bench.py --config config2 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r3d_bench_config2.json 2> gpurun_out/r3d_bench_config2.err; echo "c2 rc=$?"
CMD="python bench.py --config config2 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline"
ncu --set full --clock-control none --import-source on --launch-skip 2 --launch-count 8 -k regex:"k_sweep|k_finish|k_seed_r" -o gpurun_out/r3d_config2_full $CMD > gpurun_out/r3d_config2_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r3d_config2_full.ncu-rep --page raw --csv > gpurun_out/r3d_config2_raw.csv 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r3d_launches_bench_148trees.csv python bench.py --steps 1 --warmup 0 --trees 148 --no-e2e --no-cpu-baseline > gpurun_out/r3d_launches.log 2>&1; echo "launch list rc=$?"
