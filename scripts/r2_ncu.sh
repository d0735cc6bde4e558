# full ncu capture of the sweep / finishing kernels, config given as $2 (default config3), 148 trees
TAG=${1:-r2b}; CFG=${2:-config3}; TREES=${3:-148}; SKIP=${4:-60}
CMD="python bench.py --config $CFG --steps 1 --warmup 0 --trees $TREES --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err &&
ncu --set full --clock-control none --import-source on --launch-skip $SKIP --launch-count 8 -k regex:"k_sweep|k_finish|k_seed_r" -o gpurun_out/${TAG}_full $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/${TAG}_ncu.log; ls -la gpurun_out/${TAG}_full.ncu-rep
