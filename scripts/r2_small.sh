# small-shard efficiency (gpurun -- 'bash scripts/r2_small.sh TAG')
TAG=${1:-r2i}
show() { python -c "
import json;d=json.load(open('$1'));r=d['roofline'];print('$2', round(d['value'],1), round(d['ms_per_step'],2), {k:round(v*d['ms_per_step'],2) for k,v in r['kernel_share_of_step'].items()}, d['round_pipeline'])"; }
for t in 125 250 1000; do
python bench.py --trees $t --no-e2e --no-cpu-baseline --steps 3 > gpurun_out/${TAG}_${t}.json 2>/dev/null; show gpurun_out/${TAG}_${t}.json "trees=$t"
done
for c in config2 config5; do
python bench.py --config $c --no-e2e --no-cpu-baseline --steps 5 > gpurun_out/${TAG}_$c.json 2>/dev/null; show gpurun_out/${TAG}_$c.json "$c"
done
