cd $GRAFT_REPO_ROOT
for rep in 1 2; do
python bench.py --steps 20 --warmup 5 --extras 0 --cpu-baseline 0 > gpurun_out/bench_ab3_new_$rep.json 2>/dev/null
for v in $VARIANTS; do
B2S_LIB=$GRAFT_REPO_ROOT/asimov-mjlab_b200/_variants/lib_$v.so python bench.py --steps 20 --warmup 5 --extras 0 --cpu-baseline 0 > gpurun_out/bench_ab3_${v}_$rep.json 2>/dev/null
done; done
python bench.py --steps 300 --warmup 20 --extras 0 --cpu-baseline 0 > gpurun_out/bench_ab3_new_300.json 2>/dev/null
B2S_LIB=$GRAFT_REPO_ROOT/asimov-mjlab_b200/_variants/lib_base.so python bench.py --steps 300 --warmup 20 --extras 0 --cpu-baseline 0 > gpurun_out/bench_ab3_base_300.json 2>/dev/null
python - <<'PY'
import json,glob,os
for f in sorted(glob.glob('gpurun_out/bench_ab3_*.json')):
    try:
        j=json.loads(open(f).read().strip().splitlines()[-1]); print(os.path.basename(f), round(j['value']/1e6,3), {k:round(v*1e3,1) for k,v in j['roofline']['stage_ms'].items()})
    except Exception as e: print(f, 'ERR', e)
PY
