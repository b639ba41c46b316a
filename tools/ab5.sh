cd $GRAFT_REPO_ROOT
for wl in $WLS; do for rep in 1 2; do
python bench.py --steps 40 --warmup 10 --extras 0 --cpu-baseline 0 --workload $wl > gpurun_out/bench_ab5_new_${wl}_$rep.json 2>/dev/null
B2S_LIB=$GRAFT_REPO_ROOT/asimov-mjlab_b200/_variants/lib_${BASE:-base}.so python bench.py --steps 40 --warmup 10 --extras 0 --cpu-baseline 0 --workload $wl > gpurun_out/bench_ab5_${BASE:-base}_${wl}_$rep.json 2>/dev/null
done; done
python - <<'PY'
import json,glob,os
for f in sorted(glob.glob('gpurun_out/bench_ab5_*.json')):
    try:
        j=json.loads(open(f).read().strip().splitlines()[-1]); print(os.path.basename(f), round(j['value']/1e6,3), {k:round(v*1e3,1) for k,v in j['roofline']['stage_ms'].items()})
    except Exception as e: print(f, 'ERR', e)
PY
