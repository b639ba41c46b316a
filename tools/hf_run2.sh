cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -q -k "height or hfield or convex or asimov" > gpurun_out/gputest_hf.log 2>&1; echo "pytest exit $? $(tail -1 gpurun_out/gputest_hf.log)"; grep FAILED gpurun_out/gputest_hf.log
B2S_LIB=$GRAFT_REPO_ROOT/asimov-mjlab_b200/_variants/lib_coltime.so WL=asimov_toe_full_flat NSTEP=45 python tools/exp_col_time.py 2>&1 | tail -4
for wl in asimov_toe_full_flat; do
python bench.py --steps 40 --warmup 10 --extras 0 --cpu-baseline 0 --workload $wl > gpurun_out/bench_hf_new_$wl.json 2>/dev/null
done
python - <<'PY'
import json,glob,os
for f in sorted(glob.glob('gpurun_out/bench_hf_*.json')):
    try:
        j=json.loads(open(f).read().strip().splitlines()[-1]); print(os.path.basename(f), round(j['value']/1e6,3), {k:round(v*1e3,1) for k,v in j['roofline']['stage_ms'].items()})
    except Exception as e: print(f, 'ERR', e)
PY
