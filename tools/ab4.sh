cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -q > gpurun_out/gputest_ab4.log 2>&1; echo "pytest exit $? $(tail -1 gpurun_out/gputest_ab4.log)"; grep FAILED gpurun_out/gputest_ab4.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
for rep in 1 2; do
python bench.py --steps 20 --warmup 5 --extras 0 --cpu-baseline 0 > gpurun_out/bench_ab4_new_$rep.json 2>gpurun_out/bench_ab4_new_$rep.err
done
python - <<'PY'
import json,glob,os
for f in sorted(glob.glob('gpurun_out/bench_ab4_*.json')):
    try:
        j=json.loads(open(f).read().strip().splitlines()[-1]); print(os.path.basename(f), round(j['value']/1e6,3), round(j['e2e']['value']/1e6,3), j['gpu_launches'], {k:round(v*1e3,1) for k,v in j['roofline']['stage_ms'].items()})
    except Exception as e: print(f, 'ERR', e)
PY
