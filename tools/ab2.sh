cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -q > gpurun_out/gputest_ab2.log 2>&1; echo "pytest exit $? $(tail -1 gpurun_out/gputest_ab2.log)"; grep FAILED gpurun_out/gputest_ab2.log
for ov in 1 0; do for rep in 1 2; do
python bench.py --steps 20 --warmup 5 --extras 0 --cpu-baseline 0 --overlap $ov > gpurun_out/bench_ab2_ov${ov}_$rep.json 2>/dev/null
done; done
python - <<'PY'
import json,glob,os
for f in sorted(glob.glob('gpurun_out/bench_ab2_*.json')):
    j=json.loads(open(f).read().strip().splitlines()[-1]); print(os.path.basename(f), round(j['value']/1e6,3), {k:round(v*1e3,1) for k,v in j['roofline']['stage_ms'].items()})
PY
