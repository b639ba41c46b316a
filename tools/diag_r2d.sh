# diagnostic run (under gpurun): headline bench with extras, collision phase timing of the convex workloads
cd $GRAFT_REPO_ROOT
python bench.py --steps 20 --warmup 5 --cpu-baseline 0 > gpurun_out/bench_diag.json 2> gpurun_out/bench_diag.err; echo "bench exit $?"
B2S_LIB=$GRAFT_REPO_ROOT/asimov-mjlab_b200/_variants/lib_coltime.so python tools/exp_col_time.py > gpurun_out/coltime.log 2>&1; echo "coltime exit $?"
cat gpurun_out/coltime.log | tail -8
python - <<'PY'
import json
j=json.loads(open('gpurun_out/bench_diag.json').read().strip().splitlines()[-1])
print(round(j['value']/1e6,3), j['roofline']['stage_ms'])
print({k:round(v['value']/1e6,2) for k,v in j['workloads'].items()})
PY
