cd $GRAFT_REPO_ROOT
for rep in 1 2 3; do for ov in 1 2; do
python bench.py --steps 20 --warmup 5 --extras 0 --cpu-baseline 0 --overlap $ov > gpurun_out/bench_ab6_ov${ov}_$rep.json 2>/dev/null
done; done
python - <<'PY'
import json,glob,os
for f in sorted(glob.glob('gpurun_out/bench_ab6_*.json')):
    j=json.loads(open(f).read().strip().splitlines()[-1]); print(os.path.basename(f), round(j['value']/1e6,3), round(j['ms_per_step'],4))
PY
