cd $GRAFT_REPO_ROOT
for wl in $WLS; do for ov in 1 2; do for rep in 1 2; do
python bench.py --steps 40 --warmup 10 --extras 0 --cpu-baseline 0 --workload $wl --overlap $ov > gpurun_out/bench_ab7_${wl}_ov${ov}_$rep.json 2>/dev/null
done; done; done
python - <<'PY'
import json,glob,os
for f in sorted(glob.glob('gpurun_out/bench_ab7_*.json')):
    j=json.loads(open(f).read().strip().splitlines()[-1]); print(os.path.basename(f), round(j['value']/1e6,3))
PY
