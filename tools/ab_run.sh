# A/B of the in-tree library against a variant (under gpurun): GPU suite on the in-tree build, then the driver-shaped bench on both
cd $GRAFT_REPO_ROOT
TAG=${TAG:-ab}
if [ -z "$NOTEST" ]; then python -m pytest tests -m gpu -q > gpurun_out/gputest_$TAG.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/gputest_$TAG.log; fi
for rep in 1 2; do
python bench.py --steps 20 --warmup 5 --extras 0 --cpu-baseline 0 > gpurun_out/bench_${TAG}_new$rep.json 2> gpurun_out/bench_${TAG}_new$rep.err; echo "new exit $?"
B2S_LIB=$GRAFT_REPO_ROOT/asimov-mjlab_b200/_variants/lib_${BASE:-base}.so python bench.py --steps 20 --warmup 5 --extras 0 --cpu-baseline 0 > gpurun_out/bench_${TAG}_base$rep.json 2> gpurun_out/bench_${TAG}_base$rep.err; echo "base exit $?"
done
python bench.py --steps 300 --warmup 20 --extras 0 --cpu-baseline 0 > gpurun_out/bench_${TAG}_new300.json 2>/dev/null
B2S_LIB=$GRAFT_REPO_ROOT/asimov-mjlab_b200/_variants/lib_${BASE:-base}.so python bench.py --steps 300 --warmup 20 --extras 0 --cpu-baseline 0 > gpurun_out/bench_${TAG}_base300.json 2>/dev/null
python - <<'PY'
import json,glob,os
t=os.environ.get('TAG','ab')
for f in sorted(glob.glob(f'gpurun_out/bench_{t}_*.json')):
    try:
        j=json.loads(open(f).read().strip().splitlines()[-1]); print(os.path.basename(f), round(j['value']/1e6,3), 'solve', round(j['roofline']['stage_ms']['solve']*1e3,1))
    except Exception as e: print(f, 'ERR', e)
PY
