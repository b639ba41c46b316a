# round-2 evidence run (under gpurun): GPU tests, driver-shaped bench, long bench, ncu launch list + --set full
cd $GRAFT_REPO_ROOT
TAG=${TAG:-r2c}
python -m pytest tests -m gpu -x -q > gpurun_out/gputest_$TAG.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/gputest_$TAG.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_${TAG}_k20.json 2> gpurun_out/bench_${TAG}_k20.err; echo "bench k20 exit $?"
python bench.py --extras 0 > gpurun_out/bench_${TAG}.json 2> gpurun_out/bench_${TAG}.err; echo "bench exit $?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_${TAG}_ref.json 2> gpurun_out/bench_${TAG}_ref.err; echo "ref exit $?"
TAG=$TAG bash tools/prof_r2.sh
