# ncu evidence of one sub-step in the standing regime (run under gpurun): launch list + --set full with sources
# (a G1 step is 8 launches that match the filter: 7 stage kernels + the solver's launch for worlds with many rows; put_data 1,
# forward 7, 40 steps before the captured one: skip 328)
cd $GRAFT_REPO_ROOT
TAG=${TAG:-r2a}
export STAND=1 STEPS=4
K='regex:k_(position|dynamics|collision|constraint|solve|sensors|integrate)'
python tools/prof_run.py > gpurun_out/prof_plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s ${SKIP:-328} -c 32 --csv --log-file gpurun_out/launches_$TAG.csv python tools/prof_run.py > gpurun_out/ncu1_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k "$K" -s ${SKIP:-328} -c 8 -f -o gpurun_out/step_$TAG python tools/prof_run.py > gpurun_out/ncu2_$TAG.log 2>&1
tail -3 gpurun_out/prof_plain_$TAG.log; tail -2 gpurun_out/ncu2_$TAG.log
