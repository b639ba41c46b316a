# ncu of the collision kernels of the convex workloads (under gpurun)
cd $GRAFT_REPO_ROOT
export RAND=1 STEPS=2
ASSET=asimov_toe_full_flat ncu --set full --clock-control none --import-source on -k regex:'k_convex_tasks|k_collision' -s 600 -c 3 -f -o gpurun_out/col_asimov python tools/prof_run.py > gpurun_out/ncu_col_asimov.log 2>&1
ASSET=go1_rough_hf ncu --set full --clock-control none --import-source on -k regex:'k_collision' -s 200 -c 1 -f -o gpurun_out/col_hf python tools/prof_run.py > gpurun_out/ncu_col_hf.log 2>&1
tail -3 gpurun_out/ncu_col_asimov.log gpurun_out/ncu_col_hf.log
