# GPU suite (selected tests) on several variant libraries (under gpurun): VARIANTS="a b c" KEXPR="pytest -k expression"
cd $GRAFT_REPO_ROOT
for v in $VARIANTS; do
  B2S_LIB=$GRAFT_REPO_ROOT/asimov-mjlab_b200/_variants/lib_$v.so python -m pytest tests -m gpu -q ${KEXPR:+-k "$KEXPR"} > gpurun_out/gputest_var_$v.log 2>&1; echo "$v: exit $? $(tail -1 gpurun_out/gputest_var_$v.log)"
  grep FAILED gpurun_out/gputest_var_$v.log
done
