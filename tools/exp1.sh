set -x
cd $GRAFT_REPO_ROOT
export RAND=1
for cfg in "base 300 50" "base 128 32" "minb21 300 50" "minb21 128 32" "minb24 96 24"; do
  set -- $cfg
  if [ $1 = base ]; then unset B2S_LIB; else export B2S_LIB=$GRAFT_REPO_ROOT/asimov-mjlab_b200/_variants/lib_$1.so; fi
  echo "=== $cfg"
  NJMAX=$2 NCONMAX=$3 python tools/prof_run.py 2>&1 | tail -6
done
