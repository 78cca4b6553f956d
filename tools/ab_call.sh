# GPU box: A/B of the in-tree library against every library under rl-cattle-herding_b200/variants, then the gpu tests with the
# first variant.  usage (through gpurun): bash tools/ab_call.sh "head c3 mid c4 big" [notest]
cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/rl-cattle-herding_b200/variants
what=${1:-"head c3 mid c4 big"}
( python tools/ab_bench.py $what 2>&1 | grep -v Warning | sed 's/^/base /'
  for v in $(ls $V); do CHS_LIB=$V/$v python tools/ab_bench.py $what 2>&1 | grep -v Warning | sed "s/^/var  /"; done ) > gpurun_out/ab_last.txt
grep -v "T=128\|T=512" gpurun_out/ab_last.txt
if [ "$2" != "notest" ]; then
  first=$(ls $V | head -1)
  CHS_LIB=$V/$first timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
fi
