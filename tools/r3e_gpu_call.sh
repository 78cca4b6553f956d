cd $GRAFT_REPO_ROOT
python tools/kernel_times.py 65536 split 2>&1 | grep -v Warn | tail -5
python tools/kernel_times.py 65536 fused 2>&1 | grep -v Warn | tail -3
python tools/kernel_times.py 262144 split 2>&1 | grep -v Warn | tail -5
