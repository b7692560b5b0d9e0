cd $GRAFT_REPO_ROOT
( time python -m pytest tests -q -m gpu --durations=12 -s ) > gpurun_out/t_all.log 2>&1
grep -v "^$" gpurun_out/t_all.log | tail -70
