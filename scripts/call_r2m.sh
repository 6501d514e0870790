cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2m; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "not c5_10m" > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 $O/pytest_gpu.log
for sl in 1 2; do
  export RTB200_SHADE_LAUNCHES=$sl; echo "## SHADE_LAUNCHES=$sl" >> $O/probe.log
  RTB200_WAVE_SPLIT=1 python scripts/probe_rank.py 8 0 >> $O/probe.log 2>&1
  python scripts/probe_rank.py 8 0 >> $O/probe.log 2>&1
  RTB200_WAVE_SPLIT=1 python scripts/probe_rank.py 1 0 >> $O/probe.log 2>&1
  python scripts/probe_rank.py 1 0 >> $O/probe.log 2>&1
  python scripts/probe.py c4:0 c5:0 >> $O/probe.log 2>&1
done
grep -v "stages\|literal" $O/probe.log
