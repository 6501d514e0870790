cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2k; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "not c5_10m" > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -8 $O/pytest_gpu.log
for ph in 0 5 7 3; do RTB200_WAVE_SPLIT=1 python scripts/probe_rank.py 8 $ph >> $O/probe.log 2>&1; done
RTB200_WAVE_SPLIT=2 python scripts/probe_rank.py 8 5 >> $O/probe.log 2>&1
RTB200_WAVE_SPLIT=1 python scripts/probe_rank.py 1 0 >> $O/probe.log 2>&1
python scripts/probe_rank.py 1 0 >> $O/probe.log 2>&1
python scripts/probe.py c3:0 c4:0 c5:0 c3p c2 >> $O/probe.log 2>&1
grep -v stages $O/probe.log
