cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2l; mkdir -p $O
for ph in 0 5; do RTB200_WAVE_SPLIT=1 python scripts/probe_rank.py 8 $ph >> $O/probe.log 2>&1; done
grep -v stages $O/probe.log
