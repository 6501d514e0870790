cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2i; mkdir -p $O
for sg in 1 2 3 5; do for sd in 1 2; do
  echo "## SPLIT_GEN=$sg SPLIT_DEEP=$sd" >> $O/probe.log
  export RTB200_SPLIT_GEN=$sg RTB200_SPLIT_DEEP=$sd
  RTB200_WAVE_SPLIT=1 python scripts/probe_rank.py 8 5 >> $O/probe.log 2>&1
  RTB200_WAVE_SPLIT=1 python scripts/probe_rank.py 8 0 >> $O/probe.log 2>&1
  python scripts/probe_rank.py 1 0 >> $O/probe.log 2>&1
  python scripts/probe.py c4:0 c5:0 >> $O/probe.log 2>&1
done; done
grep -v "stages" $O/probe.log
