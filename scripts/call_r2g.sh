cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2g; mkdir -p $O
for lib in "" build/tune/drainpf.so; do
  if [ -n "$lib" ]; then export RTB200_LIB=$PWD/$lib; else unset RTB200_LIB; fi
  echo "## lib=${lib:-default}" >> $O/probe.log
  for split in 1 2; do
    RTB200_WAVE_SPLIT=$split python scripts/probe_rank.py 8 5 >> $O/probe.log 2>&1
    RTB200_WAVE_SPLIT=$split python scripts/probe_rank.py 8 0 >> $O/probe.log 2>&1
  done
  RTB200_WAVE_SPLIT=1 python scripts/probe_rank.py 1 0 >> $O/probe.log 2>&1
  python scripts/probe.py c3:0 c4:0 >> $O/probe.log 2>&1
done
cat $O/probe.log
