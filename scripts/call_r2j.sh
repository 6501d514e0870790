cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2j; mkdir -p $O
for lib in "" build/tune/nofb.so; do
  if [ -n "$lib" ]; then export RTB200_LIB=$PWD/$lib; else unset RTB200_LIB; fi
  echo "## lib=${lib:-default}" >> $O/probe.log
  for ph in 0 1 2 3 4 5 6 7; do RTB200_WAVE_SPLIT=1 python scripts/probe_rank.py 8 $ph >> $O/probe.log 2>&1; done
done
grep -v "stages" $O/probe.log
