cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2d; mkdir -p $O
run() { echo "## $*" >> $O/probe.log; env "$@" timeout 120 python scripts/probe.py c3:0 >> $O/probe.log 2>&1; }
run A=base
for lib in "" build/tune/shade4.so; do
  for ext in 4 3; do
    for split in 2 3 4; do
      for sb in 8 2 1; do
        if [ -n "$lib" ]; then L="RTB200_LIB=$PWD/$lib"; else L="A=default"; fi
        run $L RTB200_EXT_BLOCKS_PER_SM=$ext RTB200_WAVE_SPLIT=$split RTB200_SHADE_BLOCKS_PER_SM=$sb
      done
    done
  done
done
run A=base
grep -E "^##|c3" $O/probe.log | paste - - | sed -E 's/ +/ /g' | cut -c1-220
