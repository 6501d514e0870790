cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2y; mkdir -p $O
for v in "" build/tune/stk28.so build/tune/stk24.so ""; do
  if [ -n "$v" ]; then export RTB200_LIB=$PWD/$v; else unset RTB200_LIB; fi
  RTB200_DEBUG=1 timeout 200 python scripts/probe.py c3:0 c4:0 c5:0 2>&1 | grep -v "^$" >> $O/probe.log
done
cat $O/probe.log
