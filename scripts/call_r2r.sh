cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2r; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "not c5_10m" > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 $O/pytest_gpu.log
for rot in 0 1; do
  echo "## RTB200_LBVH_ROTATE=$rot" >> $O/probe.log
  RTB200_LBVH_ROTATE=$rot timeout 300 python scripts/probe.py c3:0 c5:0 field:0 c3p >> $O/probe.log 2>&1
done
timeout 200 python scripts/probe.py fieldsl:0 c4:0 >> $O/probe.log 2>&1
cat $O/probe.log
