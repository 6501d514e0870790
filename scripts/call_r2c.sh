cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2c; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "not c5_10m" > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 $O/pytest_gpu.log
python scripts/e2e_profile.py c3 > $O/e2e_c3.log 2>&1; cat $O/e2e_c3.log
python scripts/e2e_profile.py c3-primary > $O/e2e_c3p.log 2>&1; cat $O/e2e_c3p.log
python scripts/e2e_profile.py c2 > $O/e2e_c2.log 2>&1; cat $O/e2e_c2.log
for v in "" build/tune/coop2.so ""; do
  if [ -n "$v" ]; then export RTB200_LIB=$PWD/$v; else unset RTB200_LIB; fi
  timeout 120 python scripts/probe.py c3:0 c4:0 >> $O/probe.log 2>&1
done
cat $O/probe.log
RTB200_LIB=$PWD/build/tune/coop2.so timeout 600 python -m pytest tests -m gpu -x -q -k "not c5_10m" > $O/pytest_coop2.log 2>&1; echo "pytest coop2 exit $?"; tail -3 $O/pytest_coop2.log
