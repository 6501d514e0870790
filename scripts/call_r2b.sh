cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2b; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "not c5_10m" > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 $O/pytest_gpu.log
python scripts/e2e_profile.py c3 > $O/e2e_c3.log 2>&1; cat $O/e2e_c3.log
python scripts/e2e_profile.py c3-primary > $O/e2e_c3p.log 2>&1; cat $O/e2e_c3p.log
python scripts/probe.py c3p c2 > $O/probe_primary.log 2>&1; RTB200_PRIMARY_FUSED=1 python scripts/probe.py c3p c2 >> $O/probe_primary.log 2>&1; cat $O/probe_primary.log
export RTB200_WAVE_SPLIT=1
python scripts/ncu_step.py c3 1 > $O/ncu_plain_c3.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wv_extend|wv_shade|wv_resolve" -s 16 -c 16 -o $O/prof_c3 python scripts/ncu_step.py c3 0 > $O/ncu_full_c3.log 2>&1
echo "ncu exit $?"; tail -3 $O/ncu_full_c3.log; ls -la $O
