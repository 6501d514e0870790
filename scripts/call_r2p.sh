cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2p; mkdir -p $O
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $O/pytest_alone.log 2>&1; echo "alone exit $?"; tail -3 $O/pytest_alone.log
timeout 1200 python -m pytest tests -x -q -m gpu --durations=5 > $O/pytest_gpu.log 2>&1; echo "full suite exit $?"; tail -10 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as e; e.smoke()" > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -1 $O/smoke.log
