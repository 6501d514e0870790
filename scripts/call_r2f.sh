cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2f; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "not c5_10m" > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 $O/pytest_gpu.log
