cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2za; mkdir -p $O
# shipped library (src_hash b20d7cd43a17d990): the whole GPU suite incl. the new filtered re-search test, then the same against the -DRT_CHECKED build
timeout 1200 python -m pytest tests -m gpu -x -q -s > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; grep "filtered re-search" $O/pytest_gpu.log; tail -2 $O/pytest_gpu.log
RTB200_LIB=$PWD/build/tune/checked.so timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -m gpu -k "not c5_10m" > $O/checked_build_pytest.log 2>&1; echo "checked build exit $?"; tail -2 $O/checked_build_pytest.log
python -c "import bench; print('src_hash', bench.source_hash())"
