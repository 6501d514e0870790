cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2z; mkdir -p $O
# warp-cooperative filtered re-search (wv_filtered_search_warp): its test first (plain and RT_CHECKED builds), then the tables
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -s -m gpu -k "filtered_research or fallback_kernel or far_and_offset" > $O/pytest_filtered.log 2>&1; echo "filtered tests exit $?"; grep "filtered re-search\|passed\|failed\|Error" $O/pytest_filtered.log | head -12
RTB200_LIB=$PWD/build/tune/checked.so timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "filtered_research or fallback_kernel or far_and_offset" > $O/pytest_filtered_checked.log 2>&1; echo "checked build exit $?"; tail -2 $O/pytest_filtered_checked.log
RTB200_WAVE_SPLIT=1 python scripts/probe_launches.py field > $O/probe_field_launch_table.log 2>&1; cat $O/probe_field_launch_table.log | cut -c1-400
for ph in 0 5; do RTB200_WAVE_SPLIT=1 python scripts/probe_rank.py 8 $ph >> $O/probe_rank_share.log 2>&1; done
python scripts/probe_rank.py 8 5 >> $O/probe_rank_share.log 2>&1
RTB200_WAVE_SPLIT=1 python scripts/probe_rank.py 1 0 >> $O/probe_rank_share.log 2>&1
cut -c1-300 $O/probe_rank_share.log
timeout 300 python scripts/probe.py c3:0 field:0 c4:0 > $O/probe.log 2>&1; cat $O/probe.log
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 $O/pytest_gpu.log
