cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2u; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 $O/pytest_gpu.log
timeout 300 python scripts/probe.py c3:0 c5:0 field:0 fieldsl:0 c4:0 > $O/probe.log 2>&1; cat $O/probe.log
for ph in 0 5; do RTB200_WAVE_SPLIT=1 python scripts/probe_rank.py 8 $ph >> $O/probe_rank_share.log 2>&1; done
python scripts/probe_rank.py 8 5 >> $O/probe_rank_share.log 2>&1
RTB200_WAVE_SPLIT=1 python scripts/probe_rank.py 1 0 >> $O/probe_rank_share.log 2>&1
cat $O/probe_rank_share.log
echo "## fused rt_trace_shade (default)" > $O/probe_primary.log; python scripts/probe.py c3p c2 >> $O/probe_primary.log 2>&1
echo "## RTB200_PRIMARY_WAVE=1: wv_extend<GEN> + wv_shade_primary" >> $O/probe_primary.log; RTB200_PRIMARY_WAVE=1 python scripts/probe.py c3p c2 >> $O/probe_primary.log 2>&1
cat $O/probe_primary.log
