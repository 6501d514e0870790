cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2o; mkdir -p $O
PART=${PART:-a}
if [ $PART = a ]; then
export RTB200_WAVE_SPLIT=1
python scripts/ncu_step.py c3 1 > $O/ncu_plain_c3.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/launches_c3.csv python scripts/ncu_step.py c3 1 > $O/ncu_launch_c3.log 2>&1
echo "launch list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wv_extend|wv_shade|wv_resolve|wv_fallback" -s 16 -c 16 -o $O/prof_c3 python scripts/ncu_step.py c3 0 > $O/ncu_full_c3.log 2>&1
echo "ncu c3 exit $?"; tail -2 $O/ncu_full_c3.log
else
export RTB200_WAVE_SPLIT=1
python scripts/ncu_step.py c4 0 > $O/ncu_plain_c4.log 2>&1 && timeout 900 ncu --set full --clock-control none -k regex:"wv_extend|wv_shade" -s 27 -c 8 -o $O/prof_c4 python scripts/ncu_step.py c4 0 > $O/ncu_full_c4.log 2>&1
echo "ncu c4 exit $?"; tail -2 $O/ncu_full_c4.log
unset RTB200_WAVE_SPLIT
python scripts/ncu_step.py c3-primary 0 > $O/ncu_plain_c3p.log 2>&1 && timeout 600 ncu --set full --clock-control none -k regex:"rt_trace_shade" -s 1 -c 1 -o $O/prof_c3p python scripts/ncu_step.py c3-primary 0 > $O/ncu_full_c3p.log 2>&1
echo "ncu c3p exit $?"
python scripts/ncu_step.py c2 0 > $O/ncu_plain_c2.log 2>&1 && timeout 600 ncu --set full --clock-control none -k regex:"rt_trace_shade" -s 1 -c 1 -o $O/prof_c2 python scripts/ncu_step.py c2 0 > $O/ncu_full_c2.log 2>&1
echo "ncu c2 exit $?"
fi
ls -la $O; cat $O/ncu_plain_c3.log 2>/dev/null
