#!/bin/bash
# ncu captures of the shipped library, summarised ON the box (the .ncu-rep files are too large to bring back):
#   launch list of one C3 step + `--set full` of one step of C3 / C4 / C5 (wave engine) and of rt_trace_shade on
#   C3-primary / C2.  Outputs: gpurun_out/$TAG/{ncu_summary.json, ncu_<wl>.txt, launches_c3.csv}
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"
TAG=${TAG:-r2final}; O=gpurun_out/$TAG; mkdir -p $O; T=/tmp/ncu_$TAG; mkdir -p $T
HASH=$(python -c "import bench; print(bench.source_hash())")
echo "src_hash $HASH" | tee $O/src_hash.txt
export NCU_SUMMARY_OUT=$PWD/$O/ncu_summary.json
cap() { # workload, kernel regex, skip, count, wave split
  wl=$1; rx=$2; sk=$3; ct=$4
  if [ "$5" = 1 ]; then export RTB200_WAVE_SPLIT=1; else unset RTB200_WAVE_SPLIT; fi
  timeout 300 python scripts/ncu_step.py $wl 0 > $O/plain_$wl.log 2>&1 || { echo "plain run $wl failed"; tail -3 $O/plain_$wl.log; return; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$rx" -s $sk -c $ct -o $T/prof_$wl python scripts/ncu_step.py $wl 0 > $O/ncu_$wl.log 2>&1
  echo "ncu $wl exit $?"
  python scripts/ncu_extract.py $T/prof_$wl.ncu-rep $wl $HASH profiles/${TAG}_ncu_$wl.txt > /dev/null 2>> $O/extract.err
  python scripts/ncu_summary.py $T/prof_$wl.ncu-rep $O/ncu_$wl.txt > /dev/null 2>> $O/extract.err
  sed -i "1i # ncu --set full --clock-control none, workload $wl, src_hash $HASH (bench.py source_hash of the build captured)" $O/ncu_$wl.txt
  rm -f $T/prof_$wl.ncu-rep
}
export RTB200_WAVE_SPLIT=1
timeout 300 python scripts/ncu_step.py c3 1 > $O/plain_c3_list.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file $O/launches_c3.csv python scripts/ncu_step.py c3 1 > $O/ncu_launch_c3.log 2>&1
echo "launch list exit $?"
cap c3 "wv_extend|wv_shade|wv_resolve|wv_fallback" 16 16 1
cap c4 "wv_extend|wv_shade|wv_resolve|wv_fallback" 28 28 1
cap c5 "wv_extend|wv_shade|wv_resolve|wv_fallback" 16 16 1
cap c3-primary "rt_trace_shade" 1 1 0
cap c2 "rt_trace_shade" 1 1 0
ls -la $O; cat $O/extract.err | tail -5
