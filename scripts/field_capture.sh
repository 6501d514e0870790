#!/bin/bash
# ncu --set full captures of one step of the field workloads with the shipped library, summarised ON the box like
# scripts/final_capture.sh (the .ncu-rep files are too large to bring back); entries are MERGED into a copy of the
# committed profiles/ncu_summary.json.  Then the two bench lines again, now with `traffic` and the L1-hit-rate roofline.
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"
TAG=${TAG:-r2field}; O=gpurun_out/$TAG; mkdir -p $O; T=/tmp/ncu_$TAG; mkdir -p $T
HASH=$(python -c "import bench; print(bench.source_hash())")
echo "src_hash $HASH" | tee $O/src_hash.txt
cp profiles/ncu_summary.json $O/ncu_summary.json
export NCU_SUMMARY_OUT=$PWD/$O/ncu_summary.json
export RTB200_WAVE_SPLIT=1
for wl in field field-smallleaf; do
  timeout 300 python scripts/ncu_step.py $wl 0 > $O/plain_$wl.log 2>&1 || { echo "plain run $wl failed"; tail -3 $O/plain_$wl.log; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wv_extend|wv_shade|wv_resolve|wv_fallback" -s 16 -c 16 -o $T/prof_$wl python scripts/ncu_step.py $wl 0 > $O/ncu_$wl.log 2>&1
  echo "ncu $wl exit $?"
  python scripts/ncu_extract.py $T/prof_$wl.ncu-rep $wl $HASH profiles/${TAG}_ncu_$wl.txt > /dev/null 2>> $O/extract.err
  python scripts/ncu_summary.py $T/prof_$wl.ncu-rep $O/ncu_$wl.txt > /dev/null 2>> $O/extract.err
  sed -i "1i # ncu --set full --clock-control none, workload $wl, src_hash $HASH (bench.py source_hash of the build captured)" $O/ncu_$wl.txt
  rm -f $T/prof_$wl.ncu-rep
done
unset RTB200_WAVE_SPLIT
cp $O/ncu_summary.json profiles/ncu_summary.json   # on the box only: the bench lines below read it
for wl in field field-smallleaf; do
  timeout 600 python bench.py --workload $wl --steps 10 > $O/bench_${wl}_n1.json 2> $O/bench_${wl}_n1.err; echo "bench $wl exit $?"
done
ls -la $O; tail -5 $O/extract.err
