#!/bin/bash
# One GPU call: tuning-variant probes (build/tune/*.so), the parity suite on the fastest one, then the bench lines.
#   TAG=r11 BASE=noflat VARIANTS="flat flatpair" WORK="c3:0 c4:0" BENCHES="c3 c3-smallleaf" bash scripts/gpu_probe.sh
# Everything lands under gpurun_out/$TAG/.
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"
TAG=${TAG:-probe}; BASE=${BASE:-base}; VARIANTS=${VARIANTS:-}; WORK=${WORK:-c3:0 c4:0}; BENCHES=${BENCHES:-c3}
O=gpurun_out/$TAG
mkdir -p $O
P=$O/probe.log
: > $P
# a variant whose probe fails or hangs (PROBE_TIMEOUT seconds) ends the variant loop: the rest share its code
for v in $BASE $VARIANTS; do
    RTB200_LIB=$PWD/build/tune/$v.so timeout ${PROBE_TIMEOUT:-40} python scripts/probe.py $WORK >> $P 2>&1
    rc=$?
    if [ $rc -ne 0 ]; then echo "probe $v exit $rc: skipping the remaining variants" | tee -a $P; break; fi
done
RTB200_LIB=$PWD/build/tune/$BASE.so timeout ${PROBE_TIMEOUT:-40} python scripts/probe.py $WORK >> $P 2>&1
cat $P
# fastest variant on c3 that is not slower than base on c4 by more than 1 %
BEST=$(python - $P $BASE <<'EOF'
import sys, collections
r = collections.defaultdict(dict)
for line in open(sys.argv[1]):
    f = line.split()
    if len(f) > 6 and f[4] == "ms":
        tag = f[0][:-3]
        r[tag][f[1]] = max(r[tag].get(f[1], 0.0), float(f[5]))
B = sys.argv[2]
base = r.get(B, {})
best, gain = B, 1.015
for tag, d in r.items():
    if tag == B or "c3" not in d or "c4" not in d or not base: continue
    g3, g4 = d["c3"] / base["c3"], d["c4"] / base["c4"]
    if g3 > gain and g4 > 0.99: best, gain = tag, g3
print(best)
EOF
)
echo "best variant: $BEST" | tee $O/best.txt
if [ "$BEST" != "$BASE" ]; then
    RTB200_LIB=$PWD/build/tune/$BEST.so timeout ${PYTEST_TIMEOUT:-420} python -m pytest tests -m gpu -x -q > $O/pytest_$BEST.log 2>&1
    echo "pytest($BEST) exit $?" | tee -a $O/best.txt
    tail -3 $O/pytest_$BEST.log
    if grep -q " passed" $O/pytest_$BEST.log && ! grep -q "failed" $O/pytest_$BEST.log; then
        export RTB200_LIB=$PWD/build/tune/$BEST.so
        echo "benches below run with $BEST" | tee -a $O/best.txt
    fi
fi
for b in $BENCHES; do
    timeout ${BENCH_TIMEOUT:-300} python bench.py --workload $b > $O/bench_${b}_n1.json 2> $O/bench_${b}_n1.err; echo "bench $b exit $?"; cut -c1-200 $O/bench_${b}_n1.json
done
if [ -z "${RTB200_LIB:-}" ] && [ -n "${FINAL_PYTEST:-}" ]; then timeout 420 python -m pytest tests -m gpu -x -q > $O/pytest_final.log 2>&1; echo "pytest(in-tree lib) exit $?"; tail -3 $O/pytest_final.log; fi
