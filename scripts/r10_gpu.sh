#!/bin/bash
# Round-10 GPU call: tuning-variant probes (build/tune/*.so, built by the flags named in the log), parity suite on the
# fastest one, then the bench lines.  Everything lands under gpurun_out/r10/.
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out/r10
mkdir -p $O
P=$O/probe.log
: > $P
for v in base pair pref pairpref shade4 base; do
    RTB200_LIB=$PWD/build/tune/$v.so timeout 120 python scripts/probe.py c3:0 c4:0 >> $P 2>&1
done
cat $P
# fastest variant on c3 that is not slower than base on c4 by more than 1 %
BEST=$(python - $P <<'EOF'
import sys, collections
r = collections.defaultdict(dict)
for line in open(sys.argv[1]):
    f = line.split()
    if len(f) > 6 and f[1] in ("c3", "c4") and f[4] == "ms":
        tag = f[0][:-3]
        r[tag][f[1]] = max(r[tag].get(f[1], 0.0), float(f[5]))
base = r.get("base", {})
best, gain = "base", 1.015
for tag, d in r.items():
    if tag == "base" or "c3" not in d or "c4" not in d or not base: continue
    g3, g4 = d["c3"] / base["c3"], d["c4"] / base["c4"]
    if g3 > gain and g4 > 0.99: best, gain = tag, g3
print(best)
EOF
)
echo "best variant: $BEST" | tee $O/best.txt
if [ "$BEST" != "base" ]; then
    RTB200_LIB=$PWD/build/tune/$BEST.so timeout 420 python -m pytest tests -m gpu -x -q > $O/pytest_$BEST.log 2>&1
    echo "pytest($BEST) exit $?" | tee -a $O/best.txt
    tail -3 $O/pytest_$BEST.log
    if grep -q " passed" $O/pytest_$BEST.log && ! grep -q "failed" $O/pytest_$BEST.log; then
        export RTB200_LIB=$PWD/build/tune/$BEST.so
        echo "benches below run with $BEST" | tee -a $O/best.txt
    fi
fi
timeout 300 python bench.py > $O/bench_c3_n1.json 2> $O/bench_c3_n1.err; echo "bench c3 exit $?"; cat $O/bench_c3_n1.json | cut -c1-400
timeout 200 python bench.py --workload c3-smallleaf --steps 10 > $O/bench_c3-smallleaf_n1.json 2> $O/bench_c3-smallleaf_n1.err; echo "bench c3-smallleaf exit $?"
timeout 200 python bench.py --workload c3-smallleaf --traversal rebuilt --steps 10 --no-cpu-baseline > $O/bench_c3-smallleaf-rebuilt_n1.json 2> $O/bench_c3-smallleaf-rebuilt_n1.err; echo "bench c3-smallleaf rebuilt exit $?"
timeout 200 python bench.py --workload c3-primary-smallleaf --steps 20 > $O/bench_c3-primary-smallleaf_n1.json 2> $O/bench_c3-primary-smallleaf_n1.err; echo "bench c3-primary-smallleaf exit $?"
if [ -z "${RTB200_LIB:-}" ]; then timeout 420 python -m pytest tests -m gpu -x -q > $O/pytest_final.log 2>&1; echo "pytest(in-tree lib) exit $?"; tail -3 $O/pytest_final.log; fi
timeout 120 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_c3_reference_arm.json 2> $O/bench_ref.err; echo "reference arm exit $?"
