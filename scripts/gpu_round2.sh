#!/bin/bash
# One GPU call of round 2: parity suite, smoke, bench (both arms), then the ncu passes of the same short program.
#   TAG=r2a [NCU=1] [WORKLOADS="c3 c4"] bash scripts/gpu_round2.sh
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"
TAG=${TAG:-r2a}; O=gpurun_out/$TAG; mkdir -p $O
if [ -n "${PYTEST:-1}" ] && [ "${PYTEST:-1}" != "0" ]; then
  timeout ${PYTEST_TIMEOUT:-900} python -m pytest tests -m gpu -x -q --durations=8 > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -15 $O/pytest_gpu.log
fi
if [ "${SMOKE:-1}" != "0" ]; then timeout 300 python -c "import __graft_entry__ as e; e.smoke()" > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $O/smoke.log; fi
for b in ${BENCHES:-c3}; do
  timeout ${BENCH_TIMEOUT:-400} python bench.py --workload $b --steps ${STEPS:-10} > $O/bench_${b}_n1.json 2> $O/bench_${b}_n1.err; echo "bench $b exit $?"; cut -c1-400 $O/bench_${b}_n1.json; tail -2 $O/bench_${b}_n1.err
done
if [ "${REFARM:-1}" != "0" ]; then
  timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_c3_reference_arm.json 2> $O/bench_ref.err; echo "reference arm exit $?"; cut -c1-200 $O/bench_c3_reference_arm.json
fi
if [ -n "${NCU:-}" ]; then
  for wlk in ${NCU_WORKLOADS:-c3}; do
    export RTB200_WAVE_SPLIT=1
    timeout 300 python scripts/ncu_step.py $wlk 1 > $O/ncu_plain_$wlk.log 2>&1 &&
    timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_$wlk.csv python scripts/ncu_step.py $wlk 1 > $O/ncu_launch_$wlk.log 2>&1
    echo "ncu launch list $wlk exit $?"
    timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"${NCU_KERNELS:-wv_extend|wv_shade|wv_resolve|rt_trace_shade}" -s ${NCU_SKIP:-0} -c ${NCU_COUNT:-64} -o $O/prof_$wlk python scripts/ncu_step.py $wlk 0 > $O/ncu_full_$wlk.log 2>&1
    echo "ncu full $wlk exit $?"; tail -2 $O/ncu_full_$wlk.log
    unset RTB200_WAVE_SPLIT
  done
fi
ls -la $O
