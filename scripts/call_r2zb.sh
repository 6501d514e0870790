cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2zb; mkdir -p $O
echo skip tests
timeout 600 python bench.py --workload c3 --steps 10 > $O/bench_c3_n1.json 2> $O/bench_c3_n1.err; echo "bench c3 exit $?"; tail -2 $O/bench_c3_n1.err | cut -c1-300
python - $O/bench_c3_n1.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print('value %.0f ms %.3f | e2e %s | serial %s | pipe %s' % (d['value'], d['ms_per_step'], d['e2e'], d['e2e_serial'], d.get('e2e_pipelined')))
PY
