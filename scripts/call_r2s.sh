cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2s; mkdir -p $O
timeout 500 python bench.py --workload c3 --steps 10 > $O/bench_c3_n1.json 2> $O/bench_c3_n1.err; echo "bench c3 exit $?"; tail -3 $O/bench_c3_n1.err
python - $O/bench_c3_n1.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print('value %.0f (%.3f ms) e2e %.0f (%.3f ms) serial %.0f (%.3f ms) parity %s rows %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e_serial']['value'], d['e2e_serial']['ms_per_step'], d.get('bit_exact'), d.get('parity_rows_checked')))
print(d['e2e'].get('image_equals_serial_lane'), d.get('primary',{}).get('value'), d.get('primary',{}).get('e2e'))
print('roofline', {k:d['roofline'][k] for k in ('bound','achieved','peak','frac','traffic')}, d['cpu_baseline'])
PY
