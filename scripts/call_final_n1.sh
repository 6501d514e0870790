cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2final_n1; mkdir -p $O
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_c3_reference_arm.json 2> $O/ref.err; echo "reference arm exit $?"
for b in c3 c3-primary c2 c4 c5 field field-smallleaf; do
  timeout 600 python bench.py --workload $b --steps 10 > $O/bench_${b}_n1.json 2> $O/bench_${b}_n1.err; echo "bench $b exit $?"; tail -2 $O/bench_${b}_n1.err | cut -c1-300
done
python - $O <<'PY'
import json,sys,glob,os
for f in sorted(glob.glob(sys.argv[1]+'/bench_*_n1.json')):
    try: d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e: print(f, 'unreadable', e); continue
    r=d['roofline']
    print('%-22s value %8.0f (%7.3f ms) e2e %8.0f serial %8.0f | %s frac %.3f traffic %s | V %.1f T %.2f | cpu %.4f | parity %s/%s' % (os.path.basename(f)[6:-8], d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e_serial']['value'], r['bound'], r['frac'], r['traffic'], r['box_tests_per_ray'], r['tri_tests_per_ray'], d.get('cpu_baseline',{}).get('value',-1), d.get('bit_exact'), d.get('parity_rows_checked')))
PY
