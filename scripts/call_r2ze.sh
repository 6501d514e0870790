cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2ze; mkdir -p $O
show() { python - $1 <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print('  value %.0f ms %.3f e2e %.0f (%s, %.3f ms) serial %.0f ident %s pipe-image %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['mode'], d['e2e']['ms_per_step'], d['e2e_serial']['value'], d['assembled_frame_bit_identical_to_single_gpu'], d['e2e'].get('image_equals_serial_lane')))
PY
}
N=${N:-1}
if [ $N = 1 ]; then
  for wl in c3 c4 c3-primary; do timeout 600 python bench.py --workload $wl --steps 10 > $O/bench_${wl}_n1.json 2> $O/bench_${wl}_n1.err; echo "bench $wl n1 exit $?"; tail -2 $O/bench_${wl}_n1.err | cut -c1-300; show $O/bench_${wl}_n1.json; done
else
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519"
  timeout 500 $TR bench.py --gpus $N --steps 10 --workload c3 > $O/bench_c3_n$N.json 2> $O/bench_c3_n$N.err; echo "n$N exit $?"; tail -3 $O/bench_c3_n$N.err | cut -c1-300; show $O/bench_c3_n$N.json
fi
