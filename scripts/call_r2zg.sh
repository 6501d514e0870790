cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2zg; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523"
timeout 300 $TR bench.py --gpus 8 --steps 10 --workload c5 > $O/bench_c5_n8.json 2> $O/bench_c5_n8.err; echo "c5 n8 exit $?"; tail -2 $O/bench_c5_n8.err | cut -c1-300
python - $O/bench_c5_n8.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print('value %.0f ms %.3f e2e %.0f (%s, %.3f ms) serial %.0f ident %s pipe-image %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['mode'], d['e2e']['ms_per_step'], d['e2e_serial']['value'], d['assembled_frame_bit_identical_to_single_gpu'], d['e2e'].get('image_equals_serial_lane')))
PY
