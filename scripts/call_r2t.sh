cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2t; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
run() { name=$1; shift; timeout 400 $TR bench.py --gpus 8 --steps 10 "$@" > $O/$name.json 2> $O/$name.err; echo "$name exit $?"; python - $O/$name.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print('  value %.0f ms %.3f e2e %.0f (%.3f ms) serial %.0f (%.3f ms) rank_ms %s ident %s pipe_img %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e_serial']['value'], d['e2e_serial']['ms_per_step'], [round(x,3) for x in d['rank_frame_ms']], d['assembled_frame_bit_identical_to_single_gpu'], d['e2e'].get('image_equals_serial_lane')))
PY
tail -2 $O/$name.err | cut -c1-300
}
run bench_c3_n8 --workload c3
run bench_c3_n8_nccl --workload c3 --gather nccl
