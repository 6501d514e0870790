cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2final_n8; mkdir -p $O
run() { n=$1; name=$2; shift 2; TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511"
  timeout 500 $TR bench.py --gpus $n --steps 10 "$@" > $O/$name.json 2> $O/$name.err; echo "$name exit $?"; python - $O/$name.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print('  value %.0f ms %.3f e2e %.0f (%s, %.3f ms) serial %.0f rank_ms %s ident %s tiles %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['mode'], d['e2e']['ms_per_step'], d['e2e_serial']['value'], [round(x,3) for x in d['rank_frame_ms']], d['assembled_frame_bit_identical_to_single_gpu'], (d['run'].get('tile_balance') or {}).get('tiles_per_rank')))
PY
}
run 8 bench_c3_n8 --workload c3
run 8 bench_c3_n8_nccl --workload c3 --gather nccl
run 8 bench_c5_n8 --workload c5
run 8 bench_c4_n8 --workload c4
run 4 bench_c3_n4 --workload c3
run 2 bench_c3_n2 --workload c3
