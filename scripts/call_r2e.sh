cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2e; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR bench.py --gpus 8 --steps 10 --workload c3 > $O/bench_c3_n8.json 2> $O/bench_c3_n8.err; echo "c3 n8 exit $?"; cut -c1-300 $O/bench_c3_n8.json
timeout 400 $TR bench.py --gpus 8 --steps 10 --workload c5 > $O/bench_c5_n8.json 2> $O/bench_c5_n8.err; echo "c5 n8 exit $?"; cut -c1-300 $O/bench_c5_n8.json
RTB200_WAVE_SPLIT=1 timeout 300 $TR bench.py --gpus 8 --steps 10 --workload c3 > $O/bench_c3_n8_split1.json 2> $O/bench_c3_n8_split1.err; echo "c3 n8 split1 exit $?"; cut -c1-200 $O/bench_c3_n8_split1.json
timeout 300 $TR bench.py --gpus 8 --steps 10 --workload c3 --gather nccl > $O/bench_c3_n8_nccl.json 2> $O/bench_c3_n8_nccl.err; echo "c3 n8 nccl exit $?"; cut -c1-200 $O/bench_c3_n8_nccl.json
tail -3 $O/*.err
