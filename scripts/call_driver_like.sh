cd ${GRAFT_REPO_ROOT:-/root/repo}; O=gpurun_out/r2drv; mkdir -p $O
N=${N:-1}
if [ $N = 1 ]; then
  SECONDS=0; python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > $O/ref_n1.json 2> $O/ref_n1.err; echo "ref n1 exit $? after ${SECONDS}s"; SECONDS=0
  python bench.py --gpus 1 --steps 20 --warmup 5 > $O/b200_n1.json 2> $O/b200_n1.err; echo "b200 n1 exit $? after ${SECONDS}s"; SECONDS=0
  python bench.py > $O/b200_default.json 2> $O/b200_default.err; echo "b200 default flags exit $? after ${SECONDS}s"
  cut -c1-300 $O/ref_n1.json; cut -c1-300 $O/b200_n1.json
else
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
  SECONDS=0; $TR bench.py --impl reference --gpus $N --steps 20 --warmup 5 > $O/ref_n$N.json 2> $O/ref_n$N.err; echo "ref n$N exit $? after ${SECONDS}s"; SECONDS=0
  $TR bench.py --gpus $N --steps 20 --warmup 5 > $O/b200_n$N.json 2> $O/b200_n$N.err; echo "b200 n$N exit $? after ${SECONDS}s"
  cut -c1-200 $O/ref_n$N.json; cut -c1-300 $O/b200_n$N.json
fi
