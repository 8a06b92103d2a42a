# GPU box, round 1f (8 GPUs): sharded parity at N=8, C3 / C4 strong scaling at N=2,4,8, C5 replicas at N=8
set -x
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo8.txt 2>&1
run() { # N port args...
  N=$1; PORT=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT "$@"
}
run 8 29601 tools/shard_check.py > gpurun_out/shard_check_n8.json 2> gpurun_out/shard_check_n8.err; echo "shard_check c3 rc $?"
run 8 29602 tools/shard_check.py --config 4 --poses 300 --points 30000 > gpurun_out/shard_check_c4_n8.json 2> gpurun_out/shard_check_c4_n8.err; echo "shard_check c4 rc $?"
cat gpurun_out/shard_check_n8.json gpurun_out/shard_check_c4_n8.json
P=29610
for W in c3 c4; do
  timeout 600 python bench.py --gpus 1 --workload $W --steps 3 --warmup 3 --no-cpu --no-others > gpurun_out/scale_${W}_n1.json 2> gpurun_out/scale_${W}_n1.err; echo "$W n1 rc $?"
  for N in 2 4 8; do
    P=$((P+1))
    run $N $P bench.py --gpus $N --workload $W --steps 3 --warmup 3 --no-cpu > gpurun_out/scale_${W}_n$N.json 2> gpurun_out/scale_${W}_n$N.err; echo "$W n$N rc $?"
  done
done
for N in 4 8; do
  P=$((P+1))
  run $N $P bench.py --gpus $N --steps 5 --warmup 3 --no-cpu > gpurun_out/scale_c5_n$N.json 2> gpurun_out/scale_c5_n$N.err; echo "c5 n$N rc $?"
done
run 8 29640 bench.py --impl reference --gpus 8 --steps 1 --warmup 1 > gpurun_out/bench_ref_n8.json 2> gpurun_out/bench_ref_n8.err; echo "ref n8 rc $?"
tail -2 gpurun_out/*.err | tail -40
ls -la gpurun_out
