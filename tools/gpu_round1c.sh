# GPU box, round 1c (2 GPUs): sharded parity, sharded C3 bench, replica C5 bench
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 600 $TR --master-port 29511 tools/shard_check.py > gpurun_out/shard_check.json 2> gpurun_out/shard_check.err; echo "shard_check rc $?"
tail -3 gpurun_out/shard_check.err; cat gpurun_out/shard_check.json
timeout 600 $TR --master-port 29512 bench.py --gpus 2 --workload c3 --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_c3_n2.json 2> gpurun_out/bench_c3_n2.err; echo "c3 n2 rc $?"
tail -3 gpurun_out/bench_c3_n2.err
timeout 600 python bench.py --gpus 1 --workload c3 --steps 5 --warmup 3 --no-cpu --no-others > gpurun_out/bench_c3_n1.json 2> gpurun_out/bench_c3_n1.err; echo "c3 n1 rc $?"
timeout 900 $TR --master-port 29513 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_c5_n2.json 2> gpurun_out/bench_c5_n2.err; echo "c5 n2 rc $?"
tail -3 gpurun_out/bench_c5_n2.err
timeout 300 $TR --master-port 29514 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err; echo "ref n2 rc $?"
ls -la gpurun_out
