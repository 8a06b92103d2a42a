# GPU box, 2 GPUs: the default bench as the driver launches it for N=2, the sharded C3 parity check, C3 strong scaling
set -x
mkdir -p gpurun_out
run() { N=$1; PORT=$2; shift 2; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT "$@"; }
run 2 29701 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu > gpurun_out/scale_c5_n2.json 2> gpurun_out/scale_c5_n2.err; echo "c5 n2 rc $?"
run 2 29702 tools/shard_check.py > gpurun_out/shard_check_n2.json 2> gpurun_out/shard_check_n2.err; echo "shard rc $?"; cat gpurun_out/shard_check_n2.json
run 2 29703 bench.py --gpus 2 --workload c3 --steps 3 --warmup 3 --no-cpu > gpurun_out/scale_c3_n2.json 2> gpurun_out/scale_c3_n2.err; echo "c3 n2 rc $?"
run 2 29704 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err; echo "ref n2 rc $?"
tail -2 gpurun_out/*.err | tail -20
