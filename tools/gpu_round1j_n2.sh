# GPU box, 2 GPUs, final state of round 1: the default bench as the driver launches it for N=2 and the sharded C3 check
set -x
mkdir -p gpurun_out
run() { N=$1; PORT=$2; shift 2; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT "$@"; }
run 2 29701 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --no-others > gpurun_out/scale_c5_n2.json 2> gpurun_out/scale_c5_n2.err; echo "c5 n2 rc $?"
run 2 29702 tools/shard_check.py > gpurun_out/shard_check_n2.json 2> gpurun_out/shard_check_n2.err; echo "shard rc $?"; cat gpurun_out/shard_check_n2.json
tail -2 gpurun_out/scale_c5_n2.err
