# GPU box, N GPUs (default 8), final build of round 1: the default bench as the driver launches it, the sharded C3 bench and the shard check
set -x
N=${1:-8}
mkdir -p gpurun_out
run() { P=$1; shift; timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P "$@"; }
run 29701 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu --no-others > gpurun_out/scale_c5_n$N.json 2> gpurun_out/scale_c5_n$N.err; echo "c5 n$N rc $?"
run 29702 bench.py --gpus $N --steps 10 --warmup 3 --workload c3 --no-cpu --no-others > gpurun_out/scale_c3_n$N.json 2> gpurun_out/scale_c3_n$N.err; echo "c3 n$N rc $?"
run 29703 tools/shard_check.py > gpurun_out/shard_check_n$N.json 2> gpurun_out/shard_check_n$N.err; echo "shard rc $?"; cat gpurun_out/shard_check_n$N.json
tail -n 2 gpurun_out/scale_c5_n$N.err; tail -n 2 gpurun_out/scale_c3_n$N.err
run 29704 bench.py --gpus $N --steps 5 --warmup 3 --workload c4 --no-cpu --no-others > gpurun_out/scale_c4_n$N.json 2> gpurun_out/scale_c4_n$N.err; echo "c4 n$N rc $?"
