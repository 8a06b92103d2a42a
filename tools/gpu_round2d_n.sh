# GPU box, N GPUs (default 2): the driver's scaling command (headline c5 replicas + the `sharded` section: c3 / c4 point-sharded
# with parity), and the two-rank sharded GPU test
set -x
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; echo "bench n$N rc $?"
tail -3 gpurun_out/scale_n$N.err
python - <<EOF
import json
d = json.load(open("gpurun_out/scale_n$N.json"))
print("c5", d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"])
print(json.dumps(d.get("sharded"), indent=1)[:6000])
EOF
timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -q > gpurun_out/pytest_sharded_n$N.log 2>&1; echo "pytest rc $?"; tail -5 gpurun_out/pytest_sharded_n$N.log
