# GPU box, round 1d: new cooperative dense solver: parity tests, C3 bench, launch list
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu --no-others > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "c3 rc $?"
tail -3 gpurun_out/bench_c3.err; cat gpurun_out/bench_c3.json
B="python bench.py --workload c3 --steps 1 --warmup 1 --no-others --no-e2e --no-cpu"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_c3.csv $B > gpurun_out/ncu_c3.log 2>&1
ls -la gpurun_out
