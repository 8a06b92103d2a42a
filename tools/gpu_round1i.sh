# GPU box, round 1i (final state of round 1): all GPU tests, smoke, reference arm, default bench
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc $?"; tail -4 gpurun_out/smoke.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc $?"
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench (no flags) rc $?"
tail -3 gpurun_out/bench.err
