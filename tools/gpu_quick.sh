# GPU box: all GPU tests, then the bench with the side configurations
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc $?"; tail -6 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc $?"; tail -3 gpurun_out/bench_quick.err
python tools/show_bench.py gpurun_out/bench_quick.json > gpurun_out/bench_quick.txt 2>&1; grep -E "^value|other c3|other c4|other photo" gpurun_out/bench_quick.txt | cut -c1-260
