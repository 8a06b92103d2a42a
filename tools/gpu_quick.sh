# GPU box: all GPU tests, then the per-kernel times of C3 and C4
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc $?"; tail -4 gpurun_out/pytest_gpu.log
for w in c3 c4; do timeout 300 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-e2e --no-others > gpurun_out/bench_$w.json 2>/dev/null; python tools/show_bench.py gpurun_out/bench_$w.json > gpurun_out/bench_$w.txt 2>&1; head -8 gpurun_out/bench_$w.txt | cut -c1-160; done
