set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_photo_ba_gpu.py -m gpu -x -q > gpurun_out/pytest_parity.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/pytest_parity.log
for w in c3 c4; do timeout 300 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-e2e --no-others > gpurun_out/bench_$w.json 2>/dev/null; python tools/show_bench.py gpurun_out/bench_$w.json > gpurun_out/bench_$w.txt 2>&1; head -6 gpurun_out/bench_$w.txt | cut -c1-160; done
