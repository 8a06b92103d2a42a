# GPU box, round 2b: the whole GPU suite (no -x), then compute-sanitizer memcheck on the small window / general parity tests
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
T="python -m pytest tests/test_gpu_parity.py -q -x -k test_c1_window_l2_matches_oracle"
$T > gpurun_out/san_plain.log 2>&1 && timeout 900 compute-sanitizer --tool ${1:-memcheck} --error-exitcode 7 --log-file gpurun_out/sanitizer_${1:-memcheck}.log $T > gpurun_out/san_run.log 2>&1; echo "sanitizer rc $?"
tail -5 gpurun_out/san_run.log; tail -8 gpurun_out/sanitizer_${1:-memcheck}.log
