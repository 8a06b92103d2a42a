set -x
mkdir -p gpurun_out
tools/_build/peaks > gpurun_out/peaks.json 2> gpurun_out/peaks.err
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_r1a.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_gpu_r1a.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_r1a.json 2> gpurun_out/bench_r1a.err
B="python bench.py --steps 1 --warmup 1 --no-others --no-e2e --no-cpu --windows 148"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_r1a.csv $B > gpurun_out/ncu1.log 2>&1
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_schur_blocks -s 4 -c 1 -o gpurun_out/prof_schur_r1a $B > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
