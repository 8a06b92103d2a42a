# GPU box, round 1g: all GPU tests, default bench + reference arm, launch list of the default bench command,
# full ncu captures of the window kernels at the benched size and of the new tracker kernels (k_direct, k_lk_*)
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc $?"
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc $?"
tail -3 gpurun_out/bench.err
B="python bench.py --steps 2 --warmup 1 --no-others --no-e2e --no-cpu"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c5.csv $B > gpurun_out/ncu1.log 2>&1
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_win_ -s 3 -c 3 -o gpurun_out/prof_win_c5 $B > gpurun_out/ncu2.log 2>&1
D="python tools/direct_bench.py --no-cpu"
$D > gpurun_out/direct_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_direct -s 30 -c 2 -o gpurun_out/prof_direct $D > gpurun_out/ncu3.log 2>&1
L="python tools/lk_bench.py --no-cpu"
$L > gpurun_out/lk_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_lk_ -s 10 -c 4 -o gpurun_out/prof_lk $L > gpurun_out/ncu4.log 2>&1
ls -la gpurun_out
