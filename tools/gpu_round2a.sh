# GPU box, round 2a (state at the start of round 2 + evidence fixes): all GPU tests, smoke, reference arm, default bench,
# launch list of the bench command, full ncu capture of the window kernels at the benched size (with the FP64
# instruction counters tools/ncu_derive.py reads)
set -x
mkdir -p gpurun_out
nvidia-smi -L
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc $?"; tail -5 gpurun_out/smoke.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc $?"
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench (no flags) rc $?"
tail -3 gpurun_out/bench.err
B="python bench.py --steps 2 --warmup 1 --no-others --no-e2e --no-cpu"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c5.csv $B > gpurun_out/ncu1.log 2>&1
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --metrics sm__inst_executed_pipe_fp64.sum,sm__sass_thread_inst_executed_op_dfma_pred_on.sum,sm__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__sass_thread_inst_executed_op_dadd_pred_on.sum --clock-control none --import-source on -k regex:k_win_ -s 3 -c 3 -o gpurun_out/prof_win_c5 $B > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out | head -40
