# GPU box, final N=1 record of round 2 (r2i build): the no-allocation entry point's test alone in a fresh process,
# then all GPU tests (it runs again, last), smoke, reference arm, default bench, launch list, full ncu capture
set -x
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "no_alloc alone rc $?" >> gpurun_out/pytest_no_alloc_alone.log
tail -5 gpurun_out/pytest_no_alloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc $?"; tail -6 gpurun_out/smoke.log
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc $?"
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench (no flags) rc $?"
tail -3 gpurun_out/bench.err
B="python bench.py --steps 2 --warmup 1 --no-others --no-e2e --no-cpu"
$B > gpurun_out/plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c5.csv $B > gpurun_out/ncu1.log 2>&1
$B > gpurun_out/plain2.log 2>&1 && timeout 900 ncu --set full --metrics sm__inst_executed_pipe_fp64.sum,sm__sass_thread_inst_executed_op_dfma_pred_on.sum,sm__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__sass_thread_inst_executed_op_dadd_pred_on.sum --clock-control none --import-source on -k regex:"k_win_linearize_schur|k_win_finalize|k_win_update|k_solve_small" -s 4 -c 4 -f -o gpurun_out/prof_win_c5 $B > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out | head -40
