# full ncu capture of the four window-path kernels of one LM trial at the benched size (a number printed under ncu is never a bench value)
B="python bench.py --steps 2 --warmup 1 --no-others --no-e2e --no-cpu"
mkdir -p gpurun_out
$B > gpurun_out/plain2.log 2>&1 && timeout 900 ncu --set full --metrics sm__inst_executed_pipe_fp64.sum,sm__sass_thread_inst_executed_op_dfma_pred_on.sum,sm__sass_thread_inst_executed_op_dmul_pred_on.sum,sm__sass_thread_inst_executed_op_dadd_pred_on.sum --clock-control none --import-source on -k regex:"k_win_linearize_schur|k_win_finalize|k_win_update|k_solve_small" -s 4 -c 4 -f -o gpurun_out/prof_win_c5 $B > gpurun_out/ncu2.log 2>&1
grep "Profiling" gpurun_out/ncu2.log
