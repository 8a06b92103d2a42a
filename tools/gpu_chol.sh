# GPU box: the large solver stand-alone (tools/chol_bench.cu): correctness on several shapes, timing, phase profile
mkdir -p gpurun_out
for a in "40 3 8 1" "97 5 3 1" "120 10 4 1" "300 20 8 0" "199 199 1 0" "500 25 8 0" "500 25 1 0" "1999 25 8 0" "1999 25 1 0" "1999 25 16 0"; do
  timeout 120 tools/_build/chol_bench $a 5; echo "rc $?"
done 2>&1 | tee gpurun_out/chol_bench.log
timeout 60 tools/_build/chol_bench_prof 199 199 1 0 5 | tee gpurun_out/chol_prof_c3.log
timeout 60 tools/_build/chol_bench_prof 1999 25 8 0 5 | tee gpurun_out/chol_prof_c4.log
