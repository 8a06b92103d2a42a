# ncu capture of the window kernels on a short run (plain run first, as the recipe demands)
set -x
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-others --no-e2e --no-cpu --windows 296"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_win.csv $B > gpurun_out/ncu1.log 2>&1
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_win_ -s 3 -c 3 -o gpurun_out/prof_win $B > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
