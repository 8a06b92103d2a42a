# GPU box: time the C5 window kernels of every variant library under tools/_build/variants
# (tile shape / residency sweeps of k_win_linearize_schur: -DVISO_WIN_OBS_CAP, -DVISO_WIN_PT_CAP, -DVISO_WIN_LS_CTAS)
mkdir -p gpurun_out
cp viso_b200/libviso_ba.so /tmp/orig.so
for v in tools/_build/variants/*.so; do
  n=$(basename $v .so)
  cp $v viso_b200/libviso_ba.so
  timeout 300 python bench.py --steps 3 --warmup 3 --no-others --no-e2e --no-cpu > gpurun_out/tune_$n.json 2> gpurun_out/tune_$n.err
  echo "== $n rc $?"; python tools/show_bench.py gpurun_out/tune_$n.json 2>/dev/null | head -6
  python -c "import json; d=json.load(open('gpurun_out/tune_$n.json')); print('final_cost %.9f' % d['final_cost'])"
done
cp /tmp/orig.so viso_b200/libviso_ba.so
