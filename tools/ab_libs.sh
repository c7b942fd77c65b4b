# A/B helper: bench.py --quick with the in-tree library and with every library under variants/
mkdir -p gpurun_out
: > gpurun_out/ab.txt
for lib in orb_slam2_test_b200/liborb_b200.so variants/*.so; do
  [ -f "$lib" ] || continue
  for r in 1 2; do BORB_LIB=$PWD/$lib timeout 300 python bench.py --quick --steps 20 --warmup 5 2>&1 | tail -1 | sed "s|^|$lib |" >> gpurun_out/ab.txt; done
done
cat gpurun_out/ab.txt
