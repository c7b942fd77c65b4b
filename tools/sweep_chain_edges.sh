# edge-tile weights of k_pyr_chain: kernels-only p50 of a single-frame call per (wx, wy)
for wy in 0.4 0.3 0.25; do for wx in 0.5 0.4 0.3; do
  echo "wx $wx wy $wy: $(BORB_PCH_WX=$wx BORB_PCH_WY=$wy timeout 100 python tools/lat_probe.py 2>&1 | head -1)"
done; done
