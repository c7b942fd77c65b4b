# edge-tile weights of k_pyr_chain: kernels-only p50 of a single-frame call per (wx, wy)
for w in "0.5 0.4" "0.4 0.3" "0.3 0.25" "0.25 0.2" "0.2 0.15" "0.7 0.6" "1.0 1.0"; do set -- $w
  echo "wx $1 wy $2: $(BORB_PCH_WX=$1 BORB_PCH_WY=$2 timeout 100 python tools/lat_probe.py 2>&1 | head -1)"
done
