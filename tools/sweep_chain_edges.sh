# edge-tile weights of k_pyr_chain: kernels-only p50 of a single-frame call per (wx, wy) [and tile size]
for wy in 0.5 0.4 0.3; do for wx in 0.6 0.5 0.4; do
  echo "wx $wx wy $wy: $(BORB_PCH_WX=$wx BORB_PCH_WY=$wy timeout 100 python tools/lat_probe.py 2>&1 | head -1)"
done; done
for t in "20 10" "16 12" "24 12" "20 14"; do set -- $t
  echo "tw $1 th $2 wx 0.5 wy 0.4: $(BORB_PCH_TW=$1 BORB_PCH_TH=$2 BORB_PCH_WX=0.5 BORB_PCH_WY=0.4 timeout 100 python tools/lat_probe.py 2>&1 | head -1)"
done
