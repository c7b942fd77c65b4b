# sweep of K2's grid (frames walked per CTA): BORB_FAST_GRID_Y = CTA rows per cell column
for gy in 29 64 96 128 192 256 512; do echo "grid_y $gy"; BORB_FAST_GRID_Y=$gy python bench.py --quick --steps 20 --warmup 5 2>/dev/null | tail -1 | cut -c60-160; done
