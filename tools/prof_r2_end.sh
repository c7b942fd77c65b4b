# end of round 2: launch lists of the single-frame shape and of the bench shape, --set full of the single-frame quadtree
mkdir -p gpurun_out
python tools/lat_one.py > gpurun_out/lat_one.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_latency_1frame.csv python tools/lat_one.py > gpurun_out/ncu_l1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_quadtree|k_pyr_chain|k_results_to_host" -s 18 -c 3 -o gpurun_out/r2_latency python tools/lat_one.py > gpurun_out/ncu_l2.log 2>&1
python bench.py --quick --steps 2 --warmup 3 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 36 -c 48 --csv --log-file gpurun_out/r2_launches_v15.csv python bench.py --quick --steps 2 --warmup 3 > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_quadtree -s 3 -c 1 -o gpurun_out/r2_quadtree_512 python bench.py --quick --steps 2 --warmup 3 > gpurun_out/ncu3.log 2>&1
ls -la gpurun_out | tail -12
