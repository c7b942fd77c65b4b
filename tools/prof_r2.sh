mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/gputests_r2.txt; cat gpurun_out/gputests_r2.txt
python bench.py > gpurun_out/bench_r2_a.json 2> gpurun_out/bench_r2_a.err; tail -c 600 gpurun_out/bench_r2_a.err; head -c 1500 gpurun_out/bench_r2_a.json
python bench.py --quick --steps 2 --warmup 3 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 36 -c 48 --csv --log-file gpurun_out/r2_launches_v12.csv python bench.py --quick --steps 2 --warmup 3 > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_fast_cells -s 3 -c 1 -o gpurun_out/r2_fast python bench.py --quick --steps 2 --warmup 3 > gpurun_out/ncu2.log 2>&1
python tools/run_bow.py > gpurun_out/run_bow.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_bow.csv python tools/run_bow.py > gpurun_out/ncu3.log 2>&1
cat gpurun_out/run_bow.log; ls -la gpurun_out | tail -12
