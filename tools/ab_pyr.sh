mkdir -p gpurun_out
for v in new old; do
  if [ $v = old ]; then export BORB_PYR_RESIZE_V3=1; else unset BORB_PYR_RESIZE_V3; fi
  BENCH_NO_KSEQ=1 python bench.py > gpurun_out/ab_$v.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/ab_$v.json"))
print("$v", round(d["value"]), d["roofline"]["stage_ms_per_launch"], round(d["e2e"]["value"]), d["latency"].get("kernels_p50_us"), d["latency"].get("p50_us"))
PY
done
