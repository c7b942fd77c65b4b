set -x
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
nproc > gpurun_out/nproc.txt; numactl -H >> gpurun_out/nproc.txt 2>&1
python tools/hostlink_ceiling.py --ns 1,2,4,8 --repeat 2 --out gpurun_out/hostlink_r2.jsonl > gpurun_out/hostlink_r2.log 2>&1
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2950$n bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/bench_r2_n$n.json 2> gpurun_out/bench_r2_n$n.err
  echo "n=$n rc=$?"
done
python bench.py --gpus 1 --steps 10 --warmup 3 > gpurun_out/bench_r2_n1.json 2> gpurun_out/bench_r2_n1.err; echo "n=1 rc=$?"
tail -3 gpurun_out/hostlink_r2.log
