# 8-GPU box: where the data-parallel optimiser step's extra time goes (fence forms, exchange stamps) + a 1-GPU line of the same box
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
p=29560
for m in last none; do
  p=$((p+1)); FLYB_PEER_FENCE=$m timeout 200 $TR --master-port $p bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu > gpurun_out/r02t_bench_n8_$m.json 2> gpurun_out/r02t_bench_n8_$m.err; echo "$m $?"
done
p=$((p+1)); timeout 150 $TR --master-port $p tools/dp_stamps.py > gpurun_out/r02t_stamps_n8.txt 2>&1; tail -6 gpurun_out/r02t_stamps_n8.txt
timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r02t_bench_n1.json 2> gpurun_out/r02t_bench_n1.err; echo "n1 $?"
