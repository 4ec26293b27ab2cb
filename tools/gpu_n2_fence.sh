# 2-GPU box: A/B of the three peer-push fence forms (bench line + exchange stamps each)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
p=29540
for m in cta last none; do
  p=$((p+1)); FLYB_PEER_FENCE=$m timeout 200 $TR --master-port $p bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02n_bench_n2_$m.json 2> gpurun_out/r02n_bench_n2_$m.err; echo "$m $?"
  p=$((p+1)); FLYB_PEER_FENCE=$m timeout 120 $TR --master-port $p tools/dp_stamps.py > gpurun_out/r02n_stamps_$m.txt 2>&1
done
FLYB_PEER_FENCE=last FLYB_TEST_WORLD=2 timeout 400 python -m pytest tests/test_gpu_dp.py -m gpu -q -s --timeout 300 > gpurun_out/r02n_dp_tests_last.log 2>&1; tail -3 gpurun_out/r02n_dp_tests_last.log
