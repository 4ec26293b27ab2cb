# 2-GPU box: the fuseded exchange (FLYB_DP_STEP=fused, default) against the split one: DP tests, bench lines, stamps
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
FLYB_TEST_WORLD=2 timeout 600 python -m pytest tests/test_gpu_dp.py -m gpu -q -s -rA --timeout 500 > gpurun_out/r02x_dp_tests_n2.log 2>&1; echo "exit $?" >> gpurun_out/r02x_dp_tests_n2.log; tail -8 gpurun_out/r02x_dp_tests_n2.log
p=29550
for m in fused split fused split; do
  p=$((p+1)); FLYB_DP_STEP=$m timeout 200 $TR --master-port $p bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu > gpurun_out/r02x_bench_n2_$m.$p.json 2> gpurun_out/r02x_bench_n2_$m.err; echo "$m $?"
done
for m in fused split; do
  p=$((p+1)); FLYB_DP_STEP=$m timeout 120 $TR --master-port $p tools/dp_stamps.py > gpurun_out/r02x_stamps_$m.txt 2>&1; tail -4 gpurun_out/r02x_stamps_$m.txt
done
timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r02x_bench_n1.json 2> gpurun_out/r02x_bench_n1.err; echo "n1 $?"
