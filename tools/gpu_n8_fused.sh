# 8-GPU box: the fused exchange (FLYB_DP_STEP=fused, default) — DP tests at 8 ranks, bench lines fused / split, stamps, 1-GPU line
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
FLYB_TEST_WORLD=8 timeout 700 python -m pytest tests/test_gpu_dp.py -m gpu -q -s -rA --timeout 600 > gpurun_out/r02y_dp_tests_n8.log 2>&1; echo "exit $?" >> gpurun_out/r02y_dp_tests_n8.log; tail -8 gpurun_out/r02y_dp_tests_n8.log
p=29570
for m in fused split; do
  p=$((p+1)); FLYB_DP_STEP=$m timeout 200 $TR --master-port $p bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu > gpurun_out/r02y_bench_n8_$m.json 2> gpurun_out/r02y_bench_n8_$m.err; echo "$m $?"
done
p=$((p+1)); timeout 150 $TR --master-port $p tools/dp_stamps.py > gpurun_out/r02y_stamps_n8.txt 2>&1; tail -4 gpurun_out/r02y_stamps_n8.txt
timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r02y_bench_n1.json 2> gpurun_out/r02y_bench_n1.err; echo "n1 $?"
