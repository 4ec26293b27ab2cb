# 8-GPU box, final build: data-parallel tests at world 8 and bench lines at 8 / 4 / 2 / 1 GPUs of the SAME box (scaling table)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
FLYB_TEST_WORLD=8 timeout 900 python -m pytest tests/test_gpu_dp.py -m gpu -q -s -rA --timeout 800 > gpurun_out/r02s_dp_tests_n8.log 2>&1; echo "exit $?" >> gpurun_out/r02s_dp_tests_n8.log; tail -4 gpurun_out/r02s_dp_tests_n8.log
p=29550
for n in 8 4 2; do
  p=$((p+1)); timeout 300 $TR --nproc-per-node $n --master-port $p bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/r02s_bench_n$n.json 2> gpurun_out/r02s_bench_n$n.err; echo "n$n $?"
done
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/r02s_bench_n1.json 2> gpurun_out/r02s_bench_n1.err; echo "n1 $?"
