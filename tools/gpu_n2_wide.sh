# 2-GPU box: DP tests incl. the wide shape; bench lines of configs[3] (42 joints) at 2 and 1 GPUs with the fused exchange
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
FLYB_TEST_WORLD=2 FLYB_SOAK_EXCHANGES=1500 timeout 600 python -m pytest tests/test_gpu_dp.py -m gpu -q -s -rA --timeout 500 > gpurun_out/r02z_dp_tests_n2.log 2>&1; echo "exit $?" >> gpurun_out/r02z_dp_tests_n2.log; tail -8 gpurun_out/r02z_dp_tests_n2.log
timeout 250 $TR --master-port 29581 bench.py --gpus 2 --config c3 --steps 5 --warmup 3 --no-cpu > gpurun_out/r02z_bench_c3_n2.json 2> gpurun_out/r02z_bench_c3_n2.err; echo "c3 n2 $?"
timeout 250 python bench.py --config c3 --steps 5 --warmup 3 --no-cpu > gpurun_out/r02z_bench_c3_n1.json 2> gpurun_out/r02z_bench_c3_n1.err; echo "c3 n1 $?"
