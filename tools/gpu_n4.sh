# 4-GPU box: data-parallel tests at world 4, bench lines at 4 / 2 / 1 GPUs of the same box (scaling table)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
FLYB_TEST_WORLD=4 timeout 600 python -m pytest tests/test_gpu_dp.py -m gpu -q -s -rA --timeout 500 > gpurun_out/r02_dp_tests_n4.log 2>&1; echo "exit $?" >> gpurun_out/r02_dp_tests_n4.log; tail -5 gpurun_out/r02_dp_tests_n4.log
timeout 300 $TR --nproc-per-node 4 --master-port 29531 bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/r02m_bench_n4.json 2> gpurun_out/r02m_bench_n4.err; echo "n4 $?"
timeout 300 $TR --nproc-per-node 2 --master-port 29532 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02m_bench_n2.json 2> gpurun_out/r02m_bench_n2.err; echo "n2 $?"
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/r02m_bench_n1.json 2> gpurun_out/r02m_bench_n1.err; echo "n1 $?"
