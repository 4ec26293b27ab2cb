mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
FLYB_TEST_WORLD=8 timeout 900 python -m pytest tests/test_gpu_dp.py -m gpu -q -s -rA --timeout 800 > gpurun_out/r02_dp_tests_n8.log 2>&1; echo "exit $?" >> gpurun_out/r02_dp_tests_n8.log; tail -6 gpurun_out/r02_dp_tests_n8.log
timeout 300 $TR --master-port 29521 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r02l_bench_n8.json 2> gpurun_out/r02l_bench_n8.err; echo "c1 $?"
timeout 300 $TR --master-port 29522 bench.py --gpus 8 --steps 10 --warmup 3 --config c2 > gpurun_out/r02l_bench_c2_n8.json 2> gpurun_out/r02l_bench_c2_n8.err; echo "c2 $?"
timeout 400 $TR --master-port 29523 bench.py --gpus 8 --steps 5 --warmup 3 --config c3 > gpurun_out/r02l_bench_c3_n8.json 2> gpurun_out/r02l_bench_c3_n8.err; echo "c3 $?"
timeout 400 $TR --master-port 29524 tools/sweep.py --envs 16384,65536 --horizons 32,128 > gpurun_out/r02l_sweep_n8.jsonl 2> gpurun_out/r02l_sweep_n8.err; echo "sweep $?"
