set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01b_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/r01b_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r01b_bench.json 2> gpurun_out/r01b_bench.err; echo "bench exit $?"
timeout 300 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/r01b_bench_ref.json 2> gpurun_out/r01b_bench_ref.err
timeout 300 python tools/prof_kernels.py all --iters 20 > gpurun_out/r01b_kernel_times.txt 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r01b_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/r01b_ncu.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"policy_act|post_physics" -c 4 -o gpurun_out/r01b_act_post python tools/prof_kernels.py all --iters 1 > gpurun_out/r01b_ncu_full.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"ppo_fwd_bwd|ppo_dw|net_forward|gae" -c 6 -o gpurun_out/r01b_update python tools/prof_kernels.py all --iters 1 >> gpurun_out/r01b_ncu_full.log 2>&1
tail -3 gpurun_out/r01b_tests.log; cat gpurun_out/r01b_kernel_times.txt
