# One GPU call that produces everything profiles/ keeps for a round: tests, bench (own + reference arm), per-kernel
# times, the ncu launch list of the bench command and full-set captures of the hot kernels.  TAG=r02g bash tools/gpu_round.sh
set -x
TAG=${TAG:-r02}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/${TAG}_tests.log 2>&1; echo "tests exit $?" >> gpurun_out/${TAG}_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
timeout 400 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/${TAG}_ncu.log 2>&1
# (prof_kernels.py warms up with the two-kernel schedule: the act / post_physics launches are captured from `--which act/post`)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"env_step_tc|ppo_fwd_bwd_tc|ppo_dw_tc|ppo_reduce_tc|ppo_adam|net_forward_tc|gae_scan" -c 16 -o gpurun_out/${TAG}_hot python tools/prof_kernels.py all --iters 1 > gpurun_out/${TAG}_ncu_full.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"policy_act_tc|post_physics" --launch-skip 64 -c 4 -o gpurun_out/${TAG}_hot2 python tools/prof_kernels.py all --iters 1 > gpurun_out/${TAG}_ncu_full2.log 2>&1
tail -3 gpurun_out/${TAG}_tests.log
