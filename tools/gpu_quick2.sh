# collision-focused parity subset with the in-tree library, then variant benches
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py -m gpu -x -q -k "agent_stop or collision_stress or golden_rollout or agent_removed or closed_form" > gpurun_out/pytest_quick2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_quick2.log
tail -4 gpurun_out/pytest_quick2.log
bash tools/gpu_variants.sh "$@"
