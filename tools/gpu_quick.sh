# quick kernel-iteration check: lidar/kNN parity subset with the in-tree library, then variant benches
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py -m gpu -x -q -k "delta_lidar or dense_map or knn_index or golden_rollout or expert_replay_linear or collision_stress" --durations=5 > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_quick.log
tail -12 gpurun_out/pytest_quick.log
bash tools/gpu_variants.sh "$@"
