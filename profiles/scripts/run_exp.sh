set -x
python -m pytest tests/test_gpu_parity.py tests/test_gpu_headline_oracle.py tests/test_gpu_scale.py -x -q -m gpu -k "update or config3 or voxel or kernels_agree" > gpurun_out/exp_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/exp_tests.log
tail -5 gpurun_out/exp_tests.log
VXB_EXP_LIB=voxel_game_b200/lib/libvoxelb200_prev.so python profiles/scripts/exp_update_phases.py 64 4 > gpurun_out/exp_prev.jsonl 2> gpurun_out/exp_prev.err
python profiles/scripts/exp_update_phases.py 64 4 > gpurun_out/exp_new.jsonl 2> gpurun_out/exp_new.err
cat gpurun_out/exp_prev.jsonl gpurun_out/exp_new.jsonl
tail -3 gpurun_out/exp_new.err
