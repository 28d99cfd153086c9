# run_exp.sh NAME...: exp_update_phases.py against lib/libvoxelb200_NAME.so for each NAME ("cur" = the shipped library)
for name in "$@"; do
  if [ "$name" = cur ]; then lib=voxel_game_b200/lib/libvoxelb200.so; else lib=voxel_game_b200/lib/libvoxelb200_$name.so; fi
  VXB_EXP_LIB=$lib python profiles/scripts/exp_update_phases.py 64 4 2> gpurun_out/exp_$name.err | tee gpurun_out/exp_$name.jsonl
done
