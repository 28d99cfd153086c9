set -x
python profiles/scripts/prof_c2.py all 2>&1 | tail -4
for k in k_raymarch_primary k_render k_projectiles; do
  ncu --set full --clock-control none --import-source on -k regex:$k -c 1 -f -o gpurun_out/r02_$k python profiles/scripts/prof_c2.py all > /dev/null 2>&1
done
ncu --set full --clock-control none --import-source on -k regex:k_voxel_update_tiled --launch-skip 96 -c 2 -f -o gpurun_out/r02_update_elide python profiles/scripts/prof_update.py 2 96 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_voxel_update_tiled --launch-skip 96 -c 2 -f -o gpurun_out/r02_update_store python profiles/scripts/prof_update.py 0 96 > /dev/null 2>&1
ls -la gpurun_out/*.ncu-rep
