# Round-2 profile captures (one B200). Outputs under gpurun_out/; summaries are written from them into profiles/.
set -x
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_1gpu.json 2> gpurun_out/r02_bench_1gpu.err
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02_bench_reference_arm.json 2>/dev/null
# launch list of the benchmark command (per-launch durations + DRAM bytes; cold-cache, serialised)
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02_bench_launches.csv python bench.py --steps 8 --warmup 3 --no-cpu > /dev/null 2>&1
# update kernel, 131 072-chunk slab settled by 96 ticks: phases (0,0,0) and (0,0,1), elided and with all stores
ncu --set full --clock-control none --import-source on -k regex:k_voxel_update_tiled --launch-skip 96 -c 2 -f -o gpurun_out/r02_update_elide python profiles/scripts/prof_update.py 2 96 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_voxel_update_tiled --launch-skip 96 -c 2 -f -o gpurun_out/r02_update_store python profiles/scripts/prof_update.py 0 96 > /dev/null 2>&1
# the wavefront render: the five kernels of round 0 of one frame
ncu --set full --clock-control none --import-source on -k regex:"k_shade|k_trace|k_raymarch_primary" --launch-skip 16 -c 4 -f -o gpurun_out/r02_wave python profiles/scripts/prof_c2.py render > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_raymarch_primary -c 1 -f -o gpurun_out/r02_k_raymarch_primary python profiles/scripts/prof_c2.py rays > /dev/null 2>&1
ls -la gpurun_out/
