# build_variant.sh NAME SRC [nvcc defines...]: voxel_game_b200/lib/libvoxelb200_NAME.so = the current library with csrc/SRC.cu
# recompiled with extra -D flags (A/B experiments; profiles/scripts/exp_update_phases.py and prof_c2.py take VXB_EXP_LIB)
set -e
cd "$(dirname "$0")/../../voxel_game_b200"
name=$1; src=$2; shift 2
nvcc -Icsrc -I../include -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
  -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -cudart shared "$@" -c csrc/$src.cu -o lib/${src}_$name.o
objs=""
for o in k_update k_worldgen k_raymarch k_entities pipeline noalloc k_raster k_replicas slab; do
  if [ $o = $src ]; then objs="$objs lib/${src}_$name.o"; else objs="$objs lib/$o.o"; fi
done
nvcc -shared -gencode arch=compute_100a,code=sm_100a -cudart shared -o lib/libvoxelb200_$name.so $objs
echo built lib/libvoxelb200_$name.so
