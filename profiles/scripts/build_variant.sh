# build_variant.sh NAME [nvcc defines...]: libvoxelb200_NAME.so from the current sources with extra -D flags on k_update.cu
set -e
cd "$(dirname "$0")/../../voxel_game_b200"
name=$1; shift
nvcc -Icsrc -I../include -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
  -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -cudart shared "$@" -c csrc/k_update.cu -o lib/k_update_$name.o
nvcc -shared -gencode arch=compute_100a,code=sm_100a -cudart shared -o lib/libvoxelb200_$name.so lib/k_update_$name.o \
  lib/k_worldgen.o lib/k_raymarch.o lib/k_entities.o lib/pipeline.o lib/noalloc.o lib/k_raster.o lib/k_replicas.o lib/slab.o
echo built lib/libvoxelb200_$name.so
