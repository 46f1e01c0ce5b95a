#!/bin/bash
# tools/build_variant.sh NAME [extra nvcc flags...]  -> ash_renderer_b200/variants/libashcuda_NAME.so (A/B experiments)
set -e
cd "$(dirname "$0")/../ash_renderer_b200/csrc"
NAME=$1; shift
mkdir -p ../variants build_$NAME
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -ftz=false -Xcompiler -fPIC -Xcompiler -O2 -Xcompiler -ffp-contract=off -ccbin /usr/bin/g++"
for f in ashcuda depth_pyramid occlusion_cull cull_compact_static light_cull post post_f32 post_fast post_pair post_taa_tol post_up_pair post_down_box instancing forward compute_utils view_graph exact_taps; do
  /usr/local/cuda/bin/nvcc $FLAGS "$@" -c $f.cu -o build_$NAME/$f.o &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libashcuda_$NAME.so build_$NAME/*.o -cudart static
rm -rf build_$NAME
echo built ../variants/libashcuda_$NAME.so
