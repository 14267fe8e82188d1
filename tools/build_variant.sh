# usage: bash tools/build_variant.sh NAME "-DFOO=1 ..."   -> variants/libgpt_NAME.so
# (render.cu recompiled with the extra flags, every other object taken from the normal build)
set -e
cd "$(dirname "$0")/../gpu_pathtracer_b200/csrc"
make -s -j8 >/dev/null
mkdir -p ../../variants /tmp/gpt_variant
NV=/usr/local/cuda/bin/nvcc
$NV -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC \
    -Xcompiler -ffp-contract=off -Xptxas -v $2 -c render.cu -o /tmp/gpt_variant/render_$1.o 2> /tmp/gpt_variant/render_$1.log
OBJS=$(ls _obj/*.o | grep -v render.o)
$NV -gencode arch=compute_100a,code=sm_100a -shared -o ../../variants/libgpt_$1.so $OBJS /tmp/gpt_variant/render_$1.o -lcudart -lpthread
grep -A3 "k_pt_trace_wideILb0" /tmp/gpt_variant/render_$1.log | grep registers
