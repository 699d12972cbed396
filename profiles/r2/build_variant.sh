#!/bin/bash
# build_variant.sh NAME [nvcc flags for encoder.cu ...]: a libthorcuda.so whose frame encoder is compiled with other build
# parameters (THOR_ENC_NT, THOR_ENC_MIN_CTAS, hint groups), into build/variants/NAME/ (travels to the GPU box; run the hooked
# encoder against it with LD_LIBRARY_PATH=build/variants/NAME).  The other objects are the regular build's.
set -e
cd "$(dirname "$0")/../.."
name=$1; shift
mkdir -p build/variants/$name
make -s >/dev/null
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xptxas -v "$@" \
     -c -o build/variants/$name/encoder.o thor_b200/csrc/encoder.cu 2> build/variants/$name/ptxas.log
objs=$(ls build/*.o | grep -v '/encoder.o')
nvcc -shared -cudart static -o build/variants/$name/libthorcuda.so $objs build/variants/$name/encoder.o
grep -A2 "k_encode_frame" build/variants/$name/ptxas.log | grep -o "Used [0-9]* registers\|[0-9]* bytes stack frame\|[0-9]* bytes spill stores\|[0-9]* bytes spill loads" | tr '\n' ' '
echo " -> build/variants/$name"
