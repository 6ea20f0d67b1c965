#!/bin/bash
# Private build of the library with the conv timeline probe compiled in (-DS2P_CONV_TRACE): tools/_trace/libs2p_trace.so.
# The product library (make -C s2p_b200/csrc) never defines the macro.
set -e
cd "$(dirname "$0")/.."
make -C s2p_b200/csrc -j8 > /dev/null
mkdir -p tools/_trace
for f in api conv_igemm; do
  nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -cudart static -DS2P_CONV_TRACE \
       -c s2p_b200/csrc/$f.cu -o tools/_trace/$f.o
done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o tools/_trace/libs2p_trace.so tools/_trace/api.o \
     tools/_trace/conv_igemm.o $(ls s2p_b200/build/*.o | grep -v "/api.o\|/conv_igemm.o")
echo built tools/_trace/libs2p_trace.so
