#!/bin/bash
# Development aid: builds build/libwn_ab<bits>.so variants of the library with one phase of the pipelined layer role
# removed (-DWN_ABLATE=<bits>, see wn_loop_pipe.cu) for timing on the GPU box:
#   tools/ablate.sh 0 1 2 4 8 ...   then   WN_B200_LIB=build/libwn_ab4.so python tools/sweep_pipe.py ...
cd "$(dirname "$0")/.." || exit 1
mkdir -p build
for b in "$@"; do   # <ablate bits>[:extra nvcc flags, comma separated], e.g. 0:-DWN_CHAIN_ROWS=1
  extra=""; case "$b" in *:*) extra="${b#*:}"; extra="${extra//,/ }"; b="${b%%:*}";; esac
  tag="$b$(echo "$extra" | tr -cd 'A-Za-z0-9=' )"
  nvcc $extra -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --fmad=false -Xcompiler -fPIC -shared -cudart static \
    -DWN_ABLATE=$b tacotron-2_b200/csrc/wn_api.cu tacotron-2_b200/csrc/wn_upsample.cu tacotron-2_b200/csrc/wn_loop_generic.cu \
    tacotron-2_b200/csrc/wn_loop_pipe.cu tacotron-2_b200/csrc/wn_step_full.cu tacotron-2_b200/csrc/wn_step_gemm.cu -o build/libwn_ab$tag.so &
done
wait
ls -la build/libwn_ab*.so
