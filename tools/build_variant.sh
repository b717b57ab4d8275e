#!/bin/sh
# developer tool: build an experimental libcst_reward variant next to the shipped one
#   tools/build_variant.sh NAME [-DMACRO=value ...]   ->  cst_captioning_b200/libcst_var_NAME.so
# load it with CST_REWARD_LIB=cst_captioning_b200/libcst_var_NAME.so
set -e
name=$1; shift
cd "$(dirname "$0")/../cst_captioning_b200/csrc"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared \
     -I../../include "$@" -o ../libcst_var_$name.so *.cu
