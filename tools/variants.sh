#!/bin/bash
# Development aid: build tuning variants of libdis_b200.so (extra -D flags) into tools/variants/<name>.so
# usage: tools/variants.sh name1="-DDIS_CHUNK=16" name2="-DDIS_AGE=4 -DDIS_CHUNK=16" ...
# then on the GPU box:  for v in tools/variants/*.so; do DIS_B200_LIB=$v python tools/quick_bench.py; done
set -e
cd "$(dirname "$0")/.."
mkdir -p tools/variants
SRC=optical-flow-using-dense-inverse-search_b200/csrc
for spec in "$@"; do
  name="${spec%%=*}"; flags="${spec#*=}"
  ( nvcc $flags -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -fmad=false -prec-div=true -prec-sqrt=true \
      -ftz=false -Xcompiler -fPIC,-fvisibility=hidden -Xptxas -v -shared -o tools/variants/$name.so \
      $SRC/dis_api.cu $SRC/dis_search.cu $SRC/dis_densify.cu $SRC/dis_pyramid.cu $SRC/dis_color.cu $SRC/dis_pool.cu $SRC/dis_noalloc.cu -lpthread 2> tools/variants/$name.log && echo "built $name ($flags)" ) &
done
wait
