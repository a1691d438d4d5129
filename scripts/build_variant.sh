#!/bin/bash
# build_variant.sh <suffix> <file.cu> <extra nvcc flags...>: libcdx<suffix>.so with ONE source recompiled under extra flags
# (timing experiments; load with CDX_LIB=.../libcdx<suffix>.so)
set -e
cd "$(dirname "$0")/../compliancedex_b200/csrc"
sfx=$1; src=$2; shift 2
make -s all > /dev/null
mkdir -p build/var$sfx
nvcc -O3 -lineinfo -std=c++17 -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC "$@" -c $src -o build/var$sfx/${src%.cu}.o
objs=""
for o in build/*.o; do
  if [ "$(basename $o)" = "${src%.cu}.o" ]; then objs="$objs build/var$sfx/${src%.cu}.o"; else objs="$objs $o"; fi
done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libcdx$sfx.so $objs
echo built ../libcdx$sfx.so
