#!/bin/bash
# Variant builds of the library that differ in sensor_dev.cu only (selected with AMCLB200_LIB).
# usage: build_lf_variants.sh name "-DFLAG ..." [name "-DFLAG ..."] ...
set -e
cd "$(dirname "$0")/../../navigation_b200/csrc"
make -j8 > /dev/null
mkdir -p ../lib/variants build/variants
OTHERS=$(ls build/*.o | grep -v sensor_dev.o)
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false \
       -Xcompiler -fPIC,-O2,-Wall,-fvisibility=default -I../../include -I. $flags -c sensor_dev.cu -o build/variants/sensor_$name.o
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../lib/variants/libamcl_b200_$name.so $OTHERS build/variants/sensor_$name.o \
       -Xlinker -Bsymbolic-functions -lcudart_static -lpthread -ldl -lrt
  echo "built $name ($flags)"
done
