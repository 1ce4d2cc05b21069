#!/bin/bash
# experiment builds of the density kernels: libxtb_b200_<tag>.so = product objects + density.o rebuilt with extra -D flags
# usage: build_variants.sh tag1 "-DFLAGS1" tag2 "-DFLAGS2" ...
cd "$(dirname "$0")/../../xtant-3_b200/csrc" || exit 1
make -j8 > /dev/null || exit 1
while [ $# -ge 2 ]; do
  tag=$1; flags=$2; shift 2
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -ccbin /usr/bin/g++ $flags -c -o /tmp/density_$tag.o density.cu || exit 1
  objs=$(ls *.o | grep -v '^density.o$')
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libxtb_b200_$tag.so $objs /tmp/density_$tag.o -lcudart -ldl || exit 1
  echo "built libxtb_b200_$tag.so ($flags)"
done
