#!/bin/sh
# build.sh -- compile the sm_100a shared library in-tree (called by __graft_entry__.build()).
set -e
cd "$(dirname "$0")"
SRC=qpdo_b200/csrc
OUT=qpdo_b200/lib
mkdir -p "$OUT" build
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
ARCH="-gencode arch=compute_100a,code=sm_100a"
NVFLAGS="-O3 -std=c++17 -lineinfo $ARCH -Xcompiler -fPIC -fmad=false -Iinclude -I$SRC $EXTRA_NVFLAGS"
for f in dev_core dev_solver factor linesearch batch $EXTRA_CU; do
  [ -f "$SRC/$f.cu" ] || continue
  if [ ! -f build/$f.o ] || [ "$SRC/$f.cu" -nt build/$f.o ] || [ "$SRC/dev_internal.h" -nt build/$f.o ] || [ include/qpdo_cuda.h -nt build/$f.o ] || [ include/qpdo_batch.h -nt build/$f.o ]; then
    $NVCC $NVFLAGS -c "$SRC/$f.cu" -o build/$f.o &
  fi
done
g++ -O2 -std=c++17 -fPIC -Iinclude -I$SRC -c $SRC/symbolic.cpp -o build/symbolic.o &
g++ -O2 -std=c++17 -fPIC -Iinclude -I$SRC -c $SRC/ordering.cpp -o build/ordering.o &
gcc -O2 -std=c99 -fPIC -Wall -Iinclude -c $SRC/qpdo_api.c -o build/qpdo_api.o &
gcc -O2 -std=c99 -fPIC -Wall -Iinclude -c $SRC/qpdo_io.c -o build/qpdo_io.o &
wait
$NVCC $ARCH -shared -o "$OUT/libqpdo_b200.so" build/*.o -lcudart_static -lpthread -ldl -lrt -lm
echo "built $OUT/libqpdo_b200.so"
