#!/bin/sh
# build.sh -- compile the sm_100a shared library in-tree (called by __graft_entry__.build()).
# Every compile job is waited for by pid (a failed compile fails the build), stale objects are removed before
# they are rebuilt, every header is a dependency, and the link refuses undefined symbols.
set -e
cd "$(dirname "$0")"
SRC=qpdo_b200/csrc
OUT=qpdo_b200/lib
mkdir -p "$OUT" build
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
ARCH="-gencode arch=compute_100a,code=sm_100a"
NVFLAGS="-O3 -std=c++17 -lineinfo $ARCH -Xcompiler -fPIC -fmad=false -Iinclude -I$SRC $EXTRA_NVFLAGS"
HEADERS="$SRC/dev_internal.h $SRC/symbolic.h include/qpdo.h include/qpdo_cuda.h include/qpdo_batch.h include/qpdo_io.h build.sh"
# EXTRA_NVFLAGS changes what an object means: remember the flags the objects were built with
FLAGSIG=$(printf '%s' "$NVFLAGS" | cksum | cut -d' ' -f1)
[ -f build/.flags ] && [ "$(cat build/.flags)" = "$FLAGSIG" ] || { rm -f build/*.o; printf '%s' "$FLAGSIG" > build/.flags; }

stale() {   # stale <object> <source>: true when the object must be rebuilt
  [ ! -f "$1" ] && return 0
  [ "$2" -nt "$1" ] && return 0
  for h in $HEADERS; do [ "$h" -nt "$1" ] && return 0; done
  return 1
}

PIDS=""
for f in dev_core dev_solver factor linesearch batch peak $EXTRA_CU; do
  [ -f "$SRC/$f.cu" ] || continue
  if stale build/$f.o "$SRC/$f.cu"; then
    rm -f build/$f.o
    $NVCC $NVFLAGS -c "$SRC/$f.cu" -o build/$f.o &
    PIDS="$PIDS $!"
  fi
done
for f in symbolic ordering; do
  if stale build/$f.o "$SRC/$f.cpp"; then
    rm -f build/$f.o
    g++ -O2 -std=c++17 -fPIC -Iinclude -I$SRC -c $SRC/$f.cpp -o build/$f.o &
    PIDS="$PIDS $!"
  fi
done
for f in qpdo_api qpdo_io; do
  if stale build/$f.o "$SRC/$f.c"; then
    rm -f build/$f.o
    gcc -O2 -std=c99 -fPIC -Wall -Iinclude -c $SRC/$f.c -o build/$f.o &
    PIDS="$PIDS $!"
  fi
done
for p in $PIDS; do
  wait "$p" || { echo "build.sh: a compile job failed" >&2; exit 1; }
done
$NVCC $ARCH -shared -Xlinker --no-undefined -o "$OUT/libqpdo_b200.so" build/*.o -lcudart_static -lpthread -ldl -lrt -lm
echo "built $OUT/libqpdo_b200.so"
