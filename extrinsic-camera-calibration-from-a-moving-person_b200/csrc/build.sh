#!/bin/bash
# Build libbax.so in-tree for sm_100a (nvcc cross-compiles without a GPU).  Translation units compile in parallel.
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
LOGDIR=../../profiles/ptxas
mkdir -p $LOGDIR
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xptxas -v"
OBJS=""
PIDS=""
for f in bax_problem bax_residual bax_lm bax_trial bax_syrk bax_dense bax_extra bax_trf bax_batch bax_comm rsx rsx_final rsx_static $EXTRA_SRCS; do
  if [ ! -f $f.o ] || [ $f.cu -nt $f.o ] || [ bax_internal.cuh -nt $f.o ] || [ rsx_internal.cuh -nt $f.o ] || [ rsx_fit.cuh -nt $f.o ] || [ rsx_arrow.cuh -nt $f.o ] || [ rsx_handle.cuh -nt $f.o ] || [ ../../include/rsx.h -nt $f.o ] || [ ../../include/bax.h -nt $f.o ]; then
    ( $NVCC $FLAGS -c $f.cu -o $f.o.tmp 2> $f.ptxas.raw && mv $f.o.tmp $f.o || { cat $f.ptxas.raw; rm -f $f.o.tmp $f.ptxas.raw; exit 1; }; grep -v "Compile time" $f.ptxas.raw > $LOGDIR/$f.ptxas.log; rm -f $f.ptxas.raw ) &
    PIDS="$PIDS $!"
  fi
  OBJS="$OBJS $f.o"
done
for p in $PIDS; do wait $p; done
$NVCC -shared -gencode arch=compute_100a,code=sm_100a -o ../libbax.so $OBJS -lcudart -ldl
echo "built $(cd ..; pwd)/libbax.so"
