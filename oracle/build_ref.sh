#!/usr/bin/env bash
# TEST INFRASTRUCTURE: build the UNMODIFIED reference (dirac-institute/kbmod @ 617cd33) hot path from the
# sources where they lie under /root/reference into oracle/_ref/ (git-ignored, travels with gpurun).
#   oracle/_ref/search.cpython-*.so       -- the reference's pybind11 module `search` (HAVE_CUDA, HAVE_OPENMP)
#   (the same object also holds module `ref_harness`; oracle/ref.py loads it by PyInit name)
# We do not run the reference's cmake build; the path is four source files.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${KBMOD_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
SRC="$REF/src/kbmod/search"
if [ ! -d "$SRC" ]; then echo "reference not present at $REF; keeping prebuilt oracle/_ref" >&2; exit 0; fi
mkdir -p "$OUT/obj"
PYINC="$(python3 -c 'import sysconfig; print(sysconfig.get_paths()["include"])')"
EXT="$(python3 -c 'import sysconfig; print(sysconfig.get_config_var("EXT_SUFFIX"))')"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
CUFLAGS="-std=c++17 -O3 -rdc=true -DHAVE_CUDA=1 -DHAVE_OPENMP=1 -Xcompiler -fPIC,-fopenmp -gencode arch=compute_100a,code=sm_100a -I$REF/include -I$REF/include/eigen"
for f in kernels image_kernels kernel_memory; do
  $NVCC $CUFLAGS -c "$SRC/kernels/$f.cu" -o "$OUT/obj/$f.o" &
done
g++ -O3 -std=c++17 -fPIC -fvisibility=hidden -fopenmp -DHAVE_OPENMP=1 -DHAVE_CUDA=1 \
    -I"$SRC" -I"$REF/include" -I"$REF/include/eigen" -I"$REF/lib/pybind11/include" -I"$PYINC" \
    -I/usr/local/cuda/include -c "$HERE/ref_harness.cpp" -o "$OUT/obj/ref_harness.o" &
wait
$NVCC -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -dlink \
    "$OUT/obj/kernels.o" "$OUT/obj/image_kernels.o" "$OUT/obj/kernel_memory.o" -o "$OUT/obj/dlink.o"
g++ -shared -fopenmp -o "$OUT/search$EXT" "$OUT/obj/ref_harness.o" "$OUT/obj/kernels.o" \
    "$OUT/obj/image_kernels.o" "$OUT/obj/kernel_memory.o" "$OUT/obj/dlink.o" \
    -L/usr/local/cuda/lib64 -lcudart_static -lcudadevrt -lrt -ldl -lpthread

rm -rf "$OUT/obj"
echo "built $OUT/search$EXT"
