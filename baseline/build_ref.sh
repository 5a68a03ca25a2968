#!/bin/bash
# Build the UNMODIFIED reference (/root/reference) for sm_100a into baseline/_ref/ (git-ignored).
# The .cu sources are copied byte for byte; only the build recipe differs from the reference's
# CMakeLists.txt: no find_package(MPI) (none on the image -> baseline/mpi_shim), no empty cuBQL submodule.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REFERENCE_DIR:-/root/reference}"
OUT="$HERE/_ref"
mkdir -p "$OUT/src"
cp "$REF/unorderedDataVariant.cu" "$REF/prePartitionedDataVariant.cu" "$OUT/src/"
rm -rf "$OUT/src/cukd"; cp -r "$REF/submodules/cudaKDTree/cukd" "$OUT/src/cukd"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I$HERE/mpi_shim -I$OUT/src --extended-lambda -w"
$NVCC $FLAGS "$OUT/src/unorderedDataVariant.cu" "$HERE/mpi_shim/mpi_shim.cpp" -o "$OUT/cudaMpiKNN_unorderedData"
$NVCC $FLAGS "$OUT/src/prePartitionedDataVariant.cu" "$HERE/mpi_shim/mpi_shim.cpp" -o "$OUT/cudaMpiKNN_prePartitionedData"
cmp "$REF/unorderedDataVariant.cu" "$OUT/src/unorderedDataVariant.cu" && cmp "$REF/prePartitionedDataVariant.cu" "$OUT/src/prePartitionedDataVariant.cu"
echo "reference built: $OUT/cudaMpiKNN_unorderedData $OUT/cudaMpiKNN_prePartitionedData"
