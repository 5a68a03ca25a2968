#!/bin/bash
# The "ref-i64" variant of SURVEY.md 6.3: the reference with its three int*int expressions widened to size_t
# (baseline/ref_i64.patch; nothing else changes), so that it can run the headline configuration at exactly one rank
# per GPU instead of being oversubscribed to stay under n*k < 2^31. A second, clearly labelled baseline next to the
# unmodified one — `bench.py --impl reference` always runs the UNMODIFIED build.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REFERENCE_DIR:-/root/reference}"
OUT="$HERE/_ref"
mkdir -p "$OUT/src_i64"
cp "$REF/unorderedDataVariant.cu" "$OUT/src_i64/"
rm -rf "$OUT/src_i64/cukd"; cp -r "$REF/submodules/cudaKDTree/cukd" "$OUT/src_i64/cukd"
sed -i 's/candidateLists+k\*tid,k,/candidateLists+size_t(k)*tid,k,/g; s/N\*k\*sizeof(uint64_t)/size_t(N)*k*sizeof(uint64_t)/' "$OUT/src_i64/unorderedDataVariant.cu"
test "$(diff "$REF/unorderedDataVariant.cu" "$OUT/src_i64/unorderedDataVariant.cu" | grep -c '^>')" = 3
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I$HERE/mpi_shim -I$OUT/src_i64 --extended-lambda -w"
$NVCC $FLAGS "$OUT/src_i64/unorderedDataVariant.cu" "$HERE/mpi_shim/mpi_shim.cpp" -o "$OUT/cudaMpiKNN_unorderedData_i64"
echo "ref-i64 built: $OUT/cudaMpiKNN_unorderedData_i64"
