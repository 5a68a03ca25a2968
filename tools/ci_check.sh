#!/usr/bin/env bash
# What CI runs (.github/workflows/build.yml); also handy before a commit.
#   tools/ci_check.sh build   compile the library, the executables, the C example and the host emulation of the
#                             kernels for sm_100a (no GPU needed), list every kernel's registers / spills, and
#                             check that the tensor-core experiment still contains tcgen05 / TMA instructions
#   tools/ci_check.sh test    the CPU test tier (pytest -m "not gpu")
set -euo pipefail
cd "$(dirname "$0")/.."
case "${1:-build}" in
  build)
    make -j"$(nproc)" all examples simt
    echo "== kernels: registers / spills"
    grep -h -A2 "Compiling entry function" build/*.ptxas.log | grep -E "Used|spill" | paste - - | \
      awk '{print $0}' | sort | uniq -c | sort -rn | head -60
    if grep -l "bytes spill stores" build/*.ptxas.log | xargs grep -h "spill stores" | grep -v " 0 bytes spill stores" | grep -q .; then
      echo "note: some kernels spill (listed above)"
    fi
    echo "== tensor-core / TMA mnemonics in brute_tile"
    cuobjdump -sass build/brute_tile.o | grep -cE "UTCHMMA|LDTM|UTMALDG" | xargs -I{} test {} -ge 3
    echo "build ok"
    ;;
  test)
    python3 -m pytest tests -x -q -m "not gpu"
    ;;
  *) echo "usage: $0 build|test" >&2; exit 2 ;;
esac
