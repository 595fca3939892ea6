#!/usr/bin/env bash
# md5 of the library's complete SASS: a refactor of the .cu / .cuh sources that leaves this unchanged cannot have changed any kernel
# (used for source-only moves such as csrc/common/tiling.cuh, which were made without a GPU at hand).
set -euo pipefail
LIB=${1:-$(cd "$(dirname "$0")/.." && pwd)/hyperion_b200/lib/libhyperion_b200.so}
cuobjdump -sass "$LIB" | md5sum | cut -d' ' -f1
