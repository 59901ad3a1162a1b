#!/bin/bash
# Compiles the reference's OWN C++ test of the path,
#   /root/reference/tests/grid_data/test_gridder_uvw_es_fft.cpp
# (happy paths 3-D/2-D x fp64/fp32 with its adjointness criterion + its 45
# argument-error cases), unmodified and where it lies, against THIS repo's
# headers (include/, which shadow the reference's for the four headers of the
# path) and THIS libska_sdp_func.so. The binary goes to baseline/_ref_cpp/
# (git-ignored, travels to the GPU box); tests/test_gpu_contract.py runs it.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(dirname "$HERE")"
REF="${REFERENCE_ROOT:-/root/reference}"
SRC="$REF/tests/grid_data/test_gridder_uvw_es_fft.cpp"
[ -f "$SRC" ] || { echo "no reference tree at $REF"; exit 0; }
mkdir -p "$HERE/_ref_cpp"
g++ -std=c++14 -O1 -DSDP_HAVE_CUDA -I "$ROOT/include" -I "$REF/src" "$SRC" \
    -L "$ROOT/ska-sdp-func_b200/lib" -lska_sdp_func \
    -Wl,-rpath,'$ORIGIN/../../ska-sdp-func_b200/lib' \
    -o "$HERE/_ref_cpp/test_gridder_uvw_es_fft"
ls -la "$HERE/_ref_cpp"
