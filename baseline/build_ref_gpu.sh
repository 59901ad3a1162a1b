#!/bin/bash
# Installs the UNMODIFIED reference (ska-sdp-func: C++/CUDA library + Python
# package) into baseline/_ref with its own build system (setup.py -> CMake),
# device code for sm_100 (the reference's CMake stops at compute_90, so the
# architecture list is passed through CMAKE_ARGS).
# baseline/_ref is git-ignored, not gpurun-ignored: it travels to the GPU box,
# where bench.py --impl reference-gpu and tests/test_ref_gpu.py drive it
# through its own public API. Nothing from it is used by the product.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REFERENCE_ROOT:-/root/reference}"
[ -d "$REF" ] || { echo "no reference tree at $REF"; exit 0; }
WORK="$(mktemp -d /tmp/ref_src.XXXXXX)"
cp -r "$REF"/. "$WORK"/          # the tree is read-only; the build writes into it
rm -rf "$HERE/_ref"
CMAKE_ARGS="-DCUDA_ARCH=9.0 -DCMAKE_CUDA_ARCHITECTURES=100-real -DFIND_MKL=OFF -DBUILD_TESTING=OFF" \
python -m pip install --no-index --no-build-isolation --no-deps \
    --find-links /opt/wheelhouse --target "$HERE/_ref" "$WORK"
rm -rf "$WORK"
ls "$HERE/_ref"
