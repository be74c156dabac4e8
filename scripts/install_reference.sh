#!/bin/bash
# Installs the UNMODIFIED reference rasterizer package into baseline/_ref (git-ignored; travels to the GPU box).
# The reference tree is read-only, so the build runs from a copy under /tmp.  Takes ~9 min (4 sm_100a cubins).
set -e
cd "$(dirname "$0")/.."
SRC=/root/reference/submodules/gaus_2dgs_rasterization
[ -d "$SRC" ] || { echo "no reference checkout at $SRC"; exit 1; }
rm -rf /tmp/rast_copy && cp -r "$SRC" /tmp/rast_copy
mkdir -p baseline/_ref
TORCH_CUDA_ARCH_LIST=10.0a NVCC_APPEND_FLAGS="-include cstdint" MAX_JOBS=8 \
  python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
  --target baseline/_ref /tmp/rast_copy
ls baseline/_ref
