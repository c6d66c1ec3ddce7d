#!/usr/bin/env bash
# Builds the UNMODIFIED reference (su2code/SU2) CPU solver from the sources where they lie
# under /root/reference into oracle/_ref/ (git-ignored, NOT gpurun-ignored).
# This is TEST INFRASTRUCTURE (the parity oracle + the timed CPU baseline), never the product.
# The reference tree is read-only, so it is copied to a scratch dir first; only the built
# binaries land in oracle/_ref/. No reference source is copied into the repository.
#
# Recipe (SURVEY.md header): meson (vendored in the reference) + ninja, serial+OpenMP, no MPI.
set -euo pipefail
REF=${SU2_REFERENCE:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
WORK=${SU2_REF_WORK:-/tmp/su2ref}
MARCH=${SU2_REF_MARCH:-x86-64-v3}   # AVX2: simd::Array<double> width 4; portable across the pool's hosts
if [ ! -d "$REF" ]; then echo "no reference tree at $REF (GPU box?) - using prebuilt $OUT"; exit 0; fi
mkdir -p "$OUT"
if [ -x "$OUT/SU2_CFD" ] && [ -z "${SU2_REF_FORCE:-}" ]; then echo "oracle/_ref/SU2_CFD exists"; exit 0; fi
if [ ! -d "$WORK/SU2_CFD" ]; then
  mkdir -p "$WORK"; cp -r "$REF"/. "$WORK"/; chmod -R u+w "$WORK"
fi
cd "$WORK"
touch su2preconfig.timestamp
export CC=/usr/bin/gcc CXX=/usr/bin/g++
if [ ! -f build/build.ninja ]; then
  python3 externals/meson/meson.py setup build --buildtype=release \
    -Dwith-mpi=disabled -Dwith-omp=true -Denable-cgns=false -Denable-tecio=false \
    -Dcpp_args=-march=$MARCH
fi
ninja -C build SU2_CFD/src/SU2_CFD
cp build/SU2_CFD/src/SU2_CFD "$OUT/SU2_CFD"
strip "$OUT/SU2_CFD" || true
echo "built $OUT/SU2_CFD"
