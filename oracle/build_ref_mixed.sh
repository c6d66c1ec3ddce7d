#!/usr/bin/env bash
# The reference built with -Denable-mixedprec=true (su2mixedfloat = float: Jacobian, preconditioner and every Krylov
# vector in single precision; code_config.hpp:133-139) -> oracle/_ref/SU2_CFD_mixed.  TEST INFRASTRUCTURE: the oracle of
# the mixed-precision linear-solve mode.  Same recipe as build_ref.sh, second build directory of the same
# scratch copy.
set -euo pipefail
REF=${SU2_REFERENCE:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
WORK=${SU2_REF_WORK:-/tmp/su2ref}
MARCH=${SU2_REF_MARCH:-x86-64-v3}
if [ ! -d "$REF" ]; then echo "no reference tree at $REF (GPU box?) - using prebuilt $OUT"; exit 0; fi
mkdir -p "$OUT"
if [ -x "$OUT/SU2_CFD_mixed" ] && [ -z "${SU2_REF_FORCE:-}" ]; then echo "oracle/_ref/SU2_CFD_mixed exists"; exit 0; fi
if [ ! -d "$WORK/SU2_CFD" ]; then
  mkdir -p "$WORK"; cp -r "$REF"/. "$WORK"/; chmod -R u+w "$WORK"
fi
cd "$WORK"
touch su2preconfig.timestamp
export CC=/usr/bin/gcc CXX=/usr/bin/g++
if [ ! -f build_mixed/build.ninja ]; then
  python3 externals/meson/meson.py setup build_mixed --buildtype=release \
    -Dwith-mpi=disabled -Dwith-omp=true -Denable-cgns=false -Denable-tecio=false -Denable-mixedprec=true \
    -Dcpp_args=-march=$MARCH
fi
ninja -C build_mixed SU2_CFD/src/SU2_CFD
cp build_mixed/SU2_CFD/src/SU2_CFD "$OUT/SU2_CFD_mixed"
strip "$OUT/SU2_CFD_mixed" || true
echo "built $OUT/SU2_CFD_mixed"
