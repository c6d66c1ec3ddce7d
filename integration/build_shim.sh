#!/usr/bin/env bash
# Builds SU2_CFD with the libsu2b200 binding compiled in: integration/_ref/SU2_CFD_b200 (git-ignored, NOT gpurun-ignored).
# A scratch copy of the reference under /tmp is patched (integration/apply_patch.py) and built with the reference's own
# meson + ninja, serial + OpenMP, -DHAVE_SU2B200, linked against su2_b200/libsu2b200.so (rpath relative to the binary).
# Neither /root/reference nor this repository receives any reference source.
set -euo pipefail
REF=${SU2_REFERENCE:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
REPO=$(dirname "$HERE")
OUT=$HERE/_ref
WORK=${SU2_SHIM_WORK:-/tmp/su2ref_b200}
MARCH=${SU2_REF_MARCH:-x86-64-v3}
if [ ! -d "$REF" ]; then echo "no reference tree at $REF (GPU box?) - using prebuilt $OUT"; exit 0; fi
mkdir -p "$OUT"
if [ -x "$OUT/SU2_CFD_b200" ] && [ -z "${SU2_SHIM_FORCE:-}" ] && [ "$OUT/SU2_CFD_b200" -nt "$HERE/CB200FlowBackend.hpp" ] \
   && [ "$OUT/SU2_CFD_b200" -nt "$HERE/apply_patch.py" ] && [ "$OUT/SU2_CFD_b200" -nt "$REPO/include/su2b200.h" ]; then echo "integration/_ref/SU2_CFD_b200 is up to date"; exit 0; fi
[ -f "$REPO/su2_b200/libsu2b200.so" ] || { echo "build su2_b200/libsu2b200.so first"; exit 1; }
if [ ! -d "$WORK/SU2_CFD" ]; then
  mkdir -p "$WORK"
  # sources only: the oracle's scratch tree may hold a build directory with absolute paths
  (cd "$REF" && tar cf - --exclude=./build .) | (cd "$WORK" && tar xf -)
  chmod -R u+w "$WORK"
fi
python3 "$HERE/apply_patch.py" "$WORK" "$REPO"
cd "$WORK"
touch su2preconfig.timestamp
export CC=/usr/bin/gcc CXX=/usr/bin/g++
if [ ! -f build/build.ninja ]; then
  python3 externals/meson/meson.py setup build --buildtype=release \
    -Dwith-mpi=disabled -Dwith-omp=true -Denable-cgns=false -Denable-tecio=false \
    -Dcpp_args="-march=$MARCH -DHAVE_SU2B200" \
    -Dcpp_link_args="-L$REPO/su2_b200 -lsu2b200 -Wl,-rpath,\$ORIGIN/../../su2_b200 -Wl,--no-as-needed"
fi
ninja -C build SU2_CFD/src/SU2_CFD
cp build/SU2_CFD/src/SU2_CFD "$OUT/SU2_CFD_b200"
strip "$OUT/SU2_CFD_b200" || true
echo "built $OUT/SU2_CFD_b200"
