#!/usr/bin/env bash
# Builds oracle/_ref/oracle_dump (test infrastructure): our dumper TU compiled against the reference's headers
# where they lie, linked with the static libraries that oracle/build_ref.sh produced in its scratch build tree.
set -euo pipefail
HERE=$(cd "$(dirname "$0")" && pwd)
WORK=${SU2_REF_WORK:-/tmp/su2ref}
MARCH=${SU2_REF_MARCH:-x86-64-v3}
if [ ! -f "$WORK/build/SU2_CFD/src/libSU2core.a" ]; then
  if [ -x "$HERE/_ref/oracle_dump" ]; then echo "no reference build tree; keeping prebuilt oracle_dump"; exit 0; fi
  echo "run oracle/build_ref.sh first"; exit 1
fi
B=$WORK/build
/usr/bin/g++ -std=c++17 -O2 -march=$MARCH -fopenmp -DNDEBUG -w \
  -I"$WORK" -I"$WORK/externals/CLI11" -I"$WORK/externals/eigen" -I"$WORK/externals/mel" \
  -I"$B/Common/src" -I"$WORK/Common/src" -I"$B/SU2_CFD/src" -I"$WORK/SU2_CFD/src" \
  "$HERE/oracle_dump.cpp" -o "$HERE/_ref/oracle_dump" \
  -Wl,--start-group "$B/SU2_CFD/src/libSU2core.a" "$B/Common/src/libSU2Common.a" -Wl,--end-group -fopenmp
strip "$HERE/_ref/oracle_dump" || true
echo "built $HERE/_ref/oracle_dump"
