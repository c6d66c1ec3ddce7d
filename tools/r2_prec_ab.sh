#!/bin/bash
cd "$(dirname "$0")/.."
run() { echo "== $1"; shift; env "$@" python tools/bench_prec.py 185 5 2>&1 | grep "sys=0"; }
run "U=1" SU2B200_SWEEP_U=1
run "U=1 NOFILL (no dependency waits: streaming bound of this layout)" SU2B200_SWEEP_U=1 SU2B200_SWEEP_NOFILL=1
run "U=2 NOFILL" SU2B200_SWEEP_U=2 SU2B200_SWEEP_NOFILL=1
run "U=1 NOFILL all rows in the wide kernel" SU2B200_SWEEP_U=1 SU2B200_SWEEP_NOFILL=1 SU2B200_SWEEP_WIDE=1
