#!/bin/bash
# GPU box: `rt file.svati out.png` in a FRESH process (no other process holds the device, so
# driver + context start-up is paid in full), with the phase lines of RT_STATS / RTB_TRACE /
# RTB_PARSE_TIMING.  usage: bash tools/cli_probe.sh [grid]
cd "$(dirname "$0")/.."
GRID=${1:-2237}
python - <<PY
import sys
sys.path.insert(0,'raytracing-gpu_b200'); sys.path.insert(0,'tests')
import pyrtb
from util import load_golden
load_golden("car-on-road")[0].with_resolution(3840,2160).write_svati('/tmp/c4.svati')
pyrtb.HostScene.terrain($GRID, 3840, 2160).write_svati('/tmp/c5.svati')
PY
ls -la /tmp/c4.svati /tmp/c5.svati
for f in c4 c5; do
  for i in 1 2; do
    echo "== $f run $i"
    t0=$(date +%s.%N)
    RT_STATS=1 RTB_TRACE=1 RTB_PARSE_TIMING=1 ./raytracing-gpu_b200/rt /tmp/$f.svati /tmp/o.png 2>&1 | grep -E "rtb_|parse|rror"
    python3 -c "import sys,time; print(\"wall %.3f s\" % (time.time() - float(sys.argv[1])))" $t0
  done
done
rm -f /tmp/c4.svati /tmp/c5.svati /tmp/o.png
