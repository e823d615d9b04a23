#!/bin/bash
# oracle/make_ref.sh [reference-root] -- TEST INFRASTRUCTURE.  Stages the hot-path modules of the UNMODIFIED reference under
# oracle/_ref/ (git-ignored, so it never enters history; NOT gpurun-ignored, so it travels to the GPU box like the built .so
# files) for bench.py --impl reference / cpu_baseline.kind "reference": go1_gym, go1_gym_learn, scripts/train.py (its argparse and
# config blocks are exec'd by oracle/ref_harness.py) and the actuator network.  Nothing is edited; the product never reads it.
set -e
SRC=${1:-/root/reference}
DST="$(cd "$(dirname "$0")" && pwd)/_ref"
if [ ! -d "$SRC/go1_gym" ]; then
  echo "make_ref: $SRC has no go1_gym: nothing staged (bench.py falls back to oracle/port.py)"
  exit 0
fi
rm -rf "$DST"
mkdir -p "$DST/scripts" "$DST/resources"
cp -r "$SRC/go1_gym" "$SRC/go1_gym_learn" "$DST/"
cp "$SRC/scripts/train.py" "$DST/scripts/"
cp -r "$SRC/resources/actuator_nets" "$DST/resources/"
find "$DST" -name "__pycache__" -type d -prune -exec rm -rf {} +
echo "make_ref: staged $(du -sh "$DST" | cut -f1) under $DST"
