#!/usr/bin/env bash
# Stages the UNMODIFIED reference (python package legged_gym + the two actuator-net checkpoints its task classes
# load at construction) from /root/reference into oracle/_ref/, which is git-ignored but travels to the GPU box
# with the working tree.  bench.py's CPU arm (`--impl reference`, `cpu_baseline`) imports it from there over the
# fake-Isaac-Gym shim (oracle/refshim) and times the reference's own LeggedRobot.step().  TEST / BENCH
# INFRASTRUCTURE: nothing under legged_gym_b200/ reads it, no reference source enters the git history.
#   usage: oracle/make_ref.sh [reference root, default /root/reference]
set -euo pipefail
SRC=${1:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
DST="$HERE/_ref"
[ -f "$SRC/legged_gym/envs/base/legged_robot.py" ] || { echo "no reference under $SRC"; exit 1; }
rm -rf "$DST"
mkdir -p "$DST/resources/actuator_nets"
cp -r "$SRC/legged_gym" "$DST/legged_gym"
cp "$SRC"/resources/actuator_nets/*.pt "$DST/resources/actuator_nets/"
[ -f "$SRC/LICENSE" ] && cp "$SRC/LICENSE" "$DST/LICENSE"
find "$DST" -name __pycache__ -type d -prune -exec rm -rf {} +
(cd "$SRC" && find legged_gym resources/actuator_nets -type f ! -name '*.pyc' | LC_ALL=C sort | xargs sha256sum) > "$DST/SHA256SUMS"
echo "staged $(find "$DST/legged_gym" -name '*.py' | wc -l) python files + checkpoints -> $DST"
