#!/bin/bash
# GPU side of the A/B helper: for every variant under _exp/, time config 2 (1080p, 64 spp) and
# optionally other scenes.   scripts/ab_run.sh [scene ...]
pkg=final-stage-path-tracer-2-0_b200
cp $pkg/libfs2cuda.so /tmp/libfs2cuda.main.so
for d in $pkg/_exp/*/; do
  n=$(basename $d)
  cp $d/libfs2cuda.so $pkg/libfs2cuda.so
  for scene in ${@:-mesh100k}; do
    echo "== $n ($(cat $d/flags.txt)) $scene: $(python scripts/profile_trace.py wavefront 64 1920 1080 $scene 2>&1 | grep Mpaths | tail -1)"
  done
done
cp /tmp/libfs2cuda.main.so $pkg/libfs2cuda.so
