#!/bin/bash
# like ab_run.sh, at several samples per pixel:  scripts/ab_run2.sh "8 64" scene...
pkg=final-stage-path-tracer-2-0_b200
spps=$1; shift
cp $pkg/libfs2cuda.so /tmp/libfs2cuda.main.so
for d in $pkg/_exp/*/; do
  n=$(basename $d)
  cp $d/libfs2cuda.so $pkg/libfs2cuda.so
  for scene in ${@:-mesh100k}; do for spp in $spps; do
    echo "== $n $scene ${spp}spp: $(python scripts/profile_trace.py wavefront $spp 1920 1080 $scene 2>&1 | grep Mpaths | tail -1)"
  done; done
done
cp /tmp/libfs2cuda.main.so $pkg/libfs2cuda.so
