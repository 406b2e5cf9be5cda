#!/bin/bash
# A/B on one box: tools/libsvo_old.so against the in-tree build (scratch copy on the GPU box only)
D=stereo_visual_odometry_b200
cp $D/libsvo_b200.so /tmp/new.so
for r in 1 2; do
  cp tools/libsvo_old.so $D/libsvo_b200.so; bash tools/quick_bench.sh old
  cp /tmp/new.so $D/libsvo_b200.so; bash tools/quick_bench.sh new
done
