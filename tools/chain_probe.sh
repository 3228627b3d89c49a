#!/bin/bash
# How fast does ONE pixel's sequential sample chain run at 1 / 2 / 3 / 4 CTAs per SM?  A 1/64 (1/128) tile shard of C2 gives every pixel
# its own lane, so the frame time is the longest chain (chunks = 1: no hand-offs).
for mb in 1 2 3 4; do
  python tools/shard_probe.py c2 1 0 $mb 0 64,128 | grep -v '"world": 1,'
done
