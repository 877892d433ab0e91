#!/bin/bash
for shp in "2048 256 64" "2048 256 128" "4096 512 1280" "4096 5120 1280" "3001 1280 5120"; do
  timeout 40 python tools/debug_pair.py $shp 2>&1 | grep -v Warn | tail -3; echo "exit $? for $shp"
done
