#!/bin/bash
# build times of every library under tools/probes/variants/ (1 M and 20 M triangles, second commit on a warm context)
for f in tools/probes/variants/lib_*.so; do
  echo "== $f"
  for n in 1000000 20000000; do FF_LIB_OVERRIDE=$PWD/$f python tools/probes/build1m.py $n 2>&1 | tail -1; done
done
