#!/bin/bash
# builds kernel-variant libraries into tools/probes/variants/: build_variants.sh name1 "flags1" name2 "flags2" ...
set -e
mkdir -p tools/probes/variants
rm -f tools/probes/variants/lib_*.so
while [ $# -gt 1 ]; do
  FF_EXTRA_NVCC_FLAGS="$2" python gpupathtracer_b200/build.py --force > /dev/null
  cp gpupathtracer_b200/libfirefly_b200.so tools/probes/variants/lib_$1.so
  shift 2
done
python gpupathtracer_b200/build.py --force > /dev/null
