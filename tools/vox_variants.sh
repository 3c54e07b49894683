#!/bin/bash
# scratch: run the build probe with each prebuilt voxelizer occupancy variant (build_variants/lib_<emit>_<count>.so)
cp massivevoxelraytracing_b200/libhako_b200.so /tmp/lib_orig.so
for f in build_variants/lib_*.so; do
  cp $f massivevoxelraytracing_b200/libhako_b200.so
  echo "== $f"
  python tools/build_device_probe.py 4096
done
cp /tmp/lib_orig.so massivevoxelraytracing_b200/libhako_b200.so
