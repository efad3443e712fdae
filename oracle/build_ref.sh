#!/usr/bin/env bash
# Compiles the REFERENCE's own map sources, where they lie under /root/reference, against the
# minimal MRPT stand-in (shim/mrpt) into oracle/_ref/libmmo_ref.so. Test infrastructure only.
# Nothing is copied from the reference; outputs go to oracle/_ref/ (git-ignored, ships to the GPU box).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(dirname "$HERE")"
REF="${MOLA_REFERENCE:-/root/reference}/mola_metric_maps"
OUT="$HERE/_ref"
mkdir -p "$OUT"
FLAGS="-std=c++17 -O2 -ffp-contract=off -fPIC -w -I$ROOT/shim -I$REF/include -I$REF/3rdparty/robin-map/include -I$HERE"
for f in HashedVoxelPointCloud SparseVoxelPointCloud SparseTreesPointCloud; do
  g++ $FLAGS -c "$REF/src/$f.cpp" -o "$OUT/$f.o" &
done
g++ $FLAGS -c "$HERE/ref_bridge.cpp" -o "$OUT/ref_bridge.o" &
wait
g++ -shared -o "$OUT/libmmo_ref.so" "$OUT"/HashedVoxelPointCloud.o "$OUT"/SparseVoxelPointCloud.o "$OUT"/SparseTreesPointCloud.o "$OUT/ref_bridge.o" -lpthread
rm -f "$OUT"/*.o
echo "built $OUT/libmmo_ref.so"
