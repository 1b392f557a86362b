#!/usr/bin/env bash
# oracle/build_ref.sh — builds oracle/_ref/libsplat_ref.so from the reference's OWN sources,
# read where they lie under $REF (default /root/reference).  TEST INFRASTRUCTURE.
#
# Nothing from the reference is copied into the repo: the translation unit exists only in a
# pipe.  It is the concatenation of
#   a 6-line prologue (runtime image-geometry variables),
#   include/math/sincos.hpp, include/splat/ipu_geometry.hpp, include/tileMapping/tile_config.hpp,
#   include/splat/viewport.hpp, codelets/splat/codelets.cpp lines 1-642 (the live GSplat vertex;
#   lines 644-871 are the unwired AMP vertices that use IPU inline asm),
#   oracle/ref_driver.cpp (ours).
# Edits applied in the pipe, and nothing else:
#   * the five absolute '#include </home/nf20/workspace/gaussian_splat_ipu/...>' lines are
#     dropped (their targets are concatenated instead) — codelets.cpp:9-11, ipu_geometry.hpp:7,
#     tile_config.hpp:3;
#   * '#pragma once' lines are dropped (single TU);
#   * the four geometry macros of tile_config.hpp:5-8 (1280/720/40/36) become runtime floats so
#     one binary serves every BASELINE.json resolution (all are exact in f32).
# Flags: -O2, x86-64 baseline, -ffp-contract=off => no FMA contraction; arithmetic order literal.
set -euo pipefail
REF="${REF:-/root/reference}"
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
mkdir -p "$OUT"
if [ ! -f "$REF/codelets/splat/codelets.cpp" ]; then
  echo "build_ref.sh: reference not present at $REF; keeping any prebuilt $OUT/libsplat_ref.so" >&2
  exit 0
fi
strip_abs='/#include <\/home\/nf20\//d; /^#pragma once/d'
{
  cat <<'PRO'
#include <cstdint>
#include <array>
float g_ref_imwidth = 1280.0f, g_ref_imheight = 720.0f;
float g_ref_tiles_across = 40.0f, g_ref_tiles_down = 36.0f;
#include <poplar/Vertex.hpp>
PRO
  sed -e "$strip_abs" "$REF/include/math/sincos.hpp"
  sed -e "$strip_abs" "$REF/include/splat/ipu_geometry.hpp"
  sed -e "$strip_abs" \
      -e 's/^#define IMWIDTH .*/#define IMWIDTH g_ref_imwidth/' \
      -e 's/^#define IMHEIGHT .*/#define IMHEIGHT g_ref_imheight/' \
      -e 's/^#define TILES_ACCROSS .*/#define TILES_ACCROSS g_ref_tiles_across/' \
      -e 's/^#define TILES_DOWN .*/#define TILES_DOWN g_ref_tiles_down/' \
      "$REF/include/tileMapping/tile_config.hpp"
  echo
  sed -e "$strip_abs" "$REF/include/splat/viewport.hpp"
  echo
  head -642 "$REF/codelets/splat/codelets.cpp" | sed -e "$strip_abs"
  echo
  cat "$HERE/ref_driver.cpp"
} | ${HOSTCXX:-/usr/bin/g++} -x c++ -std=c++20 -O2 -ffp-contract=off -fopenmp -fPIC -shared -w \
        -I "$HERE/poplar_stub" -I "$REF/external/glm" -I "$REF/include/math" \
        -o "$OUT/libsplat_ref.so" -
# Second library: the reference's host-side INPUT code (PLY loading through happly, camera, frustum),
# compiled UNCHANGED from its own translation units, plus oracle/ref_host_driver.cpp (ours).
${HOSTCXX:-/usr/bin/g++} -std=c++20 -O2 -ffp-contract=off -fPIC -shared -w \
    -I "$REF/include" -I "$REF/external/glm" \
    "$REF/src/splat/file_io.cpp" "$REF/src/splat/camera.cpp" "$REF/src/splat/geometry.cpp" \
    "$HERE/ref_host_driver.cpp" -o "$OUT/libhost_ref.so"
# scene inputs (data, not source) so PLY-scene parity tests can run where /root/reference is absent
mkdir -p "$OUT/data"
for p in "$REF"/data/point_cloud_*.ply; do
  [ -f "$OUT/data/$(basename "$p")" ] || cp "$p" "$OUT/data/"
done
echo "built $OUT/libsplat_ref.so"
