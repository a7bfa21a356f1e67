#!/usr/bin/env bash
# Build the UNMODIFIED reference (andrew-pa/plutracer) render path into oracle/_ref/plu_ref.
#
# TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path.
#
# The reference sources are compiled from where they lie under $PLU_REFERENCE (default
# /root/reference); nothing is copied into the repository.  Outputs go to oracle/_ref/ only
# (git-ignored, not gpurun-ignored, so the binary travels to the GPU box).
#
# The reference was written for MSVC; it needs a small include shim on Linux:
#   * backslash include paths  (scene.h:4,6,10, triangle.cpp:1) -> symlinks whose NAME holds a '\'
#   * tiny_obj_loader.h is included but absent and unused        -> empty file
#   * missing std includes (memalign, uint32_t, FLT_MAX, strcmp) -> oracle/ref_prelude.h
#   * scene.h:280 binds a temporary to a non-const lvalue ref    -> sed-patched copy in _ref/shim
# The driver (oracle/ref_driver.cpp) is a unity TU that #includes the reference .cpp files, so the
# header-level `static mt19937 RNG` (cmmn.h:240) is ONE generator and seeded runs are reproducible.
set -euo pipefail
here="$(cd "$(dirname "$0")" && pwd)"
ref="${PLU_REFERENCE:-/root/reference}"
out="$here/_ref"
if [ ! -d "$ref/src" ]; then
  echo "build_ref: reference tree $ref not present; keeping prebuilt $out/plu_ref if any" >&2
  exit 0
fi
mkdir -p "$out/shim"
shim="$out/shim"
ln -sfn "$ref/inc/surfaces/surfaces.h"  "$shim/surfaces\\surfaces.h"
# triangle.h has no include guard and the unity TU already included it: the backslash name is an empty file
rm -f "$shim/surfaces\\triangle.h"; : > "$shim/surfaces\\triangle.h"
ln -sfn "$ref/inc/lights/area_light.h"  "$shim/lights\\area_light.h"
ln -sfn "$ref/inc/textures/textures.h"  "$shim/textures\\textures.h"
: > "$shim/tiny_obj_loader.h"
# scene.h:280 `make_basic_surface(cx, cx.eval1(...).get<vector<urn::value>>(), j)` (rvalue -> T&)
sed 's|auto ss = make_basic_surface(cx, cx.eval1(objvs\[i + 1\]).get<vector<urn::value>>(), j);|auto plu_tmp_vs = cx.eval1(objvs[i + 1]).get<vector<urn::value>>(); auto ss = make_basic_surface(cx, plu_tmp_vs, j);|' \
    "$ref/inc/scene.h" > "$shim/scene.h"
grep -q plu_tmp_vs "$shim/scene.h" || { echo "build_ref: scene.h patch did not apply" >&2; exit 1; }
g++ -std=gnu++14 -O2 -fpermissive -w \
    -DPLU_REF_ROOT="\"$ref\"" \
    -include "$here/ref_prelude.h" \
    -I"$shim" -I"$ref/inc" -I"$ref/depd/glm" -I"$ref/src" -I"$here/../include" -I"$here/../plutracer_b200/csrc/host" \
    "$here/ref_driver.cpp" -o "$out/plu_ref" -lpthread
# the reference's post-process (main.cpp:76-112) + BMP writer (texture.cpp:178-194), main.cpp included as a TU
g++ -std=gnu++14 -O2 -fpermissive -w \
    -include "$here/ref_prelude.h" \
    -I"$shim" -I"$ref/inc" -I"$ref/depd/glm" -I"$ref/src" \
    "$here/ref_post_driver.cpp" -o "$out/plu_ref_post" -lpthread
# the drop-in proof: the same reference TU with plu::gpu_renderer (integration/plu_gpu_renderer.h), linked against libplugpu.so
lib="$here/../plutracer_b200/_lib"
if [ -f "$lib/libplugpu.so" ]; then
  g++ -std=gnu++14 -O2 -fpermissive -w -DPLU_WITH_GPU \
      -DPLU_REF_ROOT="\"$ref\"" \
      -include "$here/ref_prelude.h" \
      -I"$shim" -I"$ref/inc" -I"$ref/depd/glm" -I"$ref/src" -I"$here/../include" -I"$here/../plutracer_b200/csrc/host" \
      "$here/ref_driver.cpp" -o "$out/plu_ref_gpu" -L"$lib" -lplugpu -Wl,-rpath,'$ORIGIN/../../plutracer_b200/_lib' -lpthread
fi
echo "build_ref: built $out/plu_ref $out/plu_ref_post $out/plu_ref_gpu"
