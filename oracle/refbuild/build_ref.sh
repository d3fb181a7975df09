#!/usr/bin/env bash
# build_ref.sh -- compile the REFERENCE's own sources for the light-culling path, where they
# lie under /root/reference, into oracle/_ref/libhvcf_ref.so.  TEST INFRASTRUCTURE ONLY.
#
#   * HelloVulkan/Source/Camera.cpp           (the UBO producer; real C++ of the reference)
#   * Shaders/ClusteredForward/AABBGenerator.comp, LightCulling.comp (+ the .glsl they include)
#     compiled as C++ against the reference's vendored glm: the shader text is streamed through
#     sed (drops `#version`, turns `layout(...) buffer/uniform` blocks into plain globals) straight
#     into g++ on stdin -- no reference source is copied into this repo or into oracle/_ref/.
#
# -ffp-contract=off: IEEE binary32, no FMA contraction (the documented oracle semantics).
# -fsingle-precision-constant on the shader TU: GLSL literals like `1.0` are 32-bit floats.
# Skips quietly (exit 0) when /root/reference is absent (the GPU box uses the prebuilt .so).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${HVCF_REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/../_ref"
if [ ! -d "$REF/HelloVulkan/Shaders/ClusteredForward" ]; then
	echo "build_ref.sh: $REF not present; keeping prebuilt oracle/_ref" >&2
	exit 0
fi
mkdir -p "$OUT"
HV="$REF/HelloVulkan"
CXX="${HVCF_CXX:-/usr/bin/g++}"  # not $CXX: the image exports a wrapper without libgomp
COMMON=(-std=c++20 -O2 -fPIC -ffp-contract=off -fvisibility=hidden -DGLM_FORCE_DEPTH_ZERO_TO_ONE
	-I"$HERE/shim" -I"$HV/Header" -I"$REF/External/glm" -I"$HV/Shaders")

glsl_to_cxx() {
	sed -E \
		-e '/^#version/d' \
		-e 's/\.xyz\b/.xyz()/g' \
		-e '/^layout *\(local_size/d' \
		-e 's/^layout *\([^)]*\) *(readonly +)?buffer +[A-Za-z_]+ *\{ *([A-Za-z_]+) +([A-Za-z_]+) *\[ *\] *; *\} *;/static SSBO<\2> \3;/' \
		-e 's/^layout *\([^)]*\) *(readonly +)?buffer +[A-Za-z_]+ *\{ *([A-Za-z_]+) +([A-Za-z_]+) *; *\} *;/static \2* \3_ptr;\n#define \3 (*\3_ptr)/' \
		-e 's/^layout *\([^)]*\) *uniform +[A-Za-z_]+ *\{ *([A-Za-z_]+) +([A-Za-z_]+) *; *\} *;/static \1 \2;/' \
		"$1"
}

"$CXX" "${COMMON[@]}" -c "$HV/Source/Camera.cpp" -o "$OUT/ref_camera.o"

{
	echo '#include "glsl_env.hpp"'
	echo '#define main shader_main'
	echo '#define gl_WorkGroupID gl.gl_WorkGroupID'
	echo '#define gl_NumWorkGroups gl.gl_NumWorkGroups'
	echo 'namespace hvref_aabbgen { using namespace hvref_glsl; static thread_local Builtins gl;'
	glsl_to_cxx "$HV/Shaders/ClusteredForward/AABBGenerator.comp"
	echo '}'
	echo 'namespace hvref_lightcull { using namespace hvref_glsl; static thread_local Builtins gl;'
	glsl_to_cxx "$HV/Shaders/ClusteredForward/LightCulling.comp"
	echo '}'
	echo '#undef main'
	echo '#undef gl_WorkGroupID'
	echo '#undef gl_NumWorkGroups'
	cat "$HERE/ref_api.inc"
} | "$CXX" "${COMMON[@]}" -fsingle-precision-constant -DGLM_FORCE_SWIZZLE -fopenmp -I"$HERE" -x c++ -c - -o "$OUT/ref_shaders.o"

"$CXX" -shared -fopenmp -o "$OUT/libhvcf_ref.so" "$OUT/ref_shaders.o" "$OUT/ref_camera.o"
rm -f "$OUT/ref_shaders.o" "$OUT/ref_camera.o"
echo "built $OUT/libhvcf_ref.so"
