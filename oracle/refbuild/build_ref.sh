#!/usr/bin/env bash
# build_ref.sh -- compile the REFERENCE's own sources for the light-culling path, where they
# lie under /root/reference, into oracle/_ref/libhvcf_ref.so.  TEST INFRASTRUCTURE ONLY.
#
#   * HelloVulkan/Source/Camera.cpp           (the UBO producer; real C++ of the reference)
#   * Shaders/ClusteredForward/AABBGenerator.comp, LightCulling.comp, Shaders/FrustumCulling.comp (+ the .glsl they
#     include), Shaders/ClusteredForward/Radiance.glsl, Shaders/PBR/PBRHeader.glsl and the light-loop lines of
#     Shaders/ClusteredForward/Scene.frag
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
	# FrustumCulling.comp
	echo '#define gl_GlobalInvocationID gl.gl_GlobalInvocationID'
	echo 'namespace hvref_frustumcull { using namespace hvref_glsl; static thread_local Builtins gl;'
	glsl_to_cxx "$HV/Shaders/FrustumCulling.comp"
	echo '}'
	echo '#undef gl_GlobalInvocationID'
	echo '#undef main'
	echo '#undef gl_WorkGroupID'
	echo '#undef gl_NumWorkGroups'
	# The consumer: Scene.frag cannot be compiled whole (textures, buffer references), so the
	# lines of the path are streamed by number into a function body; Radiance.glsl and PBRHeader.glsl compile whole.
	SF="$HV/Shaders/ClusteredForward/Scene.frag"
	echo 'namespace hvref_scenefrag { using namespace hvref_glsl;'
	cat "$HV/Shaders/LightData.glsl"; echo
	cat "$HV/Shaders/ClusteredForward/Header.glsl"; echo
	cat "$HV/Shaders/PBR/PBRPushConstants.glsl"; echo
	echo 'struct CameraPositionOnly { vec4 position; };'
	echo 'static thread_local PBRPushConstant pc; static thread_local vec3 worldPos; static thread_local vec4 gl_FragCoord;'
	echo 'static thread_local CameraPositionOnly camUBO; static ClusterForwardUBO cfUBO;'
	echo 'static SSBO<LightData> lights; static SSBO<LightCell> lightCells; static SSBO<uint> lightIndices;'
	cat "$HV/Shaders/PBR/PBRHeader.glsl"; echo
	glsl_to_cxx "$HV/Shaders/ClusteredForward/Radiance.glsl"; echo
	sed -n '55,58p' "$SF"
	echo 'static void fragment_light_loop(vec3 albedo, vec3 N, float metallic, float roughness, vec3& LoOut, uint& clusterIdxOut) {'
	sed -n '64,78p' "$SF"
	sed -n '103p;107,108p;112,113p;117p;119,134p' "$SF" | sed -E 's/\.xyz\b/.xyz()/g'
	echo 'LoOut = Lo; clusterIdxOut = clusterIdx; }'
	echo '}'
	cat "$HERE/ref_api.inc"
} | "$CXX" "${COMMON[@]}" -fsingle-precision-constant -DGLM_FORCE_SWIZZLE -fopenmp -I"$HERE" -x c++ -c - -o "$OUT/ref_shaders.o"

"$CXX" -shared -fopenmp -o "$OUT/libhvcf_ref.so" "$OUT/ref_shaders.o" "$OUT/ref_camera.o"
rm -f "$OUT/ref_shaders.o" "$OUT/ref_camera.o"
echo "built $OUT/libhvcf_ref.so"
