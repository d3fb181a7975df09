// glsl_env.hpp -- the minimum GLSL-on-C++ environment needed to compile the reference's two
// compute shaders (HelloVulkan/Shaders/ClusteredForward/{AABBGenerator,LightCulling}.comp)
// AS THEY LIE under /root/reference, with the reference's own vendored glm supplying
// vec/mat types and arithmetic.  TEST INFRASTRUCTURE ONLY (see oracle/README.md).
#pragma once
#include <cstdint>
#include <cmath>
#include "glm/glm.hpp"

namespace hvref_glsl
{
using namespace glm; // vec2/3/4, mat4, uvec3, uint, min, max, pow

// std430 runtime-sized array: `T name[];` with name.length() and name[i]
template <typename T>
struct SSBO
{
	T* p = nullptr;
	uint n = 0;
	// std430 array stride in bytes; differs from sizeof(T) when the C++ struct lacks the GLSL
	// tail padding (LightData: 36 B of members, 48 B stride -- Shaders/LightData.glsl:1-6)
	size_t stride = sizeof(T);
	T& operator[](uint i) const
	{
		return *reinterpret_cast<T*>(reinterpret_cast<char*>(p) + static_cast<size_t>(i) * stride);
	}
	uint length() const { return n; }
};

// GLSL atomicAdd on a buffer uint: returns the old value
inline uint atomicAdd(uint& mem, uint v) { return __atomic_fetch_add(&mem, v, __ATOMIC_RELAXED); }

// compute built-ins; one "invocation" per call of the shader's main()
struct Builtins
{
	uvec3 gl_WorkGroupID{0, 0, 0};
	uvec3 gl_NumWorkGroups{1, 1, 1};
	uvec3 gl_GlobalInvocationID{0, 0, 0};
};
} // namespace hvref_glsl
