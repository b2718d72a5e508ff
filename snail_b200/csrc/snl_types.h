// snail_b200 — record layouts shared by the host flatten step and the sm_100a kernels.
//
// Reference layouts (what the host SAH builder produces, and what snl_scene_set_geometry accepts):
//   RefNode     = BVH::Node     src/bvh/tree.h:60-72      32 B
//   RefTriangle = Triangle      src/triangle.h:133-135    64 B
//   RefShTri    = ShTriangle    src/triangle.h:181-230    64 B
// Device layouts (what the kernels read, see DESIGN.md "Data layout in HBM"):
//   GpuNode  64 B  one inner node = the boxes of BOTH children + two child links (one fetch per inner node)
//   GenTri   64 B  triangle for rays with their own origin (reflection rays)
//   PreTri   48 B  triangle pre-transformed for ONE ray origin (camera or light): Triangle::Collide's
//                  sharedOrigin precomputation (src/triangle.cpp:13-18) hoisted from per-packet to per-frame
//   RefShTri 64 B  shading record, unchanged (read as 4 x 16 B)
#pragma once

#include <stdint.h>

#if defined(__CUDACC__)
#define SNL_ALIGN16 __align__(16)
#else
#define SNL_ALIGN16 alignas(16)
#endif

namespace snl {

struct RefNode {
	float bmin[3], bmax[3];
	int32_t sub;  // subNode, or first | 0x80000000 for a leaf
	union {
		struct {
			int16_t axis, firstNode;
		};
		int32_t count;
	};
};
static_assert(sizeof(RefNode) == 32, "BVH::Node is 32 bytes");

struct RefTriangle {
	float a[3], ba[3], ca[3];
	float t0, it0;
	int32_t pad;
	float plane[4];
};
static_assert(sizeof(RefTriangle) == 64, "Triangle is 64 bytes");

struct RefShTri {
	float uv[3][2];
	float nrm[3][3];
	int32_t mat;  // matId | flatNormals << 31
};
static_assert(sizeof(RefShTri) == 64, "ShTriangle is 64 bytes");

struct Float4 {
	float x, y, z, w;
};
struct Int4 {
	int32_t x, y, z, w;
};

// Child link: >= 0 inner node index; < 0 leaf: ~link = first | (count-1) << 26, 1 <= count <= 32.
// The flatten step splits larger leaves into a chain of inner nodes carrying the same box.
enum { kLeafFirstBits = 26, kLeafMaxCount = 32, kLeafFirstMask = (1 << kLeafFirstBits) - 1 };

struct SNL_ALIGN16 GpuNode {
	Float4 b0;   // child0 {min.x, max.x, min.y, max.y}
	Float4 b1;   // child1 {min.x, max.x, min.y, max.y}
	Float4 bz;   // {child0 min.z, child0 max.z, child1 min.z, child1 max.z}
	Int4 link;   // {child0 link, child1 link, 0, 0}
};
static_assert(sizeof(GpuNode) == 64, "GpuNode is 64 bytes");

struct SNL_ALIGN16 GenTri {
	Float4 a;   // a.xyz, it0
	Float4 ba;  // ba.xyz, 0
	Float4 ca;  // ca.xyz, 0
	Float4 n;   // plane.xyz (unit normal), 0
};
static_assert(sizeof(GenTri) == 64, "GenTri is 64 bytes");

struct SNL_ALIGN16 PreTri {
	Float4 u;  // tvec1 = (tvec ^ ca) * it0, tmul = -(tvec | nrm)       (tvec = origin - a)
	Float4 v;  // tvec0 = (ba ^ tvec) * it0, 0
	Float4 n;  // nrm, 0
};
static_assert(sizeof(PreTri) == 48, "PreTri is 48 bytes");

// ObjectInstance on the device: rotation rows + translation (src/dbvh/tree.h:177-178), 64 B.
struct SNL_ALIGN16 DevInstance {
	Float4 r0, r1, r2;  // rotation[0..2].xyz
	Float4 t;           // translation.xyz
};

struct DevMaterial {
	int32_t kind, flags;
	float color[3];
	float specular[3];
	float opacity;
	int32_t texture;
	int32_t texture2;  // TransparentMaterial: the opacity (map_d) texture
};

enum { kMaxMips = 16 };  // PointSampler::mipPitch[16], src/sampling/point_sampler.h:28
struct DevTexture {
	const unsigned char *mip[kMaxMips];
	int32_t pitch[kMaxMips];
	int32_t width, height, mips;
	float wMul, hMul;  // width - 1, height - 1 (PointSampler::Update, src/sampling/point_sampler.cpp:27-41)
};

struct DevLight {
	float pos[3], color[3];
	float radius, radSq, iRadius;
};

enum { kMaxLights = 16 };

// RayGenerator state (src/ray_generator.cpp:4-14) evaluated on the host with the reference's operation order.
struct RayGenParams {
	float origin[3];
	float tright[3], tup[3];
	float txyz[3][4];  // per quad lane
};

struct Counters {
	unsigned long long primary, shadow, reflection, reflection_shadow, node_visits, tri_tests, gtri_tests, transparency;
	unsigned long long transp_truncated;  // rays whose transparency chain was cut at kMaxTransp surfaces
};
enum { kCounters = 9 };
static_assert(sizeof(Counters) == kCounters * 8, "Counters is kCounters u64");

}  // namespace snl
