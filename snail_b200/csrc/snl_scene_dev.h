// snail_b200 — the by-value kernel argument describing a scene on the device (shared by host and device code).
#pragma once

#include "snl_types.h"

namespace snl {

enum { kStack = 96, kPacket = 256 };

// Per-origin data (camera or one light), rebuilt by k_pretransform when the origin moves: triangle records with
// Triangle::Collide's sharedOrigin precomputation applied, and node boxes translated by the origin.
struct PreSet {
	const PreTri *tris;
	const GpuNode *nodes;
};

struct SceneDev {
	const GpuNode *nodes;
	const GenTri *gtris;
	const RefShTri *shtris;  // null: no shading data (BVH::noShadingData)
	const int *triMat;       // BVH::materials[i].id
	const DevMaterial *mats;
	const DevTexture *texs;
	PreSet lightPre[kMaxLights];
	DevLight lights[kMaxLights];
	int nTriMat, nMats, nLights;
	float ambient[3];
	int gvDepth, gvFull, gvRefl, gvTint;  // gVals[1], [6], [7], [8]
	int hasUber;                          // any UberMaterial in the table (its Shade depends on the packet mask)
	float tint[3];
	Counters *counters;
};

}  // namespace snl
