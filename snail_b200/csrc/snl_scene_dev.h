// snail_b200 — the by-value kernel argument describing a scene on the device (shared by host and device code).
#pragma once

#include "snl_types.h"

namespace snl {

enum { kStack = 96, kPacket = 256 };

struct SceneDev {
	const GpuNode *nodes;
	const GenTri *gtris;
	const RefShTri *shtris;  // null: no shading data (BVH::noShadingData)
	const int *triMat;       // BVH::materials[i].id
	const DevMaterial *mats;
	const DevTexture *texs;
	const PreTri *lightPre[kMaxLights];
	DevLight lights[kMaxLights];
	int nTriMat, nMats, nLights;
	float ambient[3];
	int gvDepth, gvFull, gvRefl, gvTint;  // gVals[1], [6], [7], [8]
	float tint[3];
	Counters *counters;
};

}  // namespace snl
