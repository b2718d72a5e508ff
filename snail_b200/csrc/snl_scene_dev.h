// snail_b200 — the by-value kernel argument describing a scene on the device (shared by host and device code).
#pragma once

#include "snl_types.h"

namespace snl {

enum { kStack = 96, kPacket = 256 };

// Per-origin data (camera or one light), rebuilt by k_pretransform when the origin moves: triangle records with
// Triangle::Collide's sharedOrigin precomputation applied.
struct PreSet {
	const PreTri *tris;
};

struct SceneDev {
	const GpuNode *nodes;
	const GenTri *gtris;
	const RefShTri *shtris;  // null: no shading data (BVH::noShadingData)
	const int *triMat;       // BVH::materials[i].id
	const DevMaterial *mats;
	const DevTexture *texs;
	const GpuNode *topNodes;      // instanced scene (DBVH): top-level tree over the instances, leaves = instance ranges
	const DevInstance *instances; // in DBVH::elements order; null = plain BVH
	PreSet lightPre[kMaxLights];
	DevLight lights[kMaxLights];
	int nTriMat, nMats, nLights;
	float ambient[3];
	int gvDepth, gvFull, gvRefl, gvTint;  // gVals[1], [6], [7], [8]
	int gvSuper;                          // gVals[9]: 2x2 supersampling in Render()
	int hasTransp;                        // a material can start a transparency pass (fTransparency and opacity < 1)
	int reflDepth2;                       // two reflection bounces (snl_set_reflection_depth(2)) - plain scenes only
	int hasUber;                          // any UberMaterial in the table (its Shade depends on the packet mask)
	unsigned cullMask;                    // bit i: light i does not reach the whole scene box, so the per-packet light
										  // cull of Scene::RayTrace (src/scene_trace.cpp:494-499) can skip it somewhere
	float tint[3];
	Counters *counters;
};

// Multi-GPU frame (snl_comm.cu): control block at the start of the segment rank 0 shares with its peers (CUDA IPC or
// peer access), all fields updated with system-scope atomics / release stores. Two frame slots alternate.
enum { kCommBands = 8, kCommRanks = 16, kCommCtlBytes = 4096 };
struct CommCtl {
	unsigned long long stats[2][kCounters];            // Counters of the frame summed over the ranks
	unsigned bandDone[2][kCommBands];           // ranks whose share of the row band is stored
	unsigned ready[2];                          // seq + 1 once the slot has been opened for frame seq
	unsigned statsArrived[2];
	unsigned rankDone[kCommRanks];              // frames completed by each rank (seq + 1)
	unsigned error;                             // a wait timed out somewhere
};
static_assert(sizeof(CommCtl) <= kCommCtlBytes, "CommCtl must fit its page");

// Wavefront pipeline state for one frame (snl_wavefront.cu). NL = packets * 256 lanes; arrays are lane-indexed
// (lane = packet * 256 + thread) unless noted, so the shading kernels keep the reference's packet / quad structure
// while the traversal kernels pull rays from compacted queues.
struct WfBuffers {
	Float4 *rayO;    // reflection ray origin.xyz, w = 1 when the primary ray hit (sel0)
	Float4 *rayD;    // reflection ray direction.xyz
	Float4 *hit1;    // reflection hit {dist, obj (int bits), bu, bv}
	Float4 *st0;     // [4 * NL] level-0 state planes: diffuse.rgb | specular.rgb | lightDiffuse.rgb | lightSpecular.rgb
	Float4 *st1;     // [2 * NL] level-1 state: {diffuse.rgb, sel1} {specular.rgb, 0}
	Float4 *sdir;    // [nLights * NL] level-1 shadow ray {dir.xyz, tmax}
	float2 *sdot;    // [nLights * NL] {N.L, distance} of that ray (N.L <= 0: not cast)
	unsigned *rq;    // compacted lane ids of the reflection rays
	unsigned *sq;    // compacted (light * NL + lane) ids of the level-1 shadow rays
	unsigned *occl;  // bit `light` set: that level-1 shadow ray is occluded
	unsigned *qcount;  // [0] #rq  [1] #sq  [2] rq fetch cursor  [3] sq fetch cursor
	const PreSet *lightSets;  // device copy of SceneDev::lightPre (dynamically indexed by the shadow kernel)
	unsigned NL;
};

}  // namespace snl
