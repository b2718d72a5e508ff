// snail_b200 — the scene object behind the C ABI handle (host side; owns one device + one stream).
#pragma once

#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "../../include/snail_b200.h"
#include "snl_scene_dev.h"
#include "snl_host_scene.h"
#include "snl_kernels.h"

namespace snl {

void set_error(const char *fmt, ...);

// cudaMalloc'd array with grow-only capacity.
template <class T> struct DevBuf {
	T *ptr = nullptr;
	size_t cap = 0;
	DevBuf() = default;
	DevBuf(const DevBuf &) = delete;
	DevBuf &operator=(const DevBuf &) = delete;
	DevBuf(DevBuf &&o) noexcept : ptr(o.ptr), cap(o.cap) { o.ptr = nullptr, o.cap = 0; }
	~DevBuf() { release(); }  // the owner makes its device current first (snl_scene_destroy)
	cudaError_t reserve(size_t n) {
		if(n <= cap)
			return cudaSuccess;
		if(ptr)
			cudaFree(ptr);
		ptr = nullptr;
		cap = 0;
		cudaError_t e = cudaMalloc((void **)&ptr, n * sizeof(T));
		if(e == cudaSuccess)
			cap = n;
		return e;
	}
	void release() {
		if(ptr)
			cudaFree(ptr);
		ptr = nullptr;
		cap = 0;
	}
};

struct PreCacheEntry {
	float origin[3];
	bool valid = false;
	DevBuf<PreTri> buf;
	PreSet set() const { return PreSet{buf.ptr}; }
};

}  // namespace snl

struct snl_host_scene {
	snl::HostScene scene;
};

struct snl_scene {
	int device = 0;
	cudaStream_t stream = nullptr;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	// snl_render to a host image: the frame is rendered as kBands row bands on their own streams, each followed by the
	// device->host copy of its rows, so all but the last band's copy overlaps the rendering of the bands after it
	enum { kBands = 4, kMaxBands = 8 };  // default / most bands (SNL_RENDER_BANDS=n)
	cudaStream_t bandStream[kMaxBands] = {};
	cudaEvent_t bandEv[kMaxBands] = {}, forkEv = nullptr;
	float lastCallMs = 0.0f;             // device time of the last call that asked for stats (picks the band count)
	snl::Counters *hCounters = nullptr;  // pinned: the counters travel back in-stream, one synchronisation per call
	bool evPending = false;

	// k_render comes in two register budgets (snl_kernels.cu, MINB); the first frame of a workload is rendered with
	// both and the faster one is kept (snl_capi.cu render_common). SNL_RENDER_VARIANT=0|1 pins it.
	struct RenderTune {
		long long key = -1;  // W, H, packets, options, lights, triangles of the workload the choice was made for
		float ms[2] = {0.0f, 0.0f};
		int choice = -1;     // -1: not decided yet
	} tune;

	// reference-layout host copies (kept for snl_scene_copy_host_bvh and re-flattening)
	snl::HostBVH host;
	int nTris = 0, nGpuNodes = 0, gpuDepth = 0;
	bool devBuilt = false;     // tree built by snl_build.cu (host.nodes is empty)
	float devBBox[6] = {0};    // its scene box
	float buildMs = 0.0f;      // device time of the last device build
	int deviceBuilder = 0;     // SNL_BUILDER_LBVH / SNL_BUILDER_SAH: which device builder a soup goes through

	// instanced form (DBVH over ObjectInstances of this scene's BVH, snl_scene_set_instances); empty = plain BVH
	std::vector<snl::HostInstance> instances;  // in DBVH::elements order
	std::vector<snl::RefNode> topNodes;        // DBVH::nodes
	int topDepth = 0;
	snl::DevBuf<snl::GpuNode> dTopNodes;
	snl::DevBuf<snl::DevInstance> dInstances;

	snl::DevBuf<snl::GpuNode> dNodes;
	snl::DevBuf<snl::GenTri> dTris;
	snl::DevBuf<snl::RefShTri> dShTris;
	bool hasShading = false;
	snl::DevBuf<int> dTriMat;
	int nTriMat = 0;
	snl::DevBuf<snl::DevMaterial> dMats;
	int nMats = 0, hasUber = 0, hasTransp = 0;
	snl::DevBuf<snl::DevTexture> dTexs;
	std::vector<snl::DevBuf<unsigned char>> dTexData;

	std::vector<snl_light> lights;
	float ambient[3] = {0.1f, 0.1f, 0.1f};
	snl::PreCacheEntry lightPre[snl::kMaxLights];
	snl::PreCacheEntry originPre;  // camera / shared ray origin of the last call

	int gVals[16] = {0};
	int reflDepth = 1;  // bounces followed when gVals[7] is on: 1 = the reference's cap, 2 = snl_set_reflection_depth(2)

	snl::DevBuf<snl::Counters> dCounters;
	snl::DevBuf<unsigned char> dImage;  // staging image for snl_render (host output)
	snl::DevBuf<unsigned char> dScratch[4];
	std::vector<int> lastTiles;
	snl::DevBuf<int4> dPackets;
	int nPackets = 0;
	snl::DevBuf<int4> dFarmTiles;  // snl_render_farm: tile rectangles, their offsets, the packed tiles
	snl::DevBuf<int> dFarmOffsets;
	snl::DevBuf<unsigned char> dFarm;

	uint64_t launches = 0;

	std::vector<void *> ownedFrames, peerFrames;  // snl_frame_create / snl_frame_open

	// wavefront pipeline (snl_wavefront.cu): per-lane state + ray queues, grow-only
	snl::DevBuf<snl::Float4> wfRayO, wfRayD, wfHit1, wfSt0, wfSt1, wfSdir;
	snl::DevBuf<float2> wfSdot;
	snl::DevBuf<unsigned> wfRq, wfSq, wfOccl, wfCount;
	snl::DevBuf<snl::PreSet> wfLightSets;
	int tuneW = 0, tuneH = 0;  // frame size snl_render_multi last ran the variant pick for
	int smCount = 0;
	int fused = 1;  // 0 (SNL_WAVEFRONT=1): wavefront pipeline for frames with the reflection bounce
};

namespace snl {
// host-side helpers shared by snl_capi.cu and snl_comm.cu
int ensure_band_streams(snl_scene *s);
// the binned-SAH builder of snl_build_sah.cu run as plain loops on the host; returns the number of GpuNode records
int host_build_binned_sah(const GenTri *inTris, int n, std::vector<GpuNode> &nodes, std::vector<unsigned> &orderOut, int *heightOut);
int prepare_frame(snl_scene *s, const snl_camera *cam, int W, int H, int rank, SceneDev &S, RayGenParams &g);
int render_local(snl_scene *s, const snl_camera *cam, int W, int H, const int *tiles, int nTiles, unsigned char *dOut,
				  size_t pitch, int rank);
}  // namespace snl
