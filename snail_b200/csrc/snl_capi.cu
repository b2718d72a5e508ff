// snail_b200 — extern "C" entry points of libsnail_b200.so (see include/snail_b200.h for the contract and the
// reference interface each entry point replaces). No CPU fallback: every compute entry point launches the sm_100a
// kernels of snl_kernels.cu on the scene's device, or fails.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>

#include "snl_scene.h"

using namespace snl;

namespace snl {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_err, sizeof(g_err), fmt, ap);
	va_end(ap);
}

}  // namespace snl

#define SNL_CUDA(call)                                                                              \
	do {                                                                                            \
		cudaError_t e_ = (call);                                                                    \
		if(e_ != cudaSuccess) {                                                                     \
			set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
			return SNL_ERR_CUDA;                                                                    \
		}                                                                                           \
	} while(0)

#define SNL_REQUIRE(cond, code, ...) \
	do {                             \
		if(!(cond)) {                \
			set_error(__VA_ARGS__);  \
			return code;             \
		}                            \
	} while(0)

namespace {

enum { kTopStackHost = 66 };  // = kTopStack of snl_device.cuh (DBVH::maxDepth + 2)

// ---- flatten: BVH::Node[] (own box per node) -> GpuNode[] (both child boxes per inner node) -------------------

struct Flattener {
	const std::vector<RefNode> &ref;
	std::vector<GpuNode> out;
	std::vector<unsigned char> seen;  // a well-formed tree reaches every node once: a second visit is a cycle / shared child
	int maxDepth = 0;
	bool bad = false;

	explicit Flattener(const std::vector<RefNode> &r) : ref(r), seen(r.size(), 0) {}

	static int leaf_link(int first, int count) { return ~(first | ((count - 1) << kLeafFirstBits)); }

	void set_child(int g, int c, const RefNode &box, int link) {
		GpuNode &n = out[g];
		Float4 &b = c == 0 ? n.b0 : n.b1;
		b.x = box.bmin[0];
		b.y = box.bmax[0];
		b.z = box.bmin[1];
		b.w = box.bmax[1];
		if(c == 0) {
			n.bz.x = box.bmin[2];
			n.bz.y = box.bmax[2];
			n.link.x = link;
		}
		else {
			n.bz.z = box.bmin[2];
			n.bz.w = box.bmax[2];
			n.link.y = link;
		}
	}

	// link for reference node r; leaves holding more than kLeafMaxCount triangles become a chain of inner nodes
	// that repeat the leaf's box (same set of triangles tested, same result).
	int link_for(int r, int depth) {
		if(bad || r < 0 || r >= (int)ref.size() || depth > 1024 || seen[r]) {
			bad = true;
			return leaf_link(0, 1);
		}
		seen[r] = 1;
		const RefNode &n = ref[r];
		if(n.sub & 0x80000000) {
			int first = n.sub & 0x7fffffff, count = n.count;
			if(count < 1 || first + count > (1 << kLeafFirstBits)) {
				bad = true;
				return leaf_link(0, 1);
			}
			return chain(n, first, count, depth);
		}
		const int g = (int)out.size();
		out.push_back(GpuNode());
		memset(&out[g], 0, sizeof(GpuNode));
		maxDepth = std::max(maxDepth, depth + 1);
		const int sub = n.sub;
		if(sub < 0 || sub + 1 >= (int)ref.size()) {
			bad = true;
			return leaf_link(0, 1);
		}
		const int l0 = link_for(sub, depth + 1), l1 = link_for(sub + 1, depth + 1);
		set_child(g, 0, ref[sub], l0);
		set_child(g, 1, ref[sub + 1], l1);
		return g;
	}

	int chain(const RefNode &box, int first, int count, int depth) {
		if(count <= kLeafMaxCount)
			return leaf_link(first, count);
		const int g = (int)out.size();
		out.push_back(GpuNode());
		memset(&out[g], 0, sizeof(GpuNode));
		maxDepth = std::max(maxDepth, depth + 1);
		const int rest = chain(box, first + kLeafMaxCount, count - kLeafMaxCount, depth + 1);
		set_child(g, 0, box, leaf_link(first, kLeafMaxCount));
		set_child(g, 1, box, rest);
		return g;
	}

	bool run() {
		out.reserve(ref.size());
		const int root = link_for(0, 0);
		if(root < 0) {
			// the root itself is a leaf: node 0 = two links to it (the second visit cannot change the result:
			// nearest-hit updates need a strictly smaller t, any-hit is idempotent)
			out.insert(out.begin(), GpuNode());
			memset(&out[0], 0, sizeof(GpuNode));
			for(size_t i = 1; i < out.size(); i++) {  // chain nodes moved by one
				if(out[i].link.x >= 0)
					out[i].link.x++;
				if(out[i].link.y >= 0)
					out[i].link.y++;
			}
			set_child(0, 0, ref[0], root);
			set_child(0, 1, ref[0], root);
			maxDepth = std::max(maxDepth, 1);
		}
		return !bad && root <= 0;
	}
};

int sync_and_stats(snl_scene *s, snl_stats *stats) {
	if(stats) {
		if(!s->hCounters)
			SNL_CUDA(cudaMallocHost(&s->hCounters, sizeof(Counters)));
		SNL_CUDA(cudaMemcpyAsync(s->hCounters, s->dCounters.ptr, sizeof(Counters), cudaMemcpyDeviceToHost, s->stream));
	}
	SNL_CUDA(cudaStreamSynchronize(s->stream));
	if(stats) {
		const Counters c = *s->hCounters;
		memset(stats, 0, sizeof(*stats));
		stats->primary = c.primary;
		stats->shadow = c.shadow;
		stats->reflection = c.reflection;
		stats->reflection_shadow = c.reflection_shadow;
		stats->transparency = c.transparency;
		stats->transparency_truncated = c.transp_truncated;
		stats->rays = c.primary + c.shadow + c.reflection + c.reflection_shadow + c.transparency;
		stats->node_visits = c.node_visits;
		stats->tri_tests = c.tri_tests;
		stats->gtri_tests = c.gtri_tests;
		if(s->evPending) {
			float ms = 0.0f;
			if(cudaEventElapsedTime(&ms, s->ev0, s->ev1) == cudaSuccess) {
				stats->kernel_ms = ms;
				s->lastCallMs = ms;
			}
		}
	}
	s->evPending = false;
	return SNL_OK;
}

}  // namespace
namespace snl {
int ensure_pre(snl_scene *s, PreCacheEntry &e, const float origin[3]);
int make_scene_dev(snl_scene *s, SceneDev &S, const float *shareOrigin, unsigned *shared);
void make_raygen(const snl_camera &c, int W, int H, RayGenParams &g);
}
namespace {

int begin_call(snl_scene *s) {
	SNL_REQUIRE(s, SNL_ERR_INVALID, "null scene");
	SNL_CUDA(cudaSetDevice(s->device));
	SNL_REQUIRE(s->nGpuNodes > 0, SNL_ERR_STATE, "no geometry: call snl_scene_set_geometry or snl_scene_load_obj first");
	SNL_CUDA(cudaMemsetAsync(s->dCounters.ptr, 0, sizeof(Counters), s->stream));
	SNL_CUDA(cudaEventRecord(s->ev0, s->stream));
	return SNL_OK;
}

int end_call(snl_scene *s) {
	SNL_CUDA(cudaEventRecord(s->ev1, s->stream));
	s->evPending = true;
	return SNL_OK;
}

}  // namespace

// Per-origin triangle records, recomputed only when the origin changes.
int snl::ensure_pre(snl_scene *s, PreCacheEntry &e, const float origin[3]) {
	if(e.valid && memcmp(e.origin, origin, 12) == 0)
		return SNL_OK;
	SNL_CUDA(e.buf.reserve((size_t)s->nTris));
	SNL_CUDA(launch_pretransform(s->dTris.ptr, s->nTris, origin, e.buf.ptr, s->stream));
	s->launches++;
	memcpy(e.origin, origin, 12);
	e.valid = true;
	return SNL_OK;
}

// shareOrigin (the camera position of a Render() call): a light sitting exactly there - the reference's own way of
// placing a light, src/rtracer.cpp:270-281 - shares the camera's per-origin triangle records instead of getting a
// second identical copy (half the PreTri bytes competing for L1 / L2, one k_pretransform less per frame). The caller
// fills S.lightPre[i] for the lights flagged in *shared once the camera's records are up to date.
int snl::make_scene_dev(snl_scene *s, SceneDev &S, const float *shareOrigin, unsigned *shared) {
	memset(&S, 0, sizeof(S));
	S.nodes = s->dNodes.ptr;
	S.gtris = s->dTris.ptr;
	S.shtris = s->hasShading ? s->dShTris.ptr : nullptr;
	S.triMat = s->dTriMat.ptr;
	S.nTriMat = s->nTriMat;
	S.mats = s->dMats.ptr;
	S.nMats = s->nMats;
	S.texs = s->dTexs.ptr;
	S.topNodes = s->instances.empty() ? nullptr : s->dTopNodes.ptr;
	S.instances = s->instances.empty() ? nullptr : s->dInstances.ptr;
	S.nLights = (int)s->lights.size();
	for(int i = 0; i < S.nLights; i++) {
		const snl_light &l = s->lights[i];
		const char *shareEnv = getenv("SNL_SHARE_PRE");  // "0": every light gets its own records (A/B)
		if(shareOrigin && shared && memcmp(l.pos, shareOrigin, 12) == 0 && !(shareEnv && shareEnv[0] == '0'))
			*shared |= 1u << i;
		else {
			int r = ensure_pre(s, s->lightPre[i], l.pos);
			if(r)
				return r;
			S.lightPre[i] = s->lightPre[i].set();
		}
		{
			// can BoxPointDistanceSq(box of a packet's hit points, light) ever exceed radSq? Hit points lie in the scene box
			// (instanced: the top-level box), so a light whose farthest box corner is within its radius is never culled and
			// the kernels skip the packet-wide reduction for it. 1 % margin for the rounding of dir * t + origin.
			const float *bmin, *bmax;
			if(!s->instances.empty() && !s->topNodes.empty()) {
				bmin = s->topNodes[0].bmin;
				bmax = s->topNodes[0].bmax;
			}
			else if(s->devBuilt) {
				bmin = s->devBBox;
				bmax = s->devBBox + 3;
			}
			else {
				bmin = s->host.nodes[0].bmin;
				bmax = s->host.nodes[0].bmax;
			}
			float far2 = 0.0f;
			for(int a = 0; a < 3; a++) {
				const float ext = (bmax[a] - bmin[a]) * 0.01f;
				const float d = std::max(std::fabs(l.pos[a] - (bmin[a] - ext)), std::fabs(l.pos[a] - (bmax[a] + ext)));
				far2 += d * d;
			}
			const char *cullEnv = getenv("SNL_LIGHT_CULL");  // "0": never (ray counts then exceed the reference's; A/B)
			if(!(far2 <= l.radSq) && !(cullEnv && cullEnv[0] == '0'))
				S.cullMask |= 1u << i;
		}
		memcpy(S.lights[i].pos, l.pos, 12);
		memcpy(S.lights[i].color, l.color, 12);
		S.lights[i].radius = l.radius;
		S.lights[i].radSq = l.radSq;
		S.lights[i].iRadius = l.iRadius;
	}
	memcpy(S.ambient, s->ambient, 12);
	S.gvDepth = s->gVals[1];
	S.gvFull = s->gVals[6];
	S.gvRefl = s->gVals[7];
	S.gvTint = 0;
	S.gvSuper = s->gVals[9];
	S.reflDepth2 = s->reflDepth == 2 && s->gVals[7] && !s->gVals[1];
	S.hasUber = s->hasUber;
	S.hasTransp = s->hasTransp;
	S.counters = s->dCounters.ptr;
	return SNL_OK;
}

// RayGenerator::RayGenerator (src/ray_generator.cpp:4-14), same operation order.
void snl::make_raygen(const snl_camera &c, int W, int H, RayGenParams &g) {
	float invW = 1.0f / float(W), invH = 1.0f / float(H);
	invW *= float(W) / float(H);
	const float addx[4] = {0.0f, 1.0f, 0.0f, 1.0f}, addy[4] = {0.0f, 0.0f, 1.0f, 1.0f};
	for(int a = 0; a < 3; a++) {
		g.origin[a] = c.pos[a];
		g.tright[a] = c.right[a] * invW;
		g.tup[a] = c.up[a] * invH;
	}
	for(int k = 0; k < 4; k++) {
		const float taddx = addx[k] - float(W * 0.5f), taddy = addy[k] - float(H * 0.5f);
		for(int a = 0; a < 3; a++) {
			const float fp = c.front[a] * c.plane_dist;
			g.txyz[a][k] = (g.tright[a] * taddx + g.tup[a] * taddy) + fp;
		}
	}
}

static const float kRankTint[16][3] = {  // src/render.cpp:119-128
	{0.6f, 0.6f, 0.6f}, {0.6f, 0.6f, 1.0f}, {0.6f, 1.0f, 0.6f}, {0.6f, 1.0f, 1.0f}, {1.0f, 0.6f, 0.6f}, {1.0f, 0.6f, 1.0f},
	{1.0f, 1.0f, 0.6f}, {1.0f, 1.0f, 1.0f}, {0.3f, 0.3f, 0.3f}, {0.3f, 0.3f, 1.0f}, {0.3f, 0.7f, 0.3f}, {0.3f, 0.7f, 0.7f},
	{0.7f, 0.3f, 0.3f}, {0.7f, 0.3f, 0.7f}, {0.7f, 0.7f, 0.3f}, {0.7f, 1.7f, 0.7f}};

// Row-band streams + events of a scene (snl_render's banded host copy, snl_render_multi), created on first use.
int snl::ensure_band_streams(snl_scene *s) {
	if(s->forkEv)
		return SNL_OK;
	SNL_CUDA(cudaSetDevice(s->device));
	SNL_CUDA(cudaEventCreateWithFlags(&s->forkEv, cudaEventDisableTiming));
	for(int b = 0; b < snl_scene::kMaxBands; b++) {
		SNL_CUDA(cudaStreamCreateWithFlags(&s->bandStream[b], cudaStreamNonBlocking));
		SNL_CUDA(cudaEventCreateWithFlags(&s->bandEv[b], cudaEventDisableTiming));
	}
	return SNL_OK;
}

// What one Render() call needs on the device before its k_render: the scene argument (lights, options, per-origin
// triangle records of the camera and the lights, shared where a light sits at the camera) and the ray generator.
int snl::prepare_frame(snl_scene *s, const snl_camera *cam, int W, int H, int rank, SceneDev &S, RayGenParams &g) {
	unsigned sharedLights = 0;
	if(int r = make_scene_dev(s, S, cam->pos, &sharedLights))
		return r;
	SNL_REQUIRE(!(S.instances && S.gvSuper), SNL_ERR_UNSUPPORTED, "gVals[9] supersampling is not available on an instanced scene");
	SNL_REQUIRE(!(S.reflDepth2 && (S.instances || S.hasTransp || S.gvSuper)), SNL_ERR_UNSUPPORTED,
				"reflection depth 2 is only available on plain scenes (no instances, no transparent material, no gVals[9])");
	if(s->gVals[8]) {
		S.gvTint = 1;
		memcpy(S.tint, kRankTint[(unsigned)rank % 16], 12);
	}
	if(int r = ensure_pre(s, s->originPre, cam->pos))
		return r;
	for(int i = 0; i < S.nLights; i++)
		if(sharedLights >> i & 1)
			S.lightPre[i] = s->originPre.set();
	const int scale = s->gVals[9] ? 2 : 1;  // src/render.cpp:60-62
	make_raygen(*cam, W * scale, H * scale, g);
	return SNL_OK;
}

// =================================================================================================================
extern "C" {

const char *snl_last_error(void) { return g_err; }
int snl_abi_version(void) { return SNL_ABI_VERSION; }

int snl_device_count(void) {
	int n = 0;
	if(cudaGetDeviceCount(&n) != cudaSuccess) {
		cudaGetLastError();
		return 0;
	}
	return n;
}

snl_scene *snl_scene_create(int device) {
	int n = snl_device_count();
	if(n <= 0) {
		set_error("no CUDA device visible: libsnail_b200 has no CPU path");
		return nullptr;
	}
	if(device < 0 || device >= n) {
		set_error("device %d out of range (have %d)", device, n);
		return nullptr;
	}
	if(cudaSetDevice(device) != cudaSuccess) {
		set_error("cudaSetDevice(%d) failed", device);
		return nullptr;
	}
	snl_scene *s = new snl_scene;
	s->device = device;
	cudaDeviceGetAttribute(&s->smCount, cudaDevAttrMultiProcessorCount, device);
	if(s->smCount <= 0)
		s->smCount = 148;
	// SNL_WAVEFRONT=1 selects the queue-driven pipeline of snl_wavefront.cu for frames with the bounce; measured
	// slower than the fused kernel on every scene tried, so the fused k_render is the default
	const char *wfEnv = getenv("SNL_WAVEFRONT");
	s->fused = !(wfEnv && wfEnv[0] == '1');
	bool ok = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) == cudaSuccess &&
			  cudaEventCreate(&s->ev0) == cudaSuccess && cudaEventCreate(&s->ev1) == cudaSuccess &&
			  s->dCounters.reserve(1) == cudaSuccess;
	if(!ok) {
		set_error("CUDA resource creation failed: %s", cudaGetErrorString(cudaGetLastError()));
		snl_scene_destroy(s);
		return nullptr;
	}
	return s;
}

void snl_scene_destroy(snl_scene *s) {
	if(!s)
		return;
	cudaSetDevice(s->device);
	if(s->stream)
		cudaStreamSynchronize(s->stream);
	s->dTopNodes.release();
	s->dInstances.release();
	s->dNodes.release();
	s->dTris.release();
	s->dShTris.release();
	s->dTriMat.release();
	s->dMats.release();
	s->dTexs.release();
	for(auto &t : s->dTexData)
		t.release();
	for(auto &e : s->lightPre)
		e.buf.release();
	s->originPre.buf.release();
	s->dFarmTiles.release();
	s->dFarmOffsets.release();
	s->dFarm.release();
	s->dCounters.release();
	s->dImage.release();
	for(auto &b : s->dScratch)
		b.release();
	s->dPackets.release();
	s->wfRayO.release();
	s->wfRayD.release();
	s->wfHit1.release();
	s->wfSt0.release();
	s->wfSt1.release();
	s->wfSdir.release();
	s->wfSdot.release();
	s->wfRq.release();
	s->wfSq.release();
	s->wfOccl.release();
	s->wfCount.release();
	s->wfLightSets.release();
	for(void *p : s->peerFrames)
		cudaIpcCloseMemHandle(p);
	for(void *p : s->ownedFrames)
		cudaFree(p);
	if(s->ev0)
		cudaEventDestroy(s->ev0);
	if(s->ev1)
		cudaEventDestroy(s->ev1);
	if(s->forkEv)
		cudaEventDestroy(s->forkEv);
	if(s->hCounters)
		cudaFreeHost(s->hCounters);
	for(int b = 0; b < snl_scene::kMaxBands; b++) {
		if(s->bandEv[b])
			cudaEventDestroy(s->bandEv[b]);
		if(s->bandStream[b])
			cudaStreamDestroy(s->bandStream[b]);
	}
	if(s->stream)
		cudaStreamDestroy(s->stream);
	delete s;
}

static std::vector<GenTri> gen_tris(const HostBVH &h) {
	const int nTris = (int)h.tris.size();
	std::vector<GenTri> gt(nTris);
	for(int i = 0; i < nTris; i++) {
		const RefTriangle &t = h.tris[i];
		gt[i].a = {t.a[0], t.a[1], t.a[2], t.it0};
		gt[i].ba = {t.ba[0], t.ba[1], t.ba[2], 0.0f};
		gt[i].ca = {t.ca[0], t.ca[1], t.ca[2], 0.0f};
		gt[i].n = {t.plane[0], t.plane[1], t.plane[2], 0.0f};
	}
	return gt;
}

// HostBVH without nodes: the triangles are a soup in file order and the tree is built on the device (snl_build.cu).
static int build_geometry_on_device(snl_scene *s) {
	const HostBVH &h = s->host;
	const int nTris = (int)h.tris.size();
	const std::vector<GenTri> gt = gen_tris(h);
	SNL_CUDA(cudaStreamSynchronize(s->stream));
	DevBuf<GenTri> inTris;
	DevBuf<RefShTri> inSh;
	SNL_CUDA(inTris.reserve(nTris));
	SNL_CUDA(cudaMemcpy(inTris.ptr, gt.data(), (size_t)nTris * sizeof(GenTri), cudaMemcpyHostToDevice));
	s->hasShading = !h.shTris.empty();
	if(s->hasShading) {
		SNL_CUDA(inSh.reserve(nTris));
		SNL_CUDA(cudaMemcpy(inSh.ptr, h.shTris.data(), (size_t)nTris * sizeof(RefShTri), cudaMemcpyHostToDevice));
		SNL_CUDA(s->dShTris.reserve(nTris));
	}
	SNL_CUDA(s->dTris.reserve(nTris));
	SNL_CUDA(s->dNodes.reserve(std::max(nTris - 1, 1)));
	int height = 0, launches = 0, nNodes = std::max(nTris - 1, 1);
	cudaError_t err = s->deviceBuilder == SNL_BUILDER_SAH
						  ? device_build_sah(inTris.ptr, s->hasShading ? inSh.ptr : nullptr, nTris, s->dTris.ptr,
											 s->hasShading ? s->dShTris.ptr : nullptr, s->dNodes.ptr, &nNodes, &height, s->devBBox, s->stream,
											 &launches, &s->buildMs)
						  : device_build_lbvh(inTris.ptr, s->hasShading ? inSh.ptr : nullptr, nTris, s->dTris.ptr,
											  s->hasShading ? s->dShTris.ptr : nullptr, s->dNodes.ptr, &height, s->devBBox, s->stream,
											  &launches, &s->buildMs);
	inTris.release();
	inSh.release();
	if(err != cudaSuccess) {
		set_error("device BVH build failed: %s", cudaGetErrorString(err));
		return SNL_ERR_CUDA;
	}
	s->launches += launches;
	SNL_REQUIRE(height <= kStack - 2, SNL_ERR_UNSUPPORTED, "device-built tree depth %d exceeds the traversal stack (%d)",
				height, kStack - 2);
	s->nTriMat = (int)h.triMatId.size();
	SNL_CUDA(s->dTriMat.reserve(std::max(1, s->nTriMat)));
	if(s->nTriMat)
		SNL_CUDA(cudaMemcpy(s->dTriMat.ptr, h.triMatId.data(), s->nTriMat * sizeof(int), cudaMemcpyHostToDevice));
	s->nTris = nTris;
	s->nGpuNodes = nNodes;
	s->gpuDepth = height;
	s->devBuilt = true;
	for(auto &e : s->lightPre)
		e.valid = false;
	s->originPre.valid = false;
	return SNL_OK;
}

static int upload_geometry(snl_scene *s) {
	const HostBVH &h = s->host;
	// instance boxes and the top-level tree were built from the OLD tree's box: a new tree starts as a plain BVH again
	// (call snl_scene_set_instances after the geometry)
	s->instances.clear();
	s->topNodes.clear();
	s->tune.key = -1;
	if(h.nodes.empty() && !h.tris.empty())
		return build_geometry_on_device(s);
	s->devBuilt = false;
	Flattener f(h.nodes);
	SNL_REQUIRE(f.run(), SNL_ERR_INVALID, "malformed BVH::Node array (child / triangle range out of bounds)");
	SNL_REQUIRE(f.maxDepth <= kStack - 2, SNL_ERR_UNSUPPORTED, "tree depth %d exceeds the traversal stack (%d)",
				f.maxDepth, kStack - 2);
	for(const RefNode &n : h.nodes)
		if(n.sub & 0x80000000) {
			long first = n.sub & 0x7fffffff;
			SNL_REQUIRE(first + n.count <= (long)h.tris.size(), SNL_ERR_INVALID, "leaf triangle range out of bounds");
		}

	const int nTris = (int)h.tris.size();
	const std::vector<GenTri> gt = gen_tris(h);
	SNL_CUDA(cudaStreamSynchronize(s->stream));
	SNL_CUDA(s->dNodes.reserve(f.out.size()));
	SNL_CUDA(cudaMemcpy(s->dNodes.ptr, f.out.data(), f.out.size() * sizeof(GpuNode), cudaMemcpyHostToDevice));
	SNL_CUDA(s->dTris.reserve(nTris));
	SNL_CUDA(cudaMemcpy(s->dTris.ptr, gt.data(), (size_t)nTris * sizeof(GenTri), cudaMemcpyHostToDevice));
	s->hasShading = !h.shTris.empty();
	if(s->hasShading) {
		SNL_CUDA(s->dShTris.reserve(nTris));
		SNL_CUDA(cudaMemcpy(s->dShTris.ptr, h.shTris.data(), (size_t)nTris * sizeof(RefShTri), cudaMemcpyHostToDevice));
	}
	s->nTriMat = (int)h.triMatId.size();
	SNL_CUDA(s->dTriMat.reserve(std::max(1, s->nTriMat)));
	if(s->nTriMat)
		SNL_CUDA(cudaMemcpy(s->dTriMat.ptr, h.triMatId.data(), s->nTriMat * sizeof(int), cudaMemcpyHostToDevice));
	s->nTris = nTris;
	s->nGpuNodes = (int)f.out.size();
	s->gpuDepth = f.maxDepth;
	for(auto &e : s->lightPre)
		e.valid = false;
	s->originPre.valid = false;
	return SNL_OK;
}

int snl_scene_set_geometry(snl_scene *s, const void *nodes, int nNodes, const void *tris, const void *shTris, int nTris,
						   const int *triMatId, int nMatIds) {
	SNL_REQUIRE(s && nodes && tris, SNL_ERR_INVALID, "null argument");
	SNL_REQUIRE(nNodes > 0 && nTris > 0, SNL_ERR_INVALID, "empty geometry (nNodes=%d nTris=%d)", nNodes, nTris);
	SNL_REQUIRE(nTris <= (1 << kLeafFirstBits), SNL_ERR_UNSUPPORTED, "more than 2^26 triangles");
	SNL_REQUIRE(nMatIds >= 0 && (nMatIds == 0 || triMatId), SNL_ERR_INVALID, "bad material id table");
	SNL_CUDA(cudaSetDevice(s->device));
	HostBVH &h = s->host;
	h.nodes.assign((const RefNode *)nodes, (const RefNode *)nodes + nNodes);
	h.tris.assign((const RefTriangle *)tris, (const RefTriangle *)tris + nTris);
	if(shTris)
		h.shTris.assign((const RefShTri *)shTris, (const RefShTri *)shTris + nTris);
	else
		h.shTris.clear();
	h.triMatId.assign(triMatId, triMatId + nMatIds);
	h.matNames.clear();
	h.depth = 0;
	return upload_geometry(s);
}

int snl_scene_build_device(snl_scene *s, const void *tris, const void *shTris, int nTris, const int *triMatId, int nMatIds) {
	SNL_REQUIRE(s && tris, SNL_ERR_INVALID, "null argument");
	SNL_REQUIRE(nTris > 0, SNL_ERR_INVALID, "empty geometry (nTris=%d)", nTris);
	SNL_REQUIRE(nTris <= (1 << kLeafFirstBits), SNL_ERR_UNSUPPORTED, "more than 2^26 triangles");
	SNL_REQUIRE(nMatIds >= 0 && (nMatIds == 0 || triMatId), SNL_ERR_INVALID, "bad material id table");
	SNL_CUDA(cudaSetDevice(s->device));
	HostBVH &h = s->host;
	h.nodes.clear();
	h.tris.assign((const RefTriangle *)tris, (const RefTriangle *)tris + nTris);
	if(shTris)
		h.shTris.assign((const RefShTri *)shTris, (const RefShTri *)shTris + nTris);
	else
		h.shTris.clear();
	h.triMatId.assign(triMatId, triMatId + nMatIds);
	h.matNames.clear();
	h.depth = 0;
	return upload_geometry(s);
}

float snl_scene_build_ms(snl_scene *s) { return s ? s->buildMs : 0.0f; }

int snl_scene_set_device_builder(snl_scene *s, int builder) {
	SNL_REQUIRE(s, SNL_ERR_INVALID, "null scene");
	SNL_REQUIRE(builder == SNL_BUILDER_LBVH || builder == SNL_BUILDER_SAH, SNL_ERR_INVALID, "unknown builder %d", builder);
	s->deviceBuilder = builder;
	return SNL_OK;
}

int snl_host_build_binned_sah(const void *tris, int nTris, void *gpuNodes, int *nNodes, int *order, int *height) {
	SNL_REQUIRE(tris && nTris > 0 && gpuNodes && nNodes && order && height, SNL_ERR_INVALID, "bad argument");
	HostBVH h;
	h.tris.assign((const RefTriangle *)tris, (const RefTriangle *)tris + nTris);
	const std::vector<GenTri> gt = gen_tris(h);
	std::vector<GpuNode> nodes;
	std::vector<unsigned> ord;
	const int r = host_build_binned_sah(gt.data(), nTris, nodes, ord, height);
	SNL_REQUIRE(r > 0, SNL_ERR_UNSUPPORTED, "binned SAH build failed (%d: -2 node capacity, -3 deeper than the traversal stack)", r);
	*nNodes = r;
	memcpy(gpuNodes, nodes.data(), (size_t)r * sizeof(GpuNode));
	for(int i = 0; i < nTris; i++)
		order[i] = (int)ord[i];
	return SNL_OK;
}

int snl_scene_set_instances(snl_scene *s, const snl_instance *inst, int n) {
	SNL_REQUIRE(s && n >= 0 && (n == 0 || inst), SNL_ERR_INVALID, "bad argument");
	SNL_REQUIRE(s->nGpuNodes > 0, SNL_ERR_STATE, "no geometry: instances refer to the scene's BVH");
	SNL_CUDA(cudaSetDevice(s->device));
	SNL_CUDA(cudaStreamSynchronize(s->stream));
	s->tune.key = -1;
	if(n == 0) {
		s->instances.clear();
		s->topNodes.clear();
		return SNL_OK;
	}
	float bmin[3], bmax[3];
	if(int r = snl_scene_bbox(s, bmin, bmax))
		return r;
	std::vector<HostInstance> elems(n);
	for(int i = 0; i < n; i++) {
		memcpy(elems[i].rot, inst[i].rotation, 36);
		memcpy(elems[i].tr, inst[i].translation, 12);
		instance_compute_bbox(elems[i], bmin, bmax);
	}
	std::vector<RefNode> nodes;
	int depth = 0;
	std::string err;
	if(!build_dbvh(elems, nodes, depth, err)) {
		set_error("%s", err.c_str());
		return SNL_ERR_INVALID;
	}
	Flattener f(nodes);
	SNL_REQUIRE(f.run(), SNL_ERR_INVALID, "malformed DBVH");
	SNL_REQUIRE(f.maxDepth <= kTopStackHost - 2, SNL_ERR_UNSUPPORTED, "DBVH depth %d exceeds the top-level stack", f.maxDepth);
	std::vector<DevInstance> di(n);
	for(int i = 0; i < n; i++) {
		const HostInstance &e = elems[i];
		di[i].r0 = {e.rot[0][0], e.rot[0][1], e.rot[0][2], 0.0f};
		di[i].r1 = {e.rot[1][0], e.rot[1][1], e.rot[1][2], 0.0f};
		di[i].r2 = {e.rot[2][0], e.rot[2][1], e.rot[2][2], 0.0f};
		di[i].t = {e.tr[0], e.tr[1], e.tr[2], 0.0f};
	}
	SNL_CUDA(s->dTopNodes.reserve(f.out.size()));
	SNL_CUDA(cudaMemcpy(s->dTopNodes.ptr, f.out.data(), f.out.size() * sizeof(GpuNode), cudaMemcpyHostToDevice));
	SNL_CUDA(s->dInstances.reserve(n));
	SNL_CUDA(cudaMemcpy(s->dInstances.ptr, di.data(), (size_t)n * sizeof(DevInstance), cudaMemcpyHostToDevice));
	s->instances.swap(elems);
	s->topNodes.swap(nodes);
	s->topDepth = depth;
	return SNL_OK;
}

int snl_host_dbvh_build(const snl_instance *inst, int n, const float treeMin[3], const float treeMax[3], void *nodesOut,
						int *nNodes, snl_instance *elementsOut, float *bboxes6Out) {
	SNL_REQUIRE(inst && n > 0 && treeMin && treeMax && nNodes, SNL_ERR_INVALID, "bad argument");
	std::vector<HostInstance> elems(n);
	for(int i = 0; i < n; i++) {
		memcpy(elems[i].rot, inst[i].rotation, 36);
		memcpy(elems[i].tr, inst[i].translation, 12);
		instance_compute_bbox(elems[i], treeMin, treeMax);
	}
	std::vector<RefNode> nodes;
	int depth = 0;
	std::string err;
	if(!build_dbvh(elems, nodes, depth, err)) {
		set_error("%s", err.c_str());
		return SNL_ERR_INVALID;
	}
	*nNodes = (int)nodes.size();
	if(nodesOut)
		memcpy(nodesOut, nodes.data(), nodes.size() * sizeof(RefNode));
	for(int i = 0; i < n; i++) {
		if(elementsOut) {
			memcpy(elementsOut[i].rotation, elems[i].rot, 36);
			memcpy(elementsOut[i].translation, elems[i].tr, 12);
		}
		if(bboxes6Out) {
			memcpy(bboxes6Out + i * 6, elems[i].bmin, 12);
			memcpy(bboxes6Out + i * 6 + 3, elems[i].bmax, 12);
		}
	}
	return SNL_OK;
}

int snl_scene_copy_dbvh(snl_scene *s, int *nNodes, void *nodes, int *nInstances, snl_instance *elements, float *bboxes6) {
	SNL_REQUIRE(s, SNL_ERR_INVALID, "null scene");
	if(nNodes)
		*nNodes = (int)s->topNodes.size();
	if(nInstances)
		*nInstances = (int)s->instances.size();
	if(nodes)
		memcpy(nodes, s->topNodes.data(), s->topNodes.size() * sizeof(RefNode));
	for(size_t i = 0; i < s->instances.size(); i++) {
		const HostInstance &e = s->instances[i];
		if(elements) {
			memcpy(elements[i].rotation, e.rot, 36);
			memcpy(elements[i].translation, e.tr, 12);
		}
		if(bboxes6) {
			memcpy(bboxes6 + i * 6, e.bmin, 12);
			memcpy(bboxes6 + i * 6 + 3, e.bmax, 12);
		}
	}
	return SNL_OK;
}

static int upload_materials(snl_scene *s, const snl_material *mats, int nMats, const snl_texture *tex, int nTex) {
	SNL_CUDA(cudaSetDevice(s->device));
	SNL_CUDA(cudaStreamSynchronize(s->stream));
	for(int i = 0; i < nMats; i++) {
		const snl_material &m = mats[i];
		SNL_REQUIRE(m.kind >= SNL_MAT_SIMPLE && m.kind <= SNL_MAT_TRANSPARENT, SNL_ERR_INVALID, "material %d: bad kind", i);
		if(m.kind == SNL_MAT_TEX || m.kind == SNL_MAT_TRANSPARENT)
			SNL_REQUIRE(m.texture >= 0 && m.texture < nTex, SNL_ERR_INVALID, "material %d: texture index", i);
		if(m.kind == SNL_MAT_TRANSPARENT)
			SNL_REQUIRE(m.mask_texture >= 0 && m.mask_texture < nTex, SNL_ERR_INVALID, "material %d: mask texture index", i);
	}
	for(auto &t : s->dTexData)
		t.release();
	s->dTexData.clear();
	s->dTexData.resize(nTex);
	std::vector<DevTexture> dt(nTex);
	for(int i = 0; i < nTex; i++) {
		const snl_texture &t = tex[i];
		SNL_REQUIRE(t.data && t.width > 0 && t.height > 0 && t.mips > 0 && t.mips <= kMaxMips, SNL_ERR_INVALID,
					"texture %d: bad description", i);
		SNL_REQUIRE(!(t.width & (t.width - 1)) && !(t.height & (t.height - 1)), SNL_ERR_INVALID,
					"texture %d: width & height must be a power of 2", i);
		size_t total = 0;
		for(int m = 0; m < t.mips; m++)
			total += (size_t)std::max(t.width >> m, 1) * std::max(t.height >> m, 1) * 3;
		SNL_REQUIRE(total <= t.size, SNL_ERR_INVALID, "texture %d: chain is %zu bytes, %zu given", i, total, t.size);
		// +4: PointSampler reads 4 bytes per texel (src/sampling/point_sampler.cpp:171); we read 3, pad anyway
		SNL_CUDA(s->dTexData[i].reserve(total + 4));
		SNL_CUDA(cudaMemcpy(s->dTexData[i].ptr, t.data, total, cudaMemcpyHostToDevice));
		DevTexture &d = dt[i];
		memset(&d, 0, sizeof(d));
		size_t off = 0;
		for(int m = 0; m < t.mips; m++) {
			const int w = std::max(t.width >> m, 1), hh = std::max(t.height >> m, 1);
			d.mip[m] = s->dTexData[i].ptr + off;
			d.pitch[m] = w * 3;
			off += (size_t)w * hh * 3;
		}
		d.width = t.width;
		d.height = t.height;
		d.mips = t.mips;
		d.wMul = float(t.width - 1);
		d.hMul = float(t.height - 1);
	}
	SNL_CUDA(s->dTexs.reserve(std::max(1, nTex)));
	if(nTex)
		SNL_CUDA(cudaMemcpy(s->dTexs.ptr, dt.data(), nTex * sizeof(DevTexture), cudaMemcpyHostToDevice));
	std::vector<DevMaterial> dm(nMats);
	for(int i = 0; i < nMats; i++) {
		dm[i].kind = mats[i].kind;
		dm[i].flags = mats[i].flags;
		memcpy(dm[i].color, mats[i].color, 12);
		memcpy(dm[i].specular, mats[i].specular, 12);
		dm[i].opacity = mats[i].opacity;
		dm[i].texture = mats[i].texture;
		dm[i].texture2 = mats[i].mask_texture;
	}
	SNL_CUDA(s->dMats.reserve(std::max(1, nMats)));
	if(nMats)
		SNL_CUDA(cudaMemcpy(s->dMats.ptr, dm.data(), nMats * sizeof(DevMaterial), cudaMemcpyHostToDevice));
	s->nMats = nMats;
	s->hasUber = s->hasTransp = 0;
	for(int i = 0; i < nMats; i++) {
		s->hasUber |= mats[i].kind == SNL_MAT_UBER;
		// transSel needs fTransparency, and survives only where opacity < 1 (src/scene_trace.cpp:472): an Uber
		// material with d >= 1 (every material of sponza.mtl) can never start a transparency pass
		if(mats[i].flags & SNL_MATF_TRANSPARENCY)
			s->hasTransp |= mats[i].kind == SNL_MAT_TRANSPARENT || (mats[i].kind == SNL_MAT_UBER && mats[i].opacity < 1.0f);
	}
	return SNL_OK;
}

int snl_scene_set_materials(snl_scene *s, const snl_material *mats, int nMats, const snl_texture *tex, int nTex) {
	SNL_REQUIRE(s && nMats >= 0 && nTex >= 0 && (nMats == 0 || mats) && (nTex == 0 || tex), SNL_ERR_INVALID,
				"bad argument");
	return upload_materials(s, mats, nMats, tex, nTex);
}

int snl_scene_set_lights(snl_scene *s, const snl_light *lights, int nLights, const float ambient[3]) {
	SNL_REQUIRE(s && nLights >= 0 && (nLights == 0 || lights), SNL_ERR_INVALID, "bad argument");
	SNL_REQUIRE(nLights <= kMaxLights, SNL_ERR_UNSUPPORTED, "at most %d lights", (int)kMaxLights);
	s->lights.assign(lights, lights + nLights);
	if(ambient)
		memcpy(s->ambient, ambient, 12);
	return SNL_OK;
}

int snl_set_options(snl_scene *s, const int gVals[16]) {
	SNL_REQUIRE(s && gVals, SNL_ERR_INVALID, "null argument");
	SNL_REQUIRE(!gVals[5], SNL_ERR_UNSUPPORTED, "gVals[5] (packet statistics heat-map) has no per-ray equivalent");
	memcpy(s->gVals, gVals, sizeof(s->gVals));
	return SNL_OK;
}

int snl_set_reflection_depth(snl_scene *s, int depth) {
	SNL_REQUIRE(s, SNL_ERR_INVALID, "null scene");
	SNL_REQUIRE(depth == 1 || depth == 2, SNL_ERR_UNSUPPORTED, "reflection depth %d (1 = the reference's cap, 2)", depth);
	s->reflDepth = depth;
	s->tune.key = -1;
	return SNL_OK;
}

int snl_scene_counts(snl_scene *s, int *nNodes, int *nTris, int *nMatIds, int *depth, int *nMaterials) {
	SNL_REQUIRE(s, SNL_ERR_INVALID, "null scene");
	if(nNodes)
		*nNodes = s->devBuilt ? s->nGpuNodes : (int)s->host.nodes.size();
	if(nTris)
		*nTris = (int)s->host.tris.size();
	if(nMatIds)
		*nMatIds = (int)s->host.triMatId.size();
	if(depth)
		*depth = s->devBuilt ? s->gpuDepth : s->host.depth;
	if(nMaterials)
		*nMaterials = s->nMats;
	return SNL_OK;
}

int snl_scene_copy_host_bvh(snl_scene *s, void *nodes, void *tris, void *shTris, int *triMatId) {
	SNL_REQUIRE(s, SNL_ERR_INVALID, "null scene");
	SNL_REQUIRE(!s->devBuilt, SNL_ERR_STATE, "the tree was built on the device: there is no BVH::Node array to copy");
	const HostBVH &h = s->host;
	if(nodes)
		memcpy(nodes, h.nodes.data(), h.nodes.size() * sizeof(RefNode));
	if(tris)
		memcpy(tris, h.tris.data(), h.tris.size() * sizeof(RefTriangle));
	if(shTris)
		memcpy(shTris, h.shTris.data(), h.shTris.size() * sizeof(RefShTri));
	if(triMatId)
		memcpy(triMatId, h.triMatId.data(), h.triMatId.size() * sizeof(int));
	return SNL_OK;
}

int snl_scene_bbox(snl_scene *s, float bmin[3], float bmax[3]) {
	SNL_REQUIRE(s && s->nGpuNodes > 0, SNL_ERR_STATE, "no geometry");
	if(s->devBuilt) {
		memcpy(bmin, s->devBBox, 12);
		memcpy(bmax, s->devBBox + 3, 12);
		return SNL_OK;
	}
	memcpy(bmin, s->host.nodes[0].bmin, 12);
	memcpy(bmax, s->host.nodes[0].bmax, 12);
	return SNL_OK;
}

uint64_t snl_scene_launch_count(snl_scene *s) { return s ? s->launches : 0; }
int snl_scene_render_variant(snl_scene *s) { return s ? s->tune.choice : -1; }
uint64_t snl_scene_stream(snl_scene *s) { return s ? (uint64_t)(uintptr_t)s->stream : 0; }

// ---- compute entry points ---------------------------------------------------------------------------------------

}  // extern "C"
template <class T> static int scratch(snl_scene *s, int slot, size_t count, T **out) {
	SNL_CUDA(s->dScratch[slot].reserve(std::max<size_t>(count * sizeof(T), 16)));
	*out = reinterpret_cast<T *>(s->dScratch[slot].ptr);
	return SNL_OK;
}
extern "C" {

#define SNL_TRY(expr)    \
	do {                 \
		int r_ = (expr); \
		if(r_)           \
			return r_;   \
	} while(0)

static int trace_primary_impl(snl_scene *s, const snl_camera *cam, int W, int H, int *instId, int *triId, float *t, float *bary2);

int snl_trace_primary(snl_scene *s, const snl_camera *cam, int W, int H, int *triId, float *t, float *bary2) {
	return trace_primary_impl(s, cam, W, H, nullptr, triId, t, bary2);
}

int snl_trace_primary_inst(snl_scene *s, const snl_camera *cam, int W, int H, int *instId, int *triId, float *t, float *bary2) {
	SNL_REQUIRE(s && !s->instances.empty(), SNL_ERR_STATE, "no instances: call snl_scene_set_instances first");
	return trace_primary_impl(s, cam, W, H, instId, triId, t, bary2);
}

static int trace_primary_impl(snl_scene *s, const snl_camera *cam, int W, int H, int *instId, int *triId, float *t, float *bary2) {
	SNL_REQUIRE(s && cam && W > 0 && H > 0, SNL_ERR_INVALID, "bad argument");
	SNL_TRY(begin_call(s));
	const size_t n = (size_t)W * H;
	int *dId = nullptr;
	float *dT = nullptr, *dB = nullptr;
	SNL_TRY(scratch(s, 0, n, &dId));
	SNL_TRY(scratch(s, 1, n, &dT));
	SNL_TRY(scratch(s, 2, n * 2, &dB));
	SceneDev S;
	SNL_TRY(make_scene_dev(s, S, nullptr, nullptr));
	SNL_TRY(ensure_pre(s, s->originPre, cam->pos));
	RayGenParams g;
	make_raygen(*cam, W, H, g);
	int *dInst = nullptr;
	if(instId)
		SNL_TRY(scratch(s, 3, n, &dInst));
	SNL_CUDA(launch_primary(S, s->originPre.set(), g, W, H, dId, dT, dB, dInst, s->stream));
	s->launches++;
	SNL_TRY(end_call(s));
	if(instId)
		SNL_CUDA(cudaMemcpyAsync(instId, dInst, n * 4, cudaMemcpyDeviceToHost, s->stream));
	if(triId)
		SNL_CUDA(cudaMemcpyAsync(triId, dId, n * 4, cudaMemcpyDeviceToHost, s->stream));
	if(t)
		SNL_CUDA(cudaMemcpyAsync(t, dT, n * 4, cudaMemcpyDeviceToHost, s->stream));
	if(bary2)
		SNL_CUDA(cudaMemcpyAsync(bary2, dB, n * 8, cudaMemcpyDeviceToHost, s->stream));
	return sync_and_stats(s, nullptr);
}

// uploads caller rays into scratch slots 0 (origins), 1 (dirs), 2 (mask)
static int upload_rays(snl_scene *s, const float *o3, const float *d3, const unsigned char *mask, int n, int sharedOrigin,
					   float **dO, float **dD, unsigned char **dM) {
	const size_t no = sharedOrigin ? 1 : (size_t)n;
	SNL_TRY(scratch(s, 0, no * 3, dO));
	SNL_TRY(scratch(s, 1, (size_t)n * 3, dD));
	SNL_CUDA(cudaMemcpyAsync(*dO, o3, no * 12, cudaMemcpyHostToDevice, s->stream));
	SNL_CUDA(cudaMemcpyAsync(*dD, d3, (size_t)n * 12, cudaMemcpyHostToDevice, s->stream));
	*dM = nullptr;
	if(mask) {
		SNL_TRY(scratch(s, 2, (size_t)n, dM));
		SNL_CUDA(cudaMemcpyAsync(*dM, mask, (size_t)n, cudaMemcpyHostToDevice, s->stream));
	}
	return SNL_OK;
}

int snl_trace_rays(snl_scene *s, const float *o3, const float *d3, const unsigned char *mask, int n, int sharedOrigin,
				   int *triId, float *t, float *bary2) {
	SNL_REQUIRE(s && o3 && d3 && n >= 0, SNL_ERR_INVALID, "bad argument");
	SNL_REQUIRE(s->instances.empty(), SNL_ERR_UNSUPPORTED, "snl_trace_rays is not available on an instanced scene (use snl_trace_primary_inst / snl_render)");
	if(n == 0)
		return SNL_OK;
	SNL_TRY(begin_call(s));
	float *dO, *dD;
	unsigned char *dM;
	SNL_TRY(upload_rays(s, o3, d3, mask, n, sharedOrigin, &dO, &dD, &dM));
	// results: one scratch slot holding triId | t | bary
	unsigned char *res = nullptr;
	SNL_TRY(scratch(s, 3, (size_t)n * 16, &res));
	int *dId = (int *)res;
	float *dT = (float *)(res + (size_t)n * 4), *dB = (float *)(res + (size_t)n * 8);
	SceneDev S;
	SNL_TRY(make_scene_dev(s, S, nullptr, nullptr));
	if(sharedOrigin)
		SNL_TRY(ensure_pre(s, s->originPre, o3));
	SNL_CUDA(launch_trace_rays(S, s->originPre.set(), dO, dD, dM, n, sharedOrigin != 0, dId, dT, dB, s->stream));
	s->launches++;
	SNL_TRY(end_call(s));
	if(triId)
		SNL_CUDA(cudaMemcpyAsync(triId, dId, (size_t)n * 4, cudaMemcpyDeviceToHost, s->stream));
	if(t)
		SNL_CUDA(cudaMemcpyAsync(t, dT, (size_t)n * 4, cudaMemcpyDeviceToHost, s->stream));
	if(bary2)
		SNL_CUDA(cudaMemcpyAsync(bary2, dB, (size_t)n * 8, cudaMemcpyDeviceToHost, s->stream));
	return sync_and_stats(s, nullptr);
}

int snl_trace_shadow(snl_scene *s, const float origin[3], const float *d3, float *tmax_inout, int n) {
	SNL_REQUIRE(s && origin && d3 && tmax_inout && n >= 0, SNL_ERR_INVALID, "bad argument");
	SNL_REQUIRE(s->instances.empty(), SNL_ERR_UNSUPPORTED, "snl_trace_shadow is not available on an instanced scene (use snl_trace_primary_inst / snl_render)");
	if(n == 0)
		return SNL_OK;
	SNL_TRY(begin_call(s));
	float *dD, *dT;
	SNL_TRY(scratch(s, 1, (size_t)n * 3, &dD));
	SNL_TRY(scratch(s, 3, (size_t)n, &dT));
	SNL_CUDA(cudaMemcpyAsync(dD, d3, (size_t)n * 12, cudaMemcpyHostToDevice, s->stream));
	SNL_CUDA(cudaMemcpyAsync(dT, tmax_inout, (size_t)n * 4, cudaMemcpyHostToDevice, s->stream));
	SceneDev S;
	SNL_TRY(make_scene_dev(s, S, nullptr, nullptr));
	SNL_TRY(ensure_pre(s, s->originPre, origin));
	SNL_CUDA(launch_trace_shadow(S, s->originPre.set(), origin, dD, dT, n, s->stream));
	s->launches++;
	SNL_TRY(end_call(s));
	SNL_CUDA(cudaMemcpyAsync(tmax_inout, dT, (size_t)n * 4, cudaMemcpyDeviceToHost, s->stream));
	return sync_and_stats(s, nullptr);
}

int snl_raytrace(snl_scene *s, const float *o3, const float *d3, const unsigned char *mask, int n, int sharedOrigin,
				 float *rgb3, snl_stats *stats) {
	SNL_REQUIRE(s && o3 && d3 && rgb3 && n >= 0, SNL_ERR_INVALID, "bad argument");
	SNL_REQUIRE(s->instances.empty(), SNL_ERR_UNSUPPORTED, "snl_raytrace is not available on an instanced scene (use snl_trace_primary_inst / snl_render)");
	if(n == 0)
		return SNL_OK;
	SNL_TRY(begin_call(s));
	float *dO, *dD, *dC;
	unsigned char *dM;
	SNL_TRY(upload_rays(s, o3, d3, mask, n, sharedOrigin, &dO, &dD, &dM));
	SNL_TRY(scratch(s, 3, (size_t)n * 3, &dC));
	SceneDev S;
	SNL_TRY(make_scene_dev(s, S, nullptr, nullptr));
	if(sharedOrigin)
		SNL_TRY(ensure_pre(s, s->originPre, o3));
	SNL_CUDA(launch_raytrace(S, s->originPre.set(), dO, dD, dM, n, sharedOrigin != 0, dC, s->stream));
	s->launches++;
	SNL_TRY(end_call(s));
	SNL_CUDA(cudaMemcpyAsync(rgb3, dC, (size_t)n * 12, cudaMemcpyDeviceToHost, s->stream));
	return sync_and_stats(s, stats);
}

static int render_common(snl_scene *s, const snl_camera *cam, int W, int H, const int *tiles, int nTiles, unsigned char *dOut,
						 size_t pitch, int rank, unsigned char *hostOut = nullptr, size_t hostPitch = 0, bool *hostDone = nullptr);

// Tile list -> packet list (RenderTask::Work's x/y loops, src/render.cpp:67-68), cached between calls.
static int ensure_packets(snl_scene *s, const int *tiles, int nTiles, int W, int H) {
	std::vector<int> key(tiles, tiles + (size_t)nTiles * 4);
	key.push_back(W);
	key.push_back(H);
	if(key == s->lastTiles && s->dPackets.ptr)
		return SNL_OK;
	std::vector<int4> packets;
	for(int i = 0; i < nTiles; i++) {
		const int x = tiles[i * 4 + 0], y = tiles[i * 4 + 1], w = tiles[i * 4 + 2], h = tiles[i * 4 + 3];
		SNL_REQUIRE(x >= 0 && y >= 0 && w > 0 && h > 0 && x + w <= W && y + h <= H, SNL_ERR_INVALID,
					"tile %d (%d,%d,%d,%d) outside the %dx%d image", i, x, y, w, h, W, H);
		SNL_REQUIRE(x % 16 == 0 && (w % 16 == 0 || x + w == W), SNL_ERR_INVALID,
					"tile %d: x and w must be multiples of the packet width 16", i);
		for(int py = 0; py < h; py += 16)
			for(int px = 0; px < w; px += 16)
				packets.push_back(make_int4(x + px, y + py, std::min(16, w - px), std::min(16, h - py)));
	}
	SNL_CUDA(cudaStreamSynchronize(s->stream));
	SNL_CUDA(s->dPackets.reserve(std::max<size_t>(packets.size(), 1)));
	if(!packets.empty())
		SNL_CUDA(cudaMemcpy(s->dPackets.ptr, packets.data(), packets.size() * sizeof(int4), cudaMemcpyHostToDevice));
	s->nPackets = (int)packets.size();
	s->lastTiles.swap(key);
	return SNL_OK;
}

// hostOut != null (snl_render): a large full frame is rendered in row bands, each on its own stream and followed by the
// device->host copy of its rows; *hostDone tells the caller the copies are already queued (and joined into s->stream).
static int render_common(snl_scene *s, const snl_camera *cam, int W, int H, const int *tiles, int nTiles, unsigned char *dOut,
						 size_t pitch, int rank, unsigned char *hostOut, size_t hostPitch, bool *hostDone) {
	SceneDev S;
	RayGenParams g;
	SNL_TRY(prepare_frame(s, cam, W, H, rank, S, g));
	const int4 *packets = nullptr;
	if(tiles) {
		SNL_TRY(ensure_packets(s, tiles, nTiles, W, H));
		packets = s->dPackets.ptr;
	}
	const int nPk = packets ? s->nPackets : ((W + 15) / 16) * ((H + 15) / 16);
	if(S.gvRefl && !S.gvDepth && !s->fused && !S.reflDepth2 && !S.hasTransp && !S.gvSuper && !S.instances && !S.cullMask && nPk > 0) {
		// reflection bounce: wavefront pipeline with queue-driven persistent traversal of the incoherent rays
		const size_t NL = (size_t)nPk * kPacket, nl = std::max<size_t>(S.nLights, 1);
		SNL_REQUIRE(NL * nl < 0xffffffffull, SNL_ERR_UNSUPPORTED, "frame too large for 32-bit ray ids");
		WfBuffers B;
		SNL_CUDA(s->wfRayO.reserve(NL));
		SNL_CUDA(s->wfRayD.reserve(NL));
		SNL_CUDA(s->wfHit1.reserve(NL));
		SNL_CUDA(s->wfSt0.reserve(4 * NL));
		SNL_CUDA(s->wfSt1.reserve(2 * NL));
		SNL_CUDA(s->wfSdir.reserve(nl * NL));
		SNL_CUDA(s->wfSdot.reserve(nl * NL));
		SNL_CUDA(s->wfRq.reserve(NL));
		SNL_CUDA(s->wfSq.reserve(nl * NL));
		SNL_CUDA(s->wfOccl.reserve(NL));
		SNL_CUDA(s->wfCount.reserve(4));
		SNL_CUDA(s->wfLightSets.reserve(kMaxLights));
		SNL_CUDA(cudaMemcpyAsync(s->wfLightSets.ptr, S.lightPre, sizeof(S.lightPre), cudaMemcpyHostToDevice, s->stream));
		B.rayO = s->wfRayO.ptr;
		B.rayD = s->wfRayD.ptr;
		B.hit1 = s->wfHit1.ptr;
		B.st0 = s->wfSt0.ptr;
		B.st1 = s->wfSt1.ptr;
		B.sdir = s->wfSdir.ptr;
		B.sdot = s->wfSdot.ptr;
		B.rq = s->wfRq.ptr;
		B.sq = s->wfSq.ptr;
		B.occl = s->wfOccl.ptr;
		B.qcount = s->wfCount.ptr;
		B.lightSets = s->wfLightSets.ptr;
		B.NL = (unsigned)NL;
		int launches = 0;
		SNL_CUDA(launch_render_wavefront(S, s->originPre.set(), g, W, H, packets, s->nPackets, B, s->smCount, dOut, pitch,
										 s->stream, &launches));
		s->launches += launches;
		return SNL_OK;
	}
	// pick k_render's register budget for this workload: the first frame is rendered three times with each (same
	// inputs, same output buffer), the first pair is warm-up, the faster variant is kept
	snl_scene::RenderTune &T = s->tune;
	const long long key = ((long long)W << 40) ^ ((long long)H << 24) ^ ((long long)nPk << 4) ^ (S.gvRefl << 3) ^ (S.gvFull << 2) ^
						  (S.gvDepth << 1) ^ ((long long)S.nLights << 60) ^ ((long long)s->nTris << 8);
	if(T.key != key) {
		T = snl_scene::RenderTune();
		T.key = key;
		const char *pin = getenv("SNL_RENDER_VARIANT");
		if(pin && (pin[0] == '0' || pin[0] == '1'))
			T.choice = pin[0] - '0';
		else if(S.hasTransp || S.gvSuper || S.instances || S.reflDepth2 || nPk < 4 * s->smCount)
			T.choice = 0;  // only the plain kernel has the second variant; tiny frames are not worth tuning
	}
	if(T.choice < 0) {
		cudaEvent_t e0, e1;
		SNL_CUDA(cudaEventCreate(&e0));
		SNL_CUDA(cudaEventCreate(&e1));
		cudaError_t err = cudaSuccess;
		for(int it = 0; it < 6 && err == cudaSuccess; it++) {
			cudaEventRecord(e0, s->stream);
			err = launch_render(S, s->originPre.set(), g, W, H, packets, s->nPackets, dOut, pitch, it & 1, s->stream);
			cudaEventRecord(e1, s->stream);
			if(err == cudaSuccess)
				err = cudaEventSynchronize(e1);
			float ms = 0.0f;
			if(err == cudaSuccess)
				err = cudaEventElapsedTime(&ms, e0, e1);
			if(it >= 2)
				T.ms[it & 1] += ms;
			s->launches++;
		}
		cudaEventDestroy(e0);
		cudaEventDestroy(e1);
		SNL_CUDA(err);
		T.choice = T.ms[1] < T.ms[0] ? 1 : 0;
		// the trial frames counted their rays and their time: start this call's accounting over
		SNL_CUDA(cudaMemsetAsync(s->dCounters.ptr, 0, sizeof(Counters), s->stream));
		SNL_CUDA(cudaEventRecord(s->ev0, s->stream));
	}
	const int variant = T.choice;
	const char *bandsEnv = getenv("SNL_RENDER_BANDS");  // number of bands; "0" / "1": one launch, copy afterwards
	// frames of a millisecond and more hide more of the copy with eight bands (+1.2 % e2e on sponza_like); short ones lose
	// more to the extra launches than they gain (-1 % on feline): four - unless the frame itself is big: a 24.9 MB 4K frame
	// takes 0.45 ms to copy, and eight bands expose half as much of it as four (0.814 vs 0.858 ms on barracks 4K)
	const bool bigFrame = (size_t)W * H * 3 >= ((size_t)12 << 20);
	const int nBands = bandsEnv ? std::min(atoi(bandsEnv), (int)snl_scene::kMaxBands)
								: (s->lastCallMs > 0.8f || bigFrame ? (int)snl_scene::kMaxBands : (int)snl_scene::kBands);
	if(hostOut && hostDone && !packets && nPk >= 8 * s->smCount && nBands > 1) {
		SNL_TRY(ensure_band_streams(s));
		const int px = (W + 15) / 16, py = (H + 15) / 16;
		SNL_CUDA(cudaEventRecord(s->forkEv, s->stream));
		for(int b = 0; b < nBands; b++) {
			const int r0 = py * b / nBands, r1 = py * (b + 1) / nBands;
			if(r1 <= r0)
				continue;
			cudaStream_t st = s->bandStream[b];
			SNL_CUDA(cudaStreamWaitEvent(st, s->forkEv, 0));
			SNL_CUDA(launch_render(S, s->originPre.set(), g, W, H, nullptr, 0, dOut, pitch, variant, st, r0 * px, (r1 - r0) * px));
			s->launches++;
			const int y0 = r0 * 16, y1 = std::min(H, r1 * 16);
			SNL_CUDA(cudaMemcpy2DAsync(hostOut + (size_t)y0 * hostPitch, hostPitch, dOut + (size_t)y0 * pitch, pitch, (size_t)W * 3,
									   y1 - y0, cudaMemcpyDeviceToHost, st));
			SNL_CUDA(cudaEventRecord(s->bandEv[b], st));
			SNL_CUDA(cudaStreamWaitEvent(s->stream, s->bandEv[b], 0));
		}
		*hostDone = true;
		return SNL_OK;
	}
	SNL_CUDA(launch_render(S, s->originPre.set(), g, W, H, packets, s->nPackets, dOut, pitch, variant, s->stream));
	s->launches++;
	return SNL_OK;
}

int snl_render_dev(snl_scene *s, const snl_camera *cam, int W, int H, const int *tiles_xywh, int nTiles, void *bgr8_dev,
				   size_t pitch, int rank) {
	SNL_REQUIRE(s && cam && bgr8_dev && W > 0 && H > 0 && pitch >= (size_t)W * 3, SNL_ERR_INVALID, "bad argument");
	SNL_REQUIRE(!tiles_xywh || nTiles >= 0, SNL_ERR_INVALID, "bad tile list");
	SNL_TRY(begin_call(s));
	SNL_TRY(render_common(s, cam, W, H, tiles_xywh, nTiles, (unsigned char *)bgr8_dev, pitch, rank));
	return end_call(s);
}

int snl_scene_sync(snl_scene *s, snl_stats *stats) {
	SNL_REQUIRE(s, SNL_ERR_INVALID, "null scene");
	SNL_CUDA(cudaSetDevice(s->device));
	return sync_and_stats(s, stats);
}

int snl_render(snl_scene *s, const snl_camera *cam, int W, int H, const int *tiles_xywh, int nTiles,
			   unsigned char *bgr8_host, size_t pitch, int rank, snl_stats *stats) {
	SNL_REQUIRE(s && cam && bgr8_host && W > 0 && H > 0 && pitch >= (size_t)W * 3, SNL_ERR_INVALID, "bad argument");
	SNL_REQUIRE(!tiles_xywh || nTiles >= 0, SNL_ERR_INVALID, "bad tile list");
	SNL_TRY(begin_call(s));
	const size_t dpitch = (size_t)W * 3;
	SNL_CUDA(s->dImage.reserve(dpitch * H));
	bool copied = false;
	SNL_TRY(render_common(s, cam, W, H, tiles_xywh, nTiles, s->dImage.ptr, dpitch, rank, bgr8_host, pitch, &copied));
	SNL_TRY(end_call(s));
	if(copied)
		;  // row bands: every band's copy is already queued behind its kernel
	else if(!tiles_xywh)
		SNL_CUDA(cudaMemcpy2DAsync(bgr8_host, pitch, s->dImage.ptr, dpitch, dpitch, H, cudaMemcpyDeviceToHost, s->stream));
	else
		for(int i = 0; i < nTiles; i++) {
			const int x = tiles_xywh[i * 4 + 0], y = tiles_xywh[i * 4 + 1], w = tiles_xywh[i * 4 + 2], h = tiles_xywh[i * 4 + 3];
			SNL_CUDA(cudaMemcpy2DAsync(bgr8_host + (size_t)y * pitch + (size_t)x * 3, pitch,
									   s->dImage.ptr + (size_t)y * dpitch + (size_t)x * 3, dpitch, (size_t)w * 3, h,
									   cudaMemcpyDeviceToHost, s->stream));
		}
	return sync_and_stats(s, stats);
}

int snl_render_farm(snl_scene *s, const snl_camera *cam, int W, int H, unsigned char *data, const int *coords,
					const int *offsets, int nTiles, int rank, snl_stats *stats) {
	SNL_REQUIRE(s && cam && data && coords && offsets && W > 0 && H > 0 && nTiles >= 0, SNL_ERR_INVALID, "bad argument");
	size_t lo = ~(size_t)0, hi = 0, sum = 0;
	for(int i = 0; i < nTiles; i++) {
		const int w = coords[i * 4 + 2], h = coords[i * 4 + 3];
		SNL_REQUIRE(offsets[i] >= 0 && w > 0 && h > 0, SNL_ERR_INVALID, "bad tile / offset");
		const size_t bytes = (size_t)w * h * 3;
		lo = std::min(lo, (size_t)offsets[i]);
		hi = std::max(hi, (size_t)offsets[i] + bytes);
		sum += bytes;
	}
	SNL_REQUIRE(hi < ((size_t)1 << 31), SNL_ERR_INVALID, "tile buffer too large");
	SNL_TRY(begin_call(s));
	const size_t dpitch = (size_t)W * 3;
	SNL_CUDA(s->dImage.reserve(dpitch * H));
	SNL_TRY(render_common(s, cam, W, H, coords, nTiles, s->dImage.ptr, dpitch, rank));  // validates the rectangles
	if(nTiles > 0) {
		SNL_CUDA(s->dFarmTiles.reserve(nTiles));
		SNL_CUDA(s->dFarmOffsets.reserve(nTiles));
		SNL_CUDA(s->dFarm.reserve(hi));
		SNL_CUDA(cudaMemcpyAsync(s->dFarmTiles.ptr, coords, (size_t)nTiles * sizeof(int4), cudaMemcpyHostToDevice, s->stream));
		SNL_CUDA(cudaMemcpyAsync(s->dFarmOffsets.ptr, offsets, (size_t)nTiles * sizeof(int), cudaMemcpyHostToDevice, s->stream));
		SNL_CUDA(launch_farm_pack(s->dImage.ptr, dpitch, s->dFarmTiles.ptr, s->dFarmOffsets.ptr, nTiles, s->dFarm.ptr, s->stream));
		s->launches++;
	}
	SNL_TRY(end_call(s));
	if(nTiles > 0) {
		if(sum == hi - lo)  // the tiles fill [lo, hi) (server.cpp lays them out back to back): one copy
			SNL_CUDA(cudaMemcpyAsync(data + lo, s->dFarm.ptr + lo, hi - lo, cudaMemcpyDeviceToHost, s->stream));
		else
			for(int i = 0; i < nTiles; i++)
				SNL_CUDA(cudaMemcpyAsync(data + offsets[i], s->dFarm.ptr + offsets[i],
										 (size_t)coords[i * 4 + 2] * coords[i * 4 + 3] * 3, cudaMemcpyDeviceToHost, s->stream));
	}
	return sync_and_stats(s, stats);
}

// ---- frames shared between the ranks of one box (CUDA IPC over NVLink) ----------------------------------------------

int snl_frame_create(snl_scene *s, int W, int H, void **bgr8_dev, unsigned char ipc_handle[SNL_IPC_HANDLE_BYTES]) {
	SNL_REQUIRE(s && bgr8_dev && W > 0 && H > 0, SNL_ERR_INVALID, "bad argument");
	static_assert(sizeof(cudaIpcMemHandle_t) == SNL_IPC_HANDLE_BYTES, "cudaIpcMemHandle_t is 64 bytes");
	SNL_CUDA(cudaSetDevice(s->device));
	void *p = nullptr;
	SNL_CUDA(cudaMalloc(&p, (size_t)W * H * 3));
	SNL_CUDA(cudaMemset(p, 0, (size_t)W * H * 3));
	if(ipc_handle) {
		cudaIpcMemHandle_t h;
		cudaError_t e = cudaIpcGetMemHandle(&h, p);
		if(e != cudaSuccess) {
			cudaFree(p);
			set_error("cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
			return SNL_ERR_CUDA;
		}
		memcpy(ipc_handle, &h, sizeof(h));
	}
	s->ownedFrames.push_back(p);
	*bgr8_dev = p;
	return SNL_OK;
}

int snl_frame_open(snl_scene *s, const unsigned char ipc_handle[SNL_IPC_HANDLE_BYTES], void **bgr8_dev) {
	SNL_REQUIRE(s && ipc_handle && bgr8_dev, SNL_ERR_INVALID, "bad argument");
	SNL_CUDA(cudaSetDevice(s->device));
	cudaIpcMemHandle_t h;
	memcpy(&h, ipc_handle, sizeof(h));
	void *p = nullptr;
	SNL_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
	s->peerFrames.push_back(p);
	*bgr8_dev = p;
	return SNL_OK;
}

int snl_frame_close(snl_scene *s, void *p) {
	SNL_REQUIRE(s && p, SNL_ERR_INVALID, "bad argument");
	SNL_CUDA(cudaSetDevice(s->device));
	SNL_CUDA(cudaStreamSynchronize(s->stream));
	for(size_t i = 0; i < s->ownedFrames.size(); i++)
		if(s->ownedFrames[i] == p) {
			s->ownedFrames.erase(s->ownedFrames.begin() + i);
			SNL_CUDA(cudaFree(p));
			return SNL_OK;
		}
	for(size_t i = 0; i < s->peerFrames.size(); i++)
		if(s->peerFrames[i] == p) {
			s->peerFrames.erase(s->peerFrames.begin() + i);
			SNL_CUDA(cudaIpcCloseMemHandle(p));
			return SNL_OK;
		}
	set_error("not a frame of this scene");
	return SNL_ERR_INVALID;
}

int snl_frame_read(snl_scene *s, const void *bgr8_dev, int W, int H, unsigned char *bgr8_host, size_t host_pitch) {
	SNL_REQUIRE(s && bgr8_dev && bgr8_host && W > 0 && H > 0 && host_pitch >= (size_t)W * 3, SNL_ERR_INVALID, "bad argument");
	SNL_CUDA(cudaSetDevice(s->device));
	SNL_CUDA(cudaMemcpy2DAsync(bgr8_host, host_pitch, bgr8_dev, (size_t)W * 3, (size_t)W * 3, H, cudaMemcpyDeviceToHost,
							   s->stream));
	SNL_CUDA(cudaStreamSynchronize(s->stream));
	return SNL_OK;
}

static int set_host_scene(snl_scene *s, const HostScene &hs) {
	SNL_CUDA(cudaSetDevice(s->device));
	s->host = hs.bvh;
	SNL_TRY(upload_geometry(s));
	std::vector<snl_material> mats;
	for(auto &m : hs.mats)
		mats.push_back(m.m);
	std::vector<snl_texture> texs;
	for(auto &t : hs.texs)
		texs.push_back(snl_texture{t.data.data(), t.data.size(), t.width, t.height, t.mips});
	return upload_materials(s, mats.data(), (int)mats.size(), texs.data(), (int)texs.size());
}

snl_host_scene *snl_host_scene_load_obj(const char *objPath, const char *mtlPath, const char *texDir, int flipNormalsFlag,
										int buildFlags) {
	if(!objPath) {
		set_error("null path");
		return nullptr;
	}
	snl_host_scene *h = new snl_host_scene;
	std::string err;
	if(!load_obj_scene(objPath, mtlPath, texDir, flipNormalsFlag, buildFlags, h->scene, err)) {
		set_error("%s", err.c_str());
		delete h;
		return nullptr;
	}
	return h;
}

void snl_host_scene_free(snl_host_scene *h) { delete h; }

int snl_host_scene_counts(const snl_host_scene *h, int *nNodes, int *nTris, int *nMatIds, int *depth, int *nMaterials,
						  int *nTextures) {
	SNL_REQUIRE(h, SNL_ERR_INVALID, "null handle");
	const HostBVH &b = h->scene.bvh;
	if(nNodes)
		*nNodes = (int)b.nodes.size();
	if(nTris)
		*nTris = (int)b.tris.size();
	if(nMatIds)
		*nMatIds = (int)b.triMatId.size();
	if(depth)
		*depth = b.depth;
	if(nMaterials)
		*nMaterials = (int)h->scene.mats.size();
	if(nTextures)
		*nTextures = (int)h->scene.texs.size();
	return SNL_OK;
}

int snl_host_scene_copy(const snl_host_scene *h, void *nodes, void *tris, void *shTris, int *triMatId) {
	SNL_REQUIRE(h, SNL_ERR_INVALID, "null handle");
	const HostBVH &b = h->scene.bvh;
	if(nodes)
		memcpy(nodes, b.nodes.data(), b.nodes.size() * sizeof(RefNode));
	if(tris)
		memcpy(tris, b.tris.data(), b.tris.size() * sizeof(RefTriangle));
	if(shTris)
		memcpy(shTris, b.shTris.data(), b.shTris.size() * sizeof(RefShTri));
	if(triMatId)
		memcpy(triMatId, b.triMatId.data(), b.triMatId.size() * sizeof(int));
	return SNL_OK;
}

int snl_host_scene_material(const snl_host_scene *h, int idx, snl_material *out) {
	SNL_REQUIRE(h && out && idx >= 0 && idx < (int)h->scene.mats.size(), SNL_ERR_INVALID, "bad argument");
	*out = h->scene.mats[idx].m;
	return SNL_OK;
}

int snl_host_scene_texture(const snl_host_scene *h, int idx, snl_texture *out) {
	SNL_REQUIRE(h && out && idx >= 0 && idx < (int)h->scene.texs.size(), SNL_ERR_INVALID, "bad argument");
	const HostTexture &t = h->scene.texs[idx];
	*out = snl_texture{t.data.data(), t.data.size(), t.width, t.height, t.mips};
	return SNL_OK;
}

int snl_scene_set_host_scene(snl_scene *s, const snl_host_scene *h) {
	SNL_REQUIRE(s && h, SNL_ERR_INVALID, "null argument");
	return set_host_scene(s, h->scene);
}

int snl_scene_load_obj(snl_scene *s, const char *objPath, const char *mtlPath, const char *texDir, int flipNormalsFlag,
					   int buildFlags) {
	SNL_REQUIRE(s && objPath, SNL_ERR_INVALID, "null argument");
	HostScene hs;
	std::string err;
	if(!load_obj_scene(objPath, mtlPath, texDir, flipNormalsFlag, buildFlags & ~SNL_BUILD_SAH_BINNED, hs, err)) {
		set_error("%s", err.c_str());
		return SNL_ERR_IO;
	}
	if(buildFlags & SNL_BUILD_DEVICE)
		s->deviceBuilder = (buildFlags & SNL_BUILD_SAH_BINNED) ? SNL_BUILDER_SAH : SNL_BUILDER_LBVH;
	return set_host_scene(s, hs);
}

}  // extern "C"

int snl::render_local(snl_scene *s, const snl_camera *cam, int W, int H, const int *tiles, int nTiles, unsigned char *dOut,
					   size_t pitch, int rank) {
	return ::render_common(s, cam, W, H, tiles, nTiles, dOut, pitch, rank, nullptr, 0, nullptr);
}
