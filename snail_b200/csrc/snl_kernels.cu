// snail_b200 — kernel entry points + launch wrappers (sm_100a). See snl_device.cuh for the device functions.
#include "snl_device.cuh"
#include "snl_kernels.h"

#include <cstring>

namespace snl {

// Per-origin records, one launch per origin and frame: Triangle::Collide's sharedOrigin precomputation
// (src/triangle.cpp:9-18) for every triangle, hoisted from per packet to per frame.
__global__ void __launch_bounds__(256) k_pretransform(const GenTri *__restrict__ gtris, int nTris, float ox, float oy, float oz,
													 PreTri *__restrict__ out) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= nTris)
		return;
	const float4 ta = ldg4(&gtris[i].a), tba = ldg4(&gtris[i].ba), tca = ldg4(&gtris[i].ca), tn = ldg4(&gtris[i].n);
	const float it0 = ta.w;
	const float tx = ox - ta.x, ty = oy - ta.y, tz = oz - ta.z;
	float4 u, v, n;
	// tvec0 = (ba ^ tvec) * it0
	v.x = (tba.y * tz - tba.z * ty) * it0;
	v.y = (tba.z * tx - tba.x * tz) * it0;
	v.z = (tba.x * ty - tba.y * tx) * it0;
	v.w = 0.0f;
	// tvec1 = (tvec ^ ca) * it0
	u.x = (ty * tca.z - tz * tca.y) * it0;
	u.y = (tz * tca.x - tx * tca.z) * it0;
	u.z = (tx * tca.y - ty * tca.x) * it0;
	u.w = -(tx * tn.x + ty * tn.y + tz * tn.z);
	n = make_float4(tn.x, tn.y, tn.z, 0.0f);
	float4 *o = reinterpret_cast<float4 *>(out + i);
	o[0] = u;
	o[1] = v;
	o[2] = n;
}

// ---- BVH::TraversePrimary on generated primary rays (snl_trace_primary) -------------------------------------
// grid = packets (16x16 pixels), block = 256.
template <bool INST>
__global__ void __launch_bounds__(kPacket) k_primary(SceneDev S, const PreSet pre, RayGenParams g, int W, int H,
												  int packetsX, int *__restrict__ triId, float *__restrict__ t,
												  float *__restrict__ bary, int *__restrict__ instId) {
	const int px0 = (blockIdx.x % packetsX) * 16, py0 = (blockIdx.x / packetsX) * 16;
	const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
	const int x = px0 + tx, y = py0 + ty;
	PacketCounts pc = {};
	TravCount tc = {0, 0, 0};
	const bool inside = x < W && y < H;
	if(inside) {
		float dx, dy, dz;
		gen_primary(g, px0, py0, tx, ty, dx, dy, dz);
		Hit h;
		h.dist = CUDART_INF_F;
		h.obj = 0;
		h.inst = 0;
		h.bu = h.bv = 0.0f;
		if(INST)  // DBVH::TraversePrimary(Context<1,0>&)
			traverse_nearest_inst<true>(S, g.origin[0], g.origin[1], g.origin[2], dx, dy, dz, safe_inv(dx), safe_inv(dy),
										safe_inv(dz), h, tc);
		else
			traverse_nearest<true>(S.nodes, pre.tris, g.origin[0], g.origin[1], g.origin[2], dx, dy, dz,
								   safe_inv(dx), safe_inv(dy), safe_inv(dz), h, tc);
		pc.primary = 1;
		const size_t o = (size_t)y * W + x;
		const bool hit = h.dist < CUDART_INF_F;
		if(triId)
			triId[o] = hit ? h.obj : -1;
		if(INST && instId)
			instId[o] = hit ? h.inst : -1;
		if(t)
			t[o] = h.dist;
		if(bary) {
			bary[o * 2 + 0] = hit ? h.bu : 0.0f;
			bary[o * 2 + 1] = hit ? h.bv : 0.0f;
		}
	}
	__syncwarp();
	flush_counts(S, pc, tc);
}

// ---- BVH::TraversePrimary on caller rays (snl_trace_rays) ---------------------------------------------------
template <bool SHARED>
__global__ void __launch_bounds__(kPacket) k_trace_rays(SceneDev S, const PreSet pre, const float *__restrict__ o3,
													 const float *__restrict__ d3, const unsigned char *__restrict__ mask,
													 int n, int *__restrict__ triId, float *__restrict__ t,
													 float *__restrict__ bary) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	PacketCounts pc = {};
	TravCount tc = {0, 0, 0};
	if(i < n) {
		const bool active = !mask || mask[i];
		Hit h;
		h.dist = CUDART_INF_F;
		h.obj = 0;
		h.bu = h.bv = 0.0f;
		if(active) {
			const float *o = SHARED ? o3 : o3 + (size_t)i * 3;
			const float ox = o[0], oy = o[1], oz = o[2];
			const float dx = d3[(size_t)i * 3 + 0], dy = d3[(size_t)i * 3 + 1], dz = d3[(size_t)i * 3 + 2];
			traverse_nearest<SHARED>(S.nodes, SHARED ? (const void *)pre.tris : (const void *)S.gtris, ox, oy, oz, dx, dy, dz,
									 safe_inv(dx), safe_inv(dy), safe_inv(dz), h, tc);
			pc.primary = 1;
		}
		const bool hit = h.dist < CUDART_INF_F;
		if(triId)
			triId[i] = hit ? h.obj : -1;
		if(t)
			t[i] = h.dist;
		if(bary) {
			bary[(size_t)i * 2 + 0] = hit ? h.bu : 0.0f;
			bary[(size_t)i * 2 + 1] = hit ? h.bv : 0.0f;
		}
	}
	__syncwarp();
	flush_counts(S, pc, tc);
}

// ---- BVH::TraverseShadow on caller rays from one origin (snl_trace_shadow) ----------------------------------
__global__ void __launch_bounds__(kPacket) k_trace_shadow(SceneDev S, const PreSet pre, float ox, float oy, float oz,
													   const float *__restrict__ d3, float *__restrict__ tmax, int n) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	PacketCounts pc = {};
	TravCount tc = {0, 0, 0};
	if(i < n) {
		const float dist = tmax[i];
		if(dist >= 0.0f) {
			const float dx = d3[(size_t)i * 3 + 0], dy = d3[(size_t)i * 3 + 1], dz = d3[(size_t)i * 3 + 2];
			if(traverse_shadow(S.nodes, pre.tris, ox, oy, oz, dx, dy, dz, safe_inv(dx), safe_inv(dy),
							   safe_inv(dz), dist, tc))
				tmax[i] = -CUDART_INF_F;
			pc.shadow = 1;
		}
	}
	__syncwarp();
	flush_counts(S, pc, tc);
}

// ---- Scene::RayTrace on caller rays, 256-ray packets (snl_raytrace) -----------------------------------------
template <bool SHARED, bool TRANSP>
__global__ void __launch_bounds__(kPacket) k_raytrace(SceneDev S, const PreSet pre, const float *__restrict__ o3,
												   const float *__restrict__ d3, const unsigned char *__restrict__ mask,
												   int n, float *__restrict__ rgb3) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	// the last packet is padded with masked-off copies of the last ray (a RayGroup always holds whole quads)
	const int src = i < n ? i : n - 1;
	// RayGroup<1,0> has no mask (src/ray_group.h:74-160)
	const bool active = i < n && (SHARED || !mask || mask[i]);
	const float *o = SHARED ? o3 : o3 + (size_t)src * 3;
	const float ox = o[0], oy = o[1], oz = o[2];
	const float dx = d3[(size_t)src * 3 + 0], dy = d3[(size_t)src * 3 + 1], dz = d3[(size_t)src * 3 + 2];
	const bool all = __syncthreads_and(active);
	packet_vote_init();
	PacketCounts pc = {};
	TravCount tc = {0, 0, 0};
	float r = 0.0f, g = 0.0f, b = 0.0f;
	raytrace_lane<0, SHARED, TRANSP>(S, pre, active, (!SHARED && !all) ? -1 : 0, ox, oy, oz, dx, dy, dz, r, g, b, pc, tc);
	if(i < n) {
		rgb3[(size_t)i * 3 + 0] = r;
		rgb3[(size_t)i * 3 + 1] = g;
		rgb3[(size_t)i * 3 + 2] = b;
	}
	__syncwarp();
	flush_counts(S, pc, tc);
}

// ConvColor (src/render.cpp:11-17): trunc(clamp(c * 255, 0, 255)).
__device__ __forceinline__ unsigned conv_channel(float c) {
	float v = c * 255.0f;
	v = v > 0.0f ? v : 0.0f;      // Max(obj, min)
	v = v < 255.0f ? v : 255.0f;  // Min(.., max)
	return (unsigned)int(v);
}

// ---- Render(): RenderTask::Work per packet (src/render.cpp:47-211) ------------------------------------------
// grid = packets; packet p covers pixels [x0, x0+16) x [y0, y0+16) clipped to (cw, ch). `packets` == null: full image.
#ifndef SNL_RENDER_MIN_BLOCKS
#define SNL_RENDER_MIN_BLOCKS 4
#endif
// FLAGS: 1 = the scene has a material that can start a transparency pass, 2 = instanced scene (DBVH), 8 = two reflection
// bounces (plain scenes only). SUPER: gVals[9], every 16x16 output packet is
// traced as four 16x16 packets of the 2W x 2H image and box-filtered (src/render.cpp:71-110).
// MINB: resident CTAs per SM the register allocation aims for. 4 (64 registers) suits scenes whose hot set sits in L1;
// 5 (48 registers, a few more spills) hides more latency on large scenes: +1.9 % on sponza_like, -2 % on feline.
// The scene picks per workload by timing both on its first frames (snl_capi.cu render_common).
// One packet of RenderTask::Work: traces and shades the 16x16 pixels at (px0, py0) and leaves their bytes in `tile`
// (SUPER: stores them itself). All 256 threads of the CTA call it.
template <int FLAGS, bool SUPER>
__device__ __forceinline__ void render_packet(const SceneDev &S, const PreSet &pre, const RayGenParams &g, int px0, int py0, int cw,
											  int ch, unsigned char *__restrict__ out, size_t pitch, unsigned char *tile,
											  PacketCounts &pc, TravCount &tc) {
	const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
	// the plain kernel settles the per-packet light cull's ray accounting at its end instead of with barriers per level
	constexpr int LFLAGS = ((FLAGS & ~8) == 0 && !SUPER) ? (4 | FLAGS) : FLAGS;
	constexpr bool DEFER = (LFLAGS & 4) != 0;
	if(DEFER && S.cullMask) {
		cull_state_init();
		__syncthreads();
	}
	// Like the reference we trace the whole 16x16 packet even when it overhangs the image or the tile
	// (src/render.cpp:112-115): overhanging lanes take part in the 16-lane shading blocks and in the packet-wide
	// selector.All() test of the reflection pass (src/scene_trace.cpp:615-617), and TreeStats counts them.
	const bool active = true;
	for(int k = 0; k < (SUPER ? 4 : 1); k++) {
		float dx, dy, dz;
		if(SUPER)  // offx / offy of src/render.cpp:72
			gen_primary(g, px0 * 2 + (k & 1) * 16, py0 * 2 + (k >> 1) * 16, tx, ty, dx, dy, dz);
		else
			gen_primary(g, px0, py0, tx, ty, dx, dy, dz);
		float r = 0.0f, gg = 0.0f, b = 0.0f;
		raytrace_lane<0, true, LFLAGS>(S, pre, active, 0, g.origin[0], g.origin[1], g.origin[2], dx, dy, dz, r, gg, b, pc, tc);
		int ox = tx, oy = ty;
		bool store = true;
		if(SUPER) {
			// (tcolors[q] + tcolors[q + line]) * 0.25: rows 2j and 2j+1 of the packet are lanes l and l+16 of one warp;
			// then col[0] + col[1] / col[2] + col[3]: even + odd pixel (src/render.cpp:92-101)
			r = (r + __shfl_down_sync(SNL_FULL, r, 16)) * 0.25f;
			gg = (gg + __shfl_down_sync(SNL_FULL, gg, 16)) * 0.25f;
			b = (b + __shfl_down_sync(SNL_FULL, b, 16)) * 0.25f;
			r = r + __shfl_down_sync(SNL_FULL, r, 1);
			gg = gg + __shfl_down_sync(SNL_FULL, gg, 1);
			b = b + __shfl_down_sync(SNL_FULL, b, 1);
			store = !(ty & 1) && !(tx & 1);
			ox = (k & 1) * 8 + (tx >> 1);
			oy = (k >> 1) * 8 + (ty >> 1);
		}
		if(S.gvTint) {  // src/render.cpp:118-132
			r = (r + 0.1f) * S.tint[0];
			gg = (gg + 0.1f) * S.tint[1];
			b = (b + 0.1f) * S.tint[2];
		}
		const unsigned cb = conv_channel(b), cg = conv_channel(gg), cr = conv_channel(r);
		if(!SUPER) {
			// stage the packet's 16 x 48 bytes; store_tile() writes them as 16-byte vectors: one store per 5.3 pixels
			// instead of three 1-byte stores per pixel (what matters when `out` is rank 0's frame behind NVLink)
			unsigned char *t = tile + ty * 48 + tx * 3;
			t[0] = (unsigned char)cb;
			t[1] = (unsigned char)cg;
			t[2] = (unsigned char)cr;
		}
		else if(store && ox < cw && oy < ch) {
			unsigned char *dst = out + (size_t)(py0 + oy) * pitch + (size_t)(px0 + ox) * 3;
			dst[0] = (unsigned char)cb;
			dst[1] = (unsigned char)cg;
			dst[2] = (unsigned char)cr;
		}
		__syncwarp();
	}
	if(DEFER && S.cullMask)
		cull_resolve(S, pc);
}

// The staged packet -> frame, 3 x 16 bytes per packet row (after a CTA barrier that follows render_packet).
__device__ __forceinline__ void store_tile(const unsigned char *tile, int px0, int py0, int cw, int ch,
										   unsigned char *__restrict__ out, size_t pitch) {
	if(threadIdx.x < 48) {
		const int row = threadIdx.x / 3, part = threadIdx.x - row * 3;
		if(row < ch) {
			unsigned char *dst = out + (size_t)(py0 + row) * pitch + (size_t)px0 * 3 + part * 16;
			const unsigned char *src = tile + row * 48 + part * 16;
			if(cw == 16 && !((size_t)dst & 15))
				*reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<const uint4 *>(src);
			else
				for(int i = 0; i < 16 && part * 16 + i < cw * 3; i++)
					dst[i] = src[i];
		}
	}
}

template <int FLAGS, bool SUPER, int MINB>
__global__ void __launch_bounds__(kPacket, MINB) k_render(SceneDev S, const PreSet pre, RayGenParams g, int W, int H,
												 int packetsX, const int4 *__restrict__ packets,
												 unsigned char *__restrict__ out, size_t pitch, int firstPacket,
												 unsigned *__restrict__ costOut) {
	__shared__ __align__(16) unsigned char tile[SUPER ? 16 : 16 * 48];
	__shared__ unsigned sStart;  // (shared, not a register: nothing extra stays live across the packet)
	const int packet = blockIdx.x + firstPacket;
	if(costOut && threadIdx.x == 0)
		sStart = (unsigned)clock();
	PacketCounts pc = {};
	TravCount tc = {0, 0, 0};
	if(S.hasUber)
		packet_vote_init();
	int px0, py0, cw, ch;
	if(packets) {
		const int4 p = packets[packet];
		px0 = p.x;
		py0 = p.y;
		cw = p.z;
		ch = p.w;
	}
	else {
		px0 = (packet % packetsX) * 16;
		py0 = (packet / packetsX) * 16;
		cw = min(16, W - px0);
		ch = min(16, H - py0);
	}
	render_packet<FLAGS, SUPER>(S, pre, g, px0, py0, cw, ch, out, pitch, tile, pc, tc);
	flush_counts(S, pc, tc);  // its barriers also publish `tile`
	if(!SUPER)
		store_tile(tile, px0, py0, cw, ch, out, pitch);
	// cycles this packet kept its CTA slot (snl_render_multi orders next frame's packets by it, longest first)
	if(costOut && threadIdx.x == 0)
		costOut[packet] = max((unsigned)clock() - sStart, 1u);
}

// ---- launch wrappers ------------------------------------------------------------------------------------------

static inline unsigned blocks_for(int n, int per) { return (unsigned)((n + per - 1) / per); }

cudaError_t launch_pretransform(const GenTri *gtris, int nTris, const float origin[3], PreTri *out, cudaStream_t st) {
	if(nTris <= 0)
		return cudaSuccess;
	k_pretransform<<<blocks_for(nTris, 256), 256, 0, st>>>(gtris, nTris, origin[0], origin[1], origin[2], out);
	return cudaGetLastError();
}

cudaError_t launch_primary(const SceneDev &S, const PreSet &pre, const RayGenParams &g, int W, int H, int *triId, float *t,
						   float *bary, int *instId, cudaStream_t st) {
	const int px = (W + 15) / 16, py = (H + 15) / 16;
	if(S.instances)
		k_primary<true><<<px * py, kPacket, 0, st>>>(S, pre, g, W, H, px, triId, t, bary, instId);
	else
		k_primary<false><<<px * py, kPacket, 0, st>>>(S, pre, g, W, H, px, triId, t, bary, instId);
	return cudaGetLastError();
}

cudaError_t launch_trace_rays(const SceneDev &S, const PreSet &pre, const float *o3, const float *d3,
							  const unsigned char *mask, int n, bool shared, int *triId, float *t, float *bary,
							  cudaStream_t st) {
	if(n <= 0)
		return cudaSuccess;
	if(shared)
		k_trace_rays<true><<<blocks_for(n, kPacket), kPacket, 0, st>>>(S, pre, o3, d3, mask, n, triId, t, bary);
	else
		k_trace_rays<false><<<blocks_for(n, kPacket), kPacket, 0, st>>>(S, pre, o3, d3, mask, n, triId, t, bary);
	return cudaGetLastError();
}

cudaError_t launch_trace_shadow(const SceneDev &S, const PreSet &pre, const float origin[3], const float *d3, float *tmax,
								int n, cudaStream_t st) {
	if(n <= 0)
		return cudaSuccess;
	k_trace_shadow<<<blocks_for(n, kPacket), kPacket, 0, st>>>(S, pre, origin[0], origin[1], origin[2], d3, tmax, n);
	return cudaGetLastError();
}

cudaError_t launch_raytrace(const SceneDev &S, const PreSet &pre, const float *o3, const float *d3,
							const unsigned char *mask, int n, bool shared, float *rgb3, cudaStream_t st) {
	if(n <= 0)
		return cudaSuccess;
	const unsigned grid = blocks_for(n, kPacket);
	if(S.hasTransp) {
		if(shared)
			k_raytrace<true, true><<<grid, kPacket, 0, st>>>(S, pre, o3, d3, mask, n, rgb3);
		else
			k_raytrace<false, true><<<grid, kPacket, 0, st>>>(S, pre, o3, d3, mask, n, rgb3);
	}
	else if(shared)
		k_raytrace<true, false><<<grid, kPacket, 0, st>>>(S, pre, o3, d3, mask, n, rgb3);
	else
		k_raytrace<false, false><<<grid, kPacket, 0, st>>>(S, pre, o3, d3, mask, n, rgb3);
	return cudaGetLastError();
}

cudaError_t launch_render(const SceneDev &S, const PreSet &pre, const RayGenParams &g, int W, int H, const int4 *packets,
						  int nPackets, unsigned char *out, size_t pitch, int variant, cudaStream_t st, int first, int count,
						  unsigned *costOut) {
	const int px = (W + 15) / 16, py = (H + 15) / 16;
	const int all = packets ? nPackets : px * py;
	const int grid = count < 0 ? all - first : count;
	if(grid <= 0 || first < 0 || first + grid > all)
		return cudaSuccess;
	const int MB = SNL_RENDER_MIN_BLOCKS;
#define SNL_LAUNCH_RENDER(F, SUP, B) k_render<F, SUP, B><<<grid, kPacket, 0, st>>>(S, pre, g, W, H, px, packets, out, pitch, first, costOut)
	if(S.instances)  // Scene<DBVH>: simple shading only, so no transparency; supersampling is refused by the caller
		SNL_LAUNCH_RENDER(2, false, MB);
	else if(S.hasTransp) {
		if(S.gvSuper)
			SNL_LAUNCH_RENDER(1, true, MB);
		else
			SNL_LAUNCH_RENDER(1, false, MB);
	}
	else if(S.gvSuper)
		SNL_LAUNCH_RENDER(0, true, MB);
	else if(S.reflDepth2)
		SNL_LAUNCH_RENDER(8, false, MB);
	else if(variant == 1)
		SNL_LAUNCH_RENDER(0, false, MB + 1);
	else
		SNL_LAUNCH_RENDER(0, false, MB);
#undef SNL_LAUNCH_RENDER
	return cudaGetLastError();
}

// ---- multi-GPU frame protocol (CommCtl in rank 0's memory, see snl_comm.cu) ------------------------------------
__global__ void k_frame_open(CommCtl *ctl, unsigned seq, int world) {
	const unsigned slot = seq & 1u;
	// the slot was last used by frame seq - 2: nobody may still be fetching from / reporting to it
	if((int)threadIdx.x < world && seq >= 2)
		wait_sys<true>(&ctl->rankDone[threadIdx.x], seq - 1u, &ctl->error);
	__syncwarp();
	if(threadIdx.x == 0)
		ctl->statsArrived[slot] = 0u;
	if(threadIdx.x < kCounters)
		ctl->stats[slot][threadIdx.x] = 0ull;
	if(threadIdx.x < kCommBands)
		ctl->bandDone[slot][threadIdx.x] = 0u;
	__syncwarp();
	if(threadIdx.x == 0) {
		__threadfence_system();
		st_release_sys(&ctl->ready[slot], seq + 1u);
	}
}

__global__ void k_frame_publish(CommCtl *ctl, const Counters *local, unsigned seq, int rank) {
	const unsigned slot = seq & 1u;
	if(threadIdx.x < kCounters) {
		const unsigned long long v = reinterpret_cast<const unsigned long long *>(local)[threadIdx.x];
		if(v)
			atomicAdd_system(&ctl->stats[slot][threadIdx.x], v);
	}
	__syncwarp();
	if(threadIdx.x == 0) {
		__threadfence_system();
		atomicAdd_system(&ctl->statsArrived[slot], 1u);
		st_release_sys(&ctl->rankDone[rank], seq + 1u);
	}
}

__global__ void k_wait_counter(const unsigned *counter, unsigned expected, unsigned *error) {
	wait_sys<true>(counter, expected, error);
}

// One CTA per row band of a rank's packet list: the band's packets in descending order of the cycles they took in the
// frame before (longest first = the tail of the band is made of its cheapest packets). Bitonic sort in shared memory;
// bands of more than 4096 packets are passed through unchanged.
__global__ void __launch_bounds__(1024) k_sort_band_packets(const int4 *__restrict__ in, const unsigned *__restrict__ cost,
															 int4 *__restrict__ out, const int *__restrict__ bandFirst) {
	__shared__ unsigned key[4096];
	__shared__ unsigned short idx[4096];
	const int first = bandFirst[blockIdx.x], n = bandFirst[blockIdx.x + 1] - first;
	if(n > 4096 || n <= 1) {
		for(int i = threadIdx.x; i < n; i += blockDim.x)
			out[first + i] = in[first + i];
		return;
	}
	int m = 2;
	while(m < n)
		m <<= 1;
	for(int i = threadIdx.x; i < m; i += blockDim.x) {
		key[i] = i < n ? cost[first + i] : 0u;  // recorded costs are >= 1: the padding sorts behind every packet
		idx[i] = (unsigned short)i;
	}
	__syncthreads();
	for(int k = 2; k <= m; k <<= 1)
		for(int j = k >> 1; j > 0; j >>= 1) {
			for(int i = threadIdx.x; i < m; i += blockDim.x) {
				const int l = i ^ j;
				if(l > i) {
					const bool desc = (i & k) == 0;
					const unsigned a = key[i], b = key[l];
					if(desc ? a < b : a > b) {
						key[i] = b;
						key[l] = a;
						const unsigned short t = idx[i];
						idx[i] = idx[l];
						idx[l] = t;
					}
				}
			}
			__syncthreads();
		}
	for(int i = threadIdx.x; i < n; i += blockDim.x)
		out[first + i] = in[first + idx[i]];
}

// Runs in-stream behind a rank's k_render of one row band: the kernel boundary has completed (and made visible) every
// pixel store of that launch, peer stores included, so one system-scope increment tells rank 0 the band share is in.
__global__ void k_band_signal(unsigned *counter) { atomicAdd_system(counter, 1u); }

cudaError_t launch_frame_open(CommCtl *ctl, unsigned seq, int world, cudaStream_t st) {
	k_frame_open<<<1, 32, 0, st>>>(ctl, seq, world);
	return cudaGetLastError();
}
cudaError_t launch_frame_publish(CommCtl *ctl, const Counters *local, unsigned seq, int rank, cudaStream_t st) {
	k_frame_publish<<<1, 32, 0, st>>>(ctl, local, seq, rank);
	return cudaGetLastError();
}
cudaError_t launch_sort_band_packets(const int4 *in, const unsigned *cost, int4 *out, const int *bandFirst, int nBands,
									 cudaStream_t st) {
	k_sort_band_packets<<<nBands, 1024, 0, st>>>(in, cost, out, bandFirst);
	return cudaGetLastError();
}
cudaError_t launch_band_signal(unsigned *counter, cudaStream_t st) {
	k_band_signal<<<1, 1, 0, st>>>(counter);
	return cudaGetLastError();
}
cudaError_t launch_wait_counter(const unsigned *counter, unsigned expected, unsigned *error, cudaStream_t st) {
	k_wait_counter<<<1, 1, 0, st>>>(counter, expected, error);
	return cudaGetLastError();
}

// Farm tile format of Render(scene, camera, resx, resy, data, coords, offsets, ...) (src/render.cpp:140-166, compress == 1):
// every tile is three w*h byte planes, colour.x, colour.y - colour.x, colour.z - colour.x (mod 256). The frame holds
// ConvColor's byte order (z, y, x per pixel, src/render.cpp:11-17), so this is a pure byte shuffle of bytes that are
// already the reference's. One CTA per tile; 4 pixels (12 frame bytes, three 32-bit words) per thread when aligned.
__global__ void __launch_bounds__(256) k_farm_pack(const unsigned char *__restrict__ frame, size_t pitch, const int4 *__restrict__ tiles,
													const int *__restrict__ offsets, unsigned char *__restrict__ out) {
	const int4 t = tiles[blockIdx.x];
	const int x = t.x, y = t.y, w = t.z, h = t.w;
	unsigned char *dst = out + offsets[blockIdx.x];
	const size_t plane = (size_t)w * h;
	const bool fast = !(w & 3) && !(x & 3) && !(pitch & 3) && !(offsets[blockIdx.x] & 3) && !((size_t)frame & 3) && !((size_t)out & 3);
	if(fast) {
		const int wq = w >> 2;
		for(int i = threadIdx.x; i < wq * h; i += blockDim.x) {
			const int ty = i / wq, tx = (i - ty * wq) * 4;
			const uint32_t *p = (const uint32_t *)(frame + (size_t)(y + ty) * pitch + (size_t)(x + tx) * 3);
			const uint32_t a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
			// bytes: z0 y0 x0 z1 | y1 x1 z2 y2 | x2 z3 y3 x3
			const uint32_t X = ((a >> 16) & 255u) | (b & 0xff00u) | ((c & 255u) << 16) | (c & 0xff000000u);
			const uint32_t Y = ((a >> 8) & 255u) | ((b & 255u) << 8) | ((b >> 24) << 16) | (((c >> 16) & 255u) << 24);
			const uint32_t Z = (a & 255u) | ((a >> 24) << 8) | (((b >> 16) & 255u) << 16) | (((c >> 8) & 255u) << 24);
			const uint32_t o = ty * w + tx;
			*(uint32_t *)(dst + o) = X;
			*(uint32_t *)(dst + plane + o) = __vsub4(Y, X);
			*(uint32_t *)(dst + 2 * plane + o) = __vsub4(Z, X);
		}
	}
	else
		for(int i = threadIdx.x; i < w * h; i += blockDim.x) {
			const int ty = i / w, tx = i - ty * w;
			const unsigned char *p = frame + (size_t)(y + ty) * pitch + (size_t)(x + tx) * 3;
			const unsigned char X = p[2];
			dst[i] = X;
			dst[plane + i] = (unsigned char)(p[1] - X);
			dst[2 * plane + i] = (unsigned char)(p[0] - X);
		}
}

cudaError_t launch_farm_pack(const unsigned char *frame, size_t pitch, const int4 *tiles, const int *offsets, int nTiles,
							 unsigned char *out, cudaStream_t st) {
	if(nTiles <= 0)
		return cudaSuccess;
	k_farm_pack<<<nTiles, 256, 0, st>>>(frame, pitch, tiles, offsets, out);
	return cudaGetLastError();
}

}  // namespace snl
