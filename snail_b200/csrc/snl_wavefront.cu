// snail_b200 — wavefront form of Render() for frames with the reflection bounce (gVals[7]).
//
// The fused k_render traces the incoherent reflection rays and their shadow rays with the static thread <-> pixel
// mapping of the packet; ncu shows those two phases are 58 % of the frame's issue slots at 14-17 active lanes of 32.
// Here they are pulled out into queue-driven PERSISTENT traversal kernels in which a lane fetches the next ray the
// moment its own is finished (per-warp chunks of a global cursor), so warps stay full; shading keeps the static
// packet mapping because Scene::RayTrace's results depend on the 4-quad block structure (src/scene_trace.cpp:146).
//
//   k_wf_primary   per packet : ray gen, primary walk, Sample + material, level-0 lights (inline, coherent),
//                               emits the reflection ray + queue entry, stores level-0 state
//   k_wf_trace<0>  persistent : nearest hit of the queued reflection rays (per-ray origin)           -> hit1
//   k_wf_shade     per packet : Scene::RayTrace<0,hasMask> shading of the reflection hits, emits their shadow rays
//   k_wf_trace<1>  persistent : any-hit of the queued shadow rays (all lights)                        -> occl bits
//   k_wf_resolve   per packet : level-1 lights, reflection blend (scene_trace.cpp:454-466), compose, ConvColor, store
//
// Arithmetic is the same device functions as the fused kernel (snl_device.cuh), so the frame is bit-identical to it.
// this file keeps its own register-cached stack top
#define SNL_STEP_TOS
#include "snl_device.cuh"
#include "snl_kernels.h"

namespace snl {

__device__ __forceinline__ void st4(Float4 *p, float x, float y, float z, float w) {
	*reinterpret_cast<float4 *>(p) = make_float4(x, y, z, w);
}
__device__ __forceinline__ float4 ld4(const Float4 *p) { return *reinterpret_cast<const float4 *>(p); }

// Appends the calling lanes with `want` set to a queue: one atomicAdd per CTA, entries keep lane order (coherent).
__device__ __forceinline__ void enqueue(bool want, unsigned value, unsigned *__restrict__ queue, unsigned *counter) {
	__shared__ unsigned warpCount[kPacket / 32];
	__shared__ unsigned ctaBase;
	const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const unsigned m = __ballot_sync(SNL_FULL, want);
	if(lane == 0)
		warpCount[warp] = __popc(m);
	__syncthreads();
	if(threadIdx.x == 0) {
		unsigned total = 0;
		for(int w = 0; w < kPacket / 32; w++) {
			const unsigned c = warpCount[w];
			warpCount[w] = total;
			total += c;
		}
		ctaBase = total ? atomicAdd(counter, total) : 0;
	}
	__syncthreads();
	if(want)
		queue[ctaBase + warpCount[warp] + __popc(m & ((1u << lane) - 1))] = value;
	__syncthreads();  // warpCount is reused by the next call
}

__device__ __forceinline__ void packet_rect(const int4 *__restrict__ packets, int packetsX, int W, int H, int &px0, int &py0,
											int &cw, int &ch) {
	if(packets) {
		const int4 p = packets[blockIdx.x];
		px0 = p.x;
		py0 = p.y;
		cw = p.z;
		ch = p.w;
	}
	else {
		px0 = (blockIdx.x % packetsX) * 16;
		py0 = (blockIdx.x / packetsX) * 16;
		cw = min(16, W - px0);
		ch = min(16, H - py0);
	}
}

// ---- stage 1 ------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kPacket) k_wf_primary(SceneDev S, const PreSet pre, RayGenParams g, int W, int H,
													 int packetsX, const int4 *__restrict__ packets, WfBuffers B) {
	int px0, py0, cw, ch;
	packet_rect(packets, packetsX, W, H, px0, py0, cw, ch);
	const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
	const unsigned id = blockIdx.x * kPacket + threadIdx.x;
	float dx, dy, dz;
	gen_primary(g, px0, py0, tx, ty, dx, dy, dz);
	PacketCounts pc = {};
	TravCount tc = {0, 0, 0};

	Hit h;
	h.dist = CUDART_INF_F;
	h.obj = 0;
	h.bu = h.bv = 0.0f;
	traverse_nearest<true>(S.nodes, pre.tris, g.origin[0], g.origin[1], g.origin[2], dx, dy, dz, safe_inv(dx), safe_inv(dy),
						   safe_inv(dz), h, tc);
	pc.primary = 1;
	__syncwarp();
	const bool sel = h.dist < CUDART_INF_F;
	Sample s;
	bool trans;  // transparency and supersampling frames always take the fused kernel
	shade_block<false>(S, sel, h, g.origin[0], g.origin[1], g.origin[2], dx, dy, dz, false, s, trans);

	// the bounce: ray + queue entry (traced by k_wf_trace<0>)
	float rx, ry, rz, rox, roy, roz;
	reflection_ray(s, dx, dy, dz, rx, ry, rz, rox, roy, roz);
	st4(B.rayO + id, rox, roy, roz, sel ? 1.0f : 0.0f);
	st4(B.rayD + id, rx, ry, rz, 0.0f);
	enqueue(sel, id, B.rq, B.qcount + 0);
	pc.reflection = sel;

	// level-0 lights: coherent shadow rays from one origin, traced inline
	float ldr = S.ambient[0], ldg = S.ambient[1], ldb = S.ambient[2];
	float lsr = 0.0f, lsg = 0.0f, lsb = 0.0f;
	for(int li = 0; li < S.nLights; li++) {
		if(trace_light(S, li, sel, s, ldr, ldg, ldb, lsr, lsg, lsb, tc))
			pc.shadow++;
		__syncwarp();
	}
	st4(B.st0 + 0 * B.NL + id, s.dr, s.dg, s.db, 0.0f);
	st4(B.st0 + 1 * B.NL + id, s.sr, s.sg, s.sb, 0.0f);
	st4(B.st0 + 2 * B.NL + id, ldr, ldg, ldb, 0.0f);
	st4(B.st0 + 3 * B.NL + id, lsr, lsg, lsb, 0.0f);
	__syncwarp();
	flush_counts(S, pc, tc);
}

// ---- stages 2 and 4: queue-driven persistent traversal -----------------------------------------------------------------
// Every lane owns one ray at a time. At the top of the loop the warp hands new rays to the lanes that finished theirs:
// rays come from a warp-private chunk [wnext, wend) of the queue, refilled with one atomicAdd on the global cursor per
// kChunk rays. Then all lanes with a ray run one round of the while-while walk (inner nodes until every lane has parked
// a leaf or run dry, then the parked leaves). SHADOW = any-hit from a light origin (the one node array + PreTri of that
// light), else nearest hit with a per-ray origin (same nodes + GenTri).
enum { kChunk = 64 };

#ifndef SNL_WF_MIN_BLOCKS
#define SNL_WF_MIN_BLOCKS 4
#endif
template <bool SHADOW>
__global__ void __launch_bounds__(256, SNL_WF_MIN_BLOCKS) k_wf_trace(SceneDev S, WfBuffers B) {
	const unsigned n = B.qcount[SHADOW ? 1 : 0];
	unsigned *cursor = B.qcount + (SHADOW ? 3 : 2);
	const unsigned *__restrict__ queue = SHADOW ? B.sq : B.rq;
	const unsigned lane = threadIdx.x & 31;
	unsigned wnext = 0, wend = 0;
	bool term = false, idle = true;
	TravCount tc = {0, 0, 0};

	unsigned id = 0;
	float ox = 0.0f, oy = 0.0f, oz = 0.0f, dx = 0.0f, dy = 0.0f, dz = 0.0f, ix = 0.0f, iy = 0.0f, iz = 0.0f;
	Hit h;
	h.dist = 0.0f;
	h.obj = 0;
	h.bu = h.bv = 0.0f;
	const GpuNode *__restrict__ nodes = S.nodes;
	const PreTri *__restrict__ ptris = nullptr;
	int stackMem[kStack + 1];
	int *const stack = stackMem + 1;
	int sp = 0, tos = kSentinel, cur = kSentinel;

	for(;;) {
		__syncwarp();
		const bool want = idle && !term;
		const unsigned m = __ballot_sync(SNL_FULL, want);
		if(m) {
			const unsigned rank = __popc(m & ((1u << lane) - 1)), need = __popc(m);
			const unsigned avail = wend - wnext;
			unsigned my = 0xffffffffu;
			if(want && rank < avail)
				my = wnext + rank;
			if(avail < need) {
				unsigned base = 0;
				if(lane == 0)
					base = atomicAdd(cursor, (unsigned)kChunk);
				base = __shfl_sync(SNL_FULL, base, 0);
				const unsigned end = base < n ? min(base + (unsigned)kChunk, n) : base;
				if(want && rank >= avail) {
					const unsigned r = base + (rank - avail);
					if(r < end)
						my = r;
					else
						term = true;
				}
				wnext = min(base + (need - avail), end);
				wend = end;
			}
			else
				wnext += need;
			if(my != 0xffffffffu) {
				id = __ldg(queue + my);
				if(SHADOW) {
					const unsigned li = id / B.NL;
					const float4 r = ld4(B.sdir + id);
					dx = r.x;
					dy = r.y;
					dz = r.z;
					h.dist = r.w;  // tmax
					const PreSet ps = B.lightSets[li];
					nodes = S.nodes;
					ox = S.lights[li].pos[0];  // the shadow rays start at the light (box_test needs the origin)
					oy = S.lights[li].pos[1];
					oz = S.lights[li].pos[2];
					ptris = ps.tris;
				}
				else {
					const float4 o = ld4(B.rayO + id), d = ld4(B.rayD + id);
					ox = o.x;
					oy = o.y;
					oz = o.z;
					dx = d.x;
					dy = d.y;
					dz = d.z;
					h.dist = CUDART_INF_F;
					h.obj = 0;
					h.bu = h.bv = 0.0f;
				}
				ix = safe_inv(dx);
				iy = safe_inv(dy);
				iz = safe_inv(dz);
				sp = 0;
				tos = kSentinel;
				cur = 0;
				idle = false;
			}
		}
		if(__all_sync(SNL_FULL, idle))
			break;  // nothing in flight and nothing left to fetch
		if(idle)
			continue;

		int leaf = 0;
		bool occluded = false;
		while((unsigned)cur < (unsigned)kSentinel) {
			SNL_COUNT(tc.nodes);
			bool h0, h1;
			float lmin0, lmin1;
			int l0, l1;
			// box_test takes origin * idir (loop-invariant: hoisted out of the walk)
			node_test(nodes + cur, ox * ix, oy * iy, oz * iz, ix, iy, iz, h.dist, h0, h1, lmin0, lmin1, l0, l1);
			SNL_TRAVERSE_STEP();
			if(!__any_sync(__activemask(), leaf == 0))
				break;
		}
		while(leaf < 0) {
			const int lf = ~leaf;
			const int first = lf & kLeafFirstMask, count = (lf >> kLeafFirstBits) + 1;
			for(int i = 0; i < count; i++) {
				if(SHADOW) {
					SNL_COUNT(tc.tris);
					const PreTri *t = ptris + first + i;
					const float4 tu = ldg4(&t->u), tv = ldg4(&t->v), tn = ldg4(&t->n);
					const float det = dx * tn.x + dy * tn.y + dz * tn.z;
					const float v = dx * tv.x + dy * tv.y + dz * tv.z;
					const float u = dx * tu.x + dy * tu.y + dz * tu.z;
					if(fminf(u, v) >= 0.0f && u + v <= det && tu.w > 0.0f && tu.w < h.dist * det)
						occluded = true;
				}
				else {
					SNL_COUNT(tc.gtris);
					tri_nearest_gen(S.gtris + first + i, first + i, ox, oy, oz, dx, dy, dz, h);
				}
			}
			if(SHADOW && occluded) {
				cur = kSentinel;
				break;
			}
			leaf = cur;
			if(cur < 0) {
				cur = tos;
				tos = stack[--sp];
			}
		}
		if(cur == kSentinel) {  // this ray is done
			if(SHADOW) {
				if(occluded)
					atomicOr(B.occl + id % B.NL, 1u << (id / B.NL));
			}
			else
				st4(B.hit1 + id, h.dist, __int_as_float(h.obj), h.bu, h.bv);
			idle = true;
		}
	}
	PacketCounts pc = {};
	__syncwarp();
	flush_counts(S, pc, tc);
}

// ---- stage 3: Scene::RayTrace<0, hasMask> shading of the reflection hits (level 1, no further bounce) ------------------
__global__ void __launch_bounds__(kPacket) k_wf_shade(SceneDev S, WfBuffers B) {
	const unsigned id = blockIdx.x * kPacket + threadIdx.x;
	const float4 o = ld4(B.rayO + id), d = ld4(B.rayD + id);
	const bool active = o.w != 0.0f;  // sel of level 0
	PacketCounts pc = {};
	TravCount tc = {0, 0, 0};
	Hit h;
	h.dist = CUDART_INF_F;
	h.obj = 0;
	h.bu = h.bv = 0.0f;
	if(active) {
		const float4 r = ld4(B.hit1 + id);
		h.dist = r.x;
		h.obj = __float_as_int(r.y);
		h.bu = r.z;
		h.bv = r.w;
	}
	// selector.All() of the level-0 packet decides RayGroup<0,0> vs <0,1> (src/scene_trace.cpp:615-617)
	const bool all = S.hasUber ? (bool)__syncthreads_and(active) : true;
	const bool sel = active && h.dist < CUDART_INF_F;
	Sample s;
	bool trans;
	shade_block<false>(S, sel, h, o.x, o.y, o.z, d.x, d.y, d.z, !all, s, trans);
	st4(B.st1 + 0 * B.NL + id, s.dr, s.dg, s.db, sel ? 1.0f : 0.0f);
	st4(B.st1 + 1 * B.NL + id, s.sr, s.sg, s.sb, 0.0f);
	B.occl[id] = 0;
	for(int li = 0; li < S.nLights; li++) {
		float fx = 0.0f, fy = 0.0f, fz = 0.0f, dot = 0.0f, distance = 0.0f;
		const bool cast = sel && light_setup(S.lights[li], s.px, s.py, s.pz, s.nx, s.ny, s.nz, fx, fy, fz, dot, distance);
		const unsigned e = li * B.NL + id;
		if(cast)
			st4(B.sdir + e, fx, fy, fz, distance * 0.9999f);
		B.sdot[e] = make_float2(cast ? dot : 0.0f, distance);
		enqueue(cast, e, B.sq, B.qcount + 1);
		pc.reflection_shadow += cast;
	}
	__syncwarp();
	flush_counts(S, pc, tc);
}

// ConvColor (src/render.cpp:11-17): trunc(clamp(c * 255, 0, 255)).
__device__ __forceinline__ unsigned wf_conv(float c) {
	float v = c * 255.0f;
	v = v > 0.0f ? v : 0.0f;
	v = v < 255.0f ? v : 255.0f;
	return (unsigned)int(v);
}

// ---- stage 5: level-1 lights, reflection blend, level-0 compose, pixel -------------------------------------------------
__global__ void __launch_bounds__(kPacket) k_wf_resolve(SceneDev S, WfBuffers B, int W, int H, int packetsX,
													 const int4 *__restrict__ packets, unsigned char *__restrict__ out,
													 size_t pitch) {
	int px0, py0, cw, ch;
	packet_rect(packets, packetsX, W, H, px0, py0, cw, ch);
	const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
	const unsigned id = blockIdx.x * kPacket + threadIdx.x;
	const bool sel0 = ld4(B.rayO + id).w != 0.0f;
	float4 d0 = ld4(B.st0 + 0 * B.NL + id);
	const float4 s0 = ld4(B.st0 + 1 * B.NL + id), ld0 = ld4(B.st0 + 2 * B.NL + id), ls0 = ld4(B.st0 + 3 * B.NL + id);
	if(sel0) {
		// colour of the reflection ray = RayTrace<0,*> output of level 1 (src/scene_trace.cpp:485-511)
		const float4 d1 = ld4(B.st1 + 0 * B.NL + id), s1 = ld4(B.st1 + 1 * B.NL + id);
		float cr = d1.x, cg = d1.y, cb = d1.z;
		if(S.nLights) {
			float ldr = S.ambient[0], ldg = S.ambient[1], ldb = S.ambient[2];
			float lsr = 0.0f, lsg = 0.0f, lsb = 0.0f;
			const unsigned occl = B.occl[id];
			for(int li = 0; li < S.nLights; li++) {
				const float2 dd = B.sdot[li * B.NL + id];
				// cast (N.L > 0), not occluded, tmax > 0
				if(dd.x > 0.0f && !((occl >> li) & 1) && dd.y * 0.9999f > 0.0f)
					light_accumulate(S.lights[li], dd.x, light_attenuation(S.lights[li], dd.y), ldr, ldg, ldb, lsr, lsg, lsb);
			}
			cr = d1.x * ldr + s1.x * lsr;
			cg = d1.y * ldg + s1.y * lsg;
			cb = d1.z * ldb + s1.z * lsb;
		}
		d0.x = d0.x + (cr - d0.x) * 0.3f;  // :454-466
		d0.y = d0.y + (cg - d0.y) * 0.3f;
		d0.z = d0.z + (cb - d0.z) * 0.3f;
	}
	float r = d0.x, gg = d0.y, b = d0.z;
	if(S.nLights) {
		r = d0.x * ld0.x + s0.x * ls0.x;
		gg = d0.y * ld0.y + s0.y * ls0.y;
		b = d0.z * ld0.z + s0.z * ls0.z;
	}
	if(S.gvTint) {  // src/render.cpp:118-132
		r = (r + 0.1f) * S.tint[0];
		gg = (gg + 0.1f) * S.tint[1];
		b = (b + 0.1f) * S.tint[2];
	}
	if(tx < cw && ty < ch) {
		unsigned char *dst = out + (size_t)(py0 + ty) * pitch + (size_t)(px0 + tx) * 3;
		dst[0] = (unsigned char)wf_conv(b);
		dst[1] = (unsigned char)wf_conv(gg);
		dst[2] = (unsigned char)wf_conv(r);
	}
}

// ---- launch ------------------------------------------------------------------------------------------------------------
cudaError_t launch_render_wavefront(const SceneDev &S, const PreSet &pre, const RayGenParams &g, int W, int H,
									const int4 *packets, int nPackets, const WfBuffers &B, int persistentCtas,
									unsigned char *out, size_t pitch, cudaStream_t st, int *launches) {
	const int px = (W + 15) / 16, py = (H + 15) / 16;
	const int grid = packets ? nPackets : px * py;
	if(grid <= 0)
		return cudaSuccess;
	cudaError_t e = cudaMemsetAsync(B.qcount, 0, 4 * sizeof(unsigned), st);
	if(e != cudaSuccess)
		return e;
	k_wf_primary<<<grid, kPacket, 0, st>>>(S, pre, g, W, H, px, packets, B);
	static int occ[2] = {0, 0};  // resident CTAs per SM of the two persistent kernels
	if(!occ[0]) {
		cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[0], k_wf_trace<false>, 256, 0);
		cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[1], k_wf_trace<true>, 256, 0);
		occ[0] = occ[0] > 0 ? occ[0] : 4;
		occ[1] = occ[1] > 0 ? occ[1] : 4;
	}
	const int sms = persistentCtas;  // number of SMs
	k_wf_trace<false><<<sms * occ[0], 256, 0, st>>>(S, B);
	k_wf_shade<<<grid, kPacket, 0, st>>>(S, B);
	*launches = 4;
	if(S.nLights) {
		k_wf_trace<true><<<sms * occ[1], 256, 0, st>>>(S, B);
		*launches = 5;
	}
	k_wf_resolve<<<grid, kPacket, 0, st>>>(S, B, W, H, px, packets, out, pitch);
	return cudaGetLastError();
}

}  // namespace snl
