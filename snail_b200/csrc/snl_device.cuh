// snail_b200 — device code of the hot path (sm_100a). Compiled with -fmad=false: every expression below is
// written in the reference's operation order so that, with IEEE division / square root, results are bit-identical
// to the reference built on veclib's scalar back-end (veclib/scalar/f32.h:102-111: Inv = 1/x, RSqrt = 1/sqrt(x)).
//
// One thread = one ray. One CTA of 256 threads = one reference packet (64 quads, src/render.cpp:52-66): thread i is
// lane i&3 of quad i>>2; a half-warp is one 4-quad shading block (src/scene_trace.cpp:98,146).
#pragma once

#include <cuda_runtime.h>
#include <math_constants.h>

#include "snl_scene_dev.h"

namespace snl {

#define SNL_FULL 0xffffffffu

__device__ __forceinline__ float4 ldg4(const Float4 *p) { return __ldg(reinterpret_cast<const float4 *>(p)); }
__device__ __forceinline__ int4 ldg4(const Int4 *p) { return __ldg(reinterpret_cast<const int4 *>(p)); }

// 256-bit read-only load (LDG.E.ENL2.256.CONSTANT on sm_100a): `p` must be 32-byte aligned. Divergent lanes pay the
// L1 one wavefront per distinct line and load instruction, so a 64-byte record costs two of these instead of four
// 128-bit loads.
__device__ __forceinline__ void ldg8(const void *p, float4 &lo, float4 &hi) {
	asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
				 : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w)
				 : "l"(p));
}

#ifdef SNL_COUNTERS
#define SNL_COUNT(var) (++(var))
#else
#define SNL_COUNT(var) ((void)0)
#endif

struct TravCount {
	unsigned nodes, tris, gtris;  // inner nodes fetched, PreTri records tested, GenTri records tested
};

// BBox::Test, one ray vs one box (src/bounding_box.cpp:61-97): hit <=> lmax >= 0 && lmin <= min(lmax, dist), for a ray
// with its own origin. The reference evaluates idir * (plane - origin); here each plane costs ONE fused multiply-add
// plane * idir - origin * idir (oi* = origin * idir, once per ray). The slab distances differ from the reference's by
// an ulp or two, which can only flip the verdict for a ray grazing the box surface; which triangle is hit, and t / u /
// v, come from the triangle test alone and stay bit-identical (tests/test_gpu_golden.py).
__device__ __forceinline__ bool box_test(float minx, float maxx, float miny, float maxy, float minz, float maxz, float oix,
										 float oiy, float oiz, float ix, float iy, float iz, float dist, float &lminOut) {
	float l1 = fmaf(minx, ix, -oix), l2 = fmaf(maxx, ix, -oix);
	float lmin = fminf(l1, l2), lmax = fmaxf(l1, l2);
	l1 = fmaf(miny, iy, -oiy);
	l2 = fmaf(maxy, iy, -oiy);
	lmin = fmaxf(lmin, fminf(l1, l2));
	lmax = fminf(lmax, fmaxf(l1, l2));
	l1 = fmaf(minz, iz, -oiz);
	l2 = fmaf(maxz, iz, -oiz);
	lmin = fmaxf(lmin, fminf(l1, l2));
	lmax = fminf(lmax, fmaxf(l1, l2));
	lminOut = lmin;
	return lmax >= 0.0f && lmin <= fminf(lmax, dist);
}

struct Hit {
	float dist;
	int obj;
	float bu, bv;
	int inst;  // instanced scenes only: Context::object (index into DBVH::elements); obj is then Context::element
};

// Triangle::Collide(Context<1,*>&) per ray (src/triangle.cpp:3-63), sharedOrigin form with the per-origin
// precomputation already applied (PreTri).
__device__ __forceinline__ void tri_nearest_pre(const PreTri *__restrict__ t, int idx, float dx, float dy, float dz, Hit &h) {
	const float4 tu = ldg4(&t->u), tv = ldg4(&t->v), tn = ldg4(&t->n);
	const float det = dx * tn.x + dy * tn.y + dz * tn.z;
	const float v = dx * tv.x + dy * tv.y + dz * tv.z;
	const float u = dx * tu.x + dy * tu.y + dz * tu.z;
	const float duv = det - u - v;
	const float uvmin = fminf(u, fminf(v, duv));
	const float uvmax = fmaxf(u, fmaxf(v, duv));
	if(uvmax <= 0.0f || uvmin >= 0.0f) {
		const float idet = 1.0f / det;
		const float dist = idet * tu.w;
		if(dist < h.dist && dist > 0.0f) {
			h.dist = dist;
			h.obj = idx;
			h.bu = u * idet;
			h.bv = v * idet;
		}
	}
}

// Triangle::Collide(Context<0,*>&) per ray (src/triangle.cpp:29-37,41-61): per-ray origin.
// SHORDER: the sharedOrigin form's operation order (tvec0 / tvec1 scaled by it0 BEFORE the dot products, :13-18) on an
// origin that is not known per frame: a shared origin seen from inside an instance (ObjectInstance::CollidePrimary
// hands BVH::TraversePrimary a Context<1,*>, src/dbvh/tree.h:118-165).
template <bool SHORDER = false>
__device__ __forceinline__ void tri_nearest_gen(const GenTri *__restrict__ t, int idx, float ox, float oy, float oz,
												float dx, float dy, float dz, Hit &h) {
	float4 ta, tba, tca, tn;
	ldg8(&t->a, ta, tba);
	ldg8(&t->ca, tca, tn);
	const float det = dx * tn.x + dy * tn.y + dz * tn.z;
	const float tx = ox - ta.x, ty = oy - ta.y, tz = oz - ta.z;
	// tvec0 = ba ^ tvec, tvec1 = tvec ^ ca   (veclib/vec3.h:101-106)
	const float t0x = tba.y * tz - tba.z * ty, t0y = tba.z * tx - tba.x * tz, t0z = tba.x * ty - tba.y * tx;
	const float t1x = ty * tca.z - tz * tca.y, t1y = tz * tca.x - tx * tca.z, t1z = tx * tca.y - ty * tca.x;
	const float tmul = -(tx * tn.x + ty * tn.y + tz * tn.z);
	const float v = SHORDER ? dx * (t0x * ta.w) + dy * (t0y * ta.w) + dz * (t0z * ta.w) : (dx * t0x + dy * t0y + dz * t0z) * ta.w;
	const float u = SHORDER ? dx * (t1x * ta.w) + dy * (t1y * ta.w) + dz * (t1z * ta.w) : (dx * t1x + dy * t1y + dz * t1z) * ta.w;
	const float duv = det - u - v;
	const float uvmin = fminf(u, fminf(v, duv));
	const float uvmax = fmaxf(u, fmaxf(v, duv));
	if(uvmax <= 0.0f || uvmin >= 0.0f) {
		const float idet = 1.0f / det;
		const float dist = idet * tmul;
		if(dist < h.dist && dist > 0.0f) {
			h.dist = dist;
			h.obj = idx;
			h.bu = u * idet;
			h.bv = v * idet;
		}
	}
}

// ---- traversal ------------------------------------------------------------------------------------------------------
// Per-thread stack, "while-while" order (Aila & Laine): a lane walks inner nodes until it reaches a leaf, parks the
// leaf and keeps walking until every lane of the warp has one parked (or has run out of nodes); then the warp tests
// triangles together. Inner-node code and triangle code each run with most lanes active instead of interleaving.
// Children are visited nearest first (by slab entry distance instead of the packet's firstNode^sign[axis]); only
// exact-t ties can resolve differently from the reference's packet order.
//
// SHARED rays (one origin for the whole launch: camera or light) test PreTri records (k_pretransform), per-ray-origin
// rays GenTri records; both walk the ONE node array with the FMA slab test above.

enum { kSentinel = 0x7fffffff };

// One step of the walk after the two box tests of the node at `cur`: go to the nearer hit child (push the farther), or
// pop; a leaf reached while none is parked gets parked and the walk continues with the next stack entry. stack[-1]
// reads (sp == 0 when tos is the sentinel) cannot happen: the sentinel ends the walk before it would be popped twice.
#ifdef SNL_STEP_TOS
#define SNL_STACK_DECL()        \
	int stackMem[kStack + 1];   \
	int *const stack = stackMem + 1; \
	int sp = 0;                 \
	int tos = kSentinel
#define SNL_POP(dst)        \
	do {                    \
		dst = tos;          \
		tos = stack[--sp];  \
	} while(0)
#define SNL_TRAVERSE_STEP()                                   \
	do {                                                      \
		const bool both_ = h0 && h1, any_ = h0 || h1;         \
		const bool swap_ = lmin1 < lmin0;                     \
		int next_ = both_ ? (swap_ ? l1 : l0) : (h0 ? l0 : l1); \
		if(both_) {                                           \
			stack[sp++] = tos;                                \
			tos = swap_ ? l0 : l1;                            \
		}                                                     \
		if(!any_) {                                           \
			next_ = tos;                                      \
			tos = stack[--sp];                                \
		}                                                     \
		if(next_ < 0 && leaf == 0) {                          \
			leaf = next_;                                     \
			next_ = tos;                                      \
			tos = stack[--sp];                                \
		}                                                     \
		cur = next_;                                          \
	} while(0)
#else
// Default: plain stack in local memory, [0] = sentinel. Fewer instructions per step than the register-cached top
// (SNL_STEP_TOS, which the wavefront kernels keep), one more dependent LDL on a double pop: 0.5 % faster on B200.
// The stack top is a POINTER into local memory: a push / pop is STL / LDL + one add, without the index -> address LEA
// (+1.8 % on sponza_like against an index).
#define SNL_STACK_DECL()   \
	int stack[kStack + 1]; \
	stack[0] = kSentinel;  \
	int *sp = stack + 1
#define SNL_PUSH(v) *sp++ = (v)
#define SNL_POP(dst) dst = *--sp
#define SNL_STEP_ONEPOP
// One pop site per step: a leaf that gets parked and a step that hit nothing both take the next stack entry. (A popped
// entry that is itself a leaf ends the inner walk even when nothing is parked yet; the triangle loop takes it from cur.)
// +3.8 % sponza_like, +3.4 % feline on B200 against a step with two pop sites.
#define SNL_TRAVERSE_STEP()                                   \
	do {                                                      \
		const bool both_ = h0 && h1, any_ = h0 || h1;         \
		const bool swap_ = lmin1 < lmin0;                     \
		int next_ = both_ ? (swap_ ? l1 : l0) : (h0 ? l0 : l1); \
		if(both_)                                             \
			SNL_PUSH(swap_ ? l0 : l1);                        \
		bool pop_ = !any_;                                    \
		if(any_ && next_ < 0 && leaf == 0) {                  \
			leaf = next_;                                     \
			pop_ = true;                                      \
		}                                                     \
		if(pop_)                                              \
			SNL_POP(next_);                                   \
		cur = next_;                                          \
	} while(0)
#endif

// Both children of one inner node. o* = origin * idir.
__device__ __forceinline__ void node_test(const GpuNode *__restrict__ n, float ox, float oy, float oz, float ix, float iy,
										  float iz, float dist, bool &h0, bool &h1, float &lmin0, float &lmin1, int &l0, int &l1) {
	float4 b0, b1, bz, lk;
	ldg8(&n->b0, b0, b1);
	ldg8(&n->bz, bz, lk);
	l0 = __float_as_int(lk.x);
	l1 = __float_as_int(lk.y);
	h0 = box_test(b0.x, b0.y, b0.z, b0.w, bz.x, bz.y, ox, oy, oz, ix, iy, iz, dist, lmin0);
	h1 = box_test(b1.x, b1.y, b1.z, b1.w, bz.z, bz.w, ox, oy, oz, ix, iy, iz, dist, lmin1);
}

// BVH::TraversePrimaryN per ray (src/bvh/traverse.cpp:14-80).
template <bool SHARED, bool SHORDER = false>
__device__ __forceinline__ void traverse_nearest(const GpuNode *__restrict__ nodes, const void *__restrict__ tris,
												 float ox, float oy, float oz, float dx, float dy, float dz, float ix,
												 float iy, float iz, Hit &h, TravCount &tc) {
	SNL_STACK_DECL();
	int cur = 0;
	const float bx = ox * ix, by = oy * iy, bz = oz * iz;  // box_test's origin * idir
	while(cur != kSentinel) {
		int leaf = 0;  // parked leaf link (< 0), 0 = none
		while((unsigned)cur < (unsigned)kSentinel) {
			SNL_COUNT(tc.nodes);
			bool h0, h1;
			float lmin0, lmin1;
			int l0, l1;
			node_test(nodes + cur, bx, by, bz, ix, iy, iz, h.dist, h0, h1, lmin0, lmin1, l0, l1);
			SNL_TRAVERSE_STEP();
			if(!__any_sync(__activemask(), leaf == 0))
				break;
		}
#ifdef SNL_STEP_ONEPOP
		if(leaf == 0 && cur < 0) {
			leaf = cur;
			SNL_POP(cur);
		}
#endif
		while(leaf < 0) {
			const int lf = ~leaf;
			const int first = lf & kLeafFirstMask, count = (lf >> kLeafFirstBits) + 1;
			for(int i = 0; i < count; i++) {
				if(SHARED) {
					SNL_COUNT(tc.tris);
					tri_nearest_pre(reinterpret_cast<const PreTri *>(tris) + first + i, first + i, dx, dy, dz, h);
				}
				else {
					SNL_COUNT(tc.gtris);
					tri_nearest_gen<SHORDER>(reinterpret_cast<const GenTri *>(tris) + first + i, first + i, ox, oy, oz, dx, dy, dz, h);
				}
			}
			leaf = cur;
			if(cur < 0)
				SNL_POP(cur);
		}
	}
}

// BVH::TraverseShadow per ray (src/bvh/traverse.cpp:82-149) with Triangle::Collide(ShadowContext&)
// (src/triangle.cpp:65-102): any hit, ONE-SIDED (only triangles with dir.nrm >= u+v >= 0 occlude), no division.
// Origin = light position; triangles = that light's PreTri records, or (GEN, instanced scenes: the light seen from
// inside an instance) the GenTri records with the sharedOrigin precomputation (src/triangle.cpp:13-18) done on the fly.
template <bool GEN = false>
__device__ __forceinline__ bool traverse_shadow(const GpuNode *__restrict__ nodes, const void *__restrict__ tris, float ox,
												float oy, float oz, float dx, float dy, float dz, float ix, float iy, float iz,
												float dist, TravCount &tc) {
	const float bx = ox * ix, by = oy * iy, bz = oz * iz;  // box_test's origin * idir
	SNL_STACK_DECL();
	int cur = 0;
	bool occluded = false;
	while(cur != kSentinel) {
		int leaf = 0;
		while((unsigned)cur < (unsigned)kSentinel) {
			SNL_COUNT(tc.nodes);
			bool h0, h1;
			float lmin0, lmin1;
			int l0, l1;
			node_test(nodes + cur, bx, by, bz, ix, iy, iz, dist, h0, h1, lmin0, lmin1, l0, l1);
			SNL_TRAVERSE_STEP();
			if(!__any_sync(__activemask(), leaf == 0))
				break;
		}
#ifdef SNL_STEP_ONEPOP
		if(leaf == 0 && cur < 0) {
			leaf = cur;
			SNL_POP(cur);
		}
#endif
		while(leaf < 0) {
			const int lf = ~leaf;
			const int first = lf & kLeafFirstMask, count = (lf >> kLeafFirstBits) + 1;
			for(int i = 0; i < count; i++) {
				SNL_COUNT(tc.tris);
				float4 tu, tv, tn;
				if(GEN) {
					const GenTri *t = reinterpret_cast<const GenTri *>(tris) + first + i;
					float4 ta, tba, tca;
					ldg8(&t->a, ta, tba);
					ldg8(&t->ca, tca, tn);
					const float tx = ox - ta.x, ty = oy - ta.y, tz = oz - ta.z;
					tv = make_float4((tba.y * tz - tba.z * ty) * ta.w, (tba.z * tx - tba.x * tz) * ta.w,
									 (tba.x * ty - tba.y * tx) * ta.w, 0.0f);
					tu = make_float4((ty * tca.z - tz * tca.y) * ta.w, (tz * tca.x - tx * tca.z) * ta.w,
									 (tx * tca.y - ty * tca.x) * ta.w, -(tx * tn.x + ty * tn.y + tz * tn.z));
				}
				else {
					const PreTri *t = reinterpret_cast<const PreTri *>(tris) + first + i;
					tu = ldg4(&t->u);
					tv = ldg4(&t->v);
					tn = ldg4(&t->n);
				}
				const float det = dx * tn.x + dy * tn.y + dz * tn.z;
				const float v = dx * tv.x + dy * tv.y + dz * tv.z;
				const float u = dx * tu.x + dy * tu.y + dz * tu.z;
				const float tmul = tu.w;
				if(fminf(u, v) >= 0.0f && u + v <= det && tmul > 0.0f && tmul < dist * det)
					occluded = true;
			}
			if(occluded) {
				cur = kSentinel;
				break;
			}
			leaf = cur;
			if(cur < 0)
				SNL_POP(cur);
		}
	}
	return occluded;
}

// SafeInv (src/rtbase.h:117-119): Inv(v + 1e-8).
__device__ __forceinline__ float safe_inv(float v) { return 1.0f / (v + 0.00000001f); }

// ---- instanced two-level BVH (DBVH, src/dbvh/tree.h, src/dbvh/traverse.cpp) ----------------------------------------
// Top level: a tree over the instances in the same GpuNode layout (leaf link = a range of DBVH::elements). An instance
// is entered by transforming the ray into its space (ObjectInstance::ITransformVec / ITransformPoint,
// src/dbvh/tree.h:32-44; the rotation is orthonormal, so t is shared) and walking the ONE bottom-level BVH there with
// the shared hit record, exactly what CollidePrimary / CollideShadow do per packet.
enum { kTopStack = 66 };  // DBVH::maxDepth + 2

__device__ __forceinline__ void inst_ray(const DevInstance *__restrict__ I, float ox, float oy, float oz, float dx, float dy,
										 float dz, float &lox, float &loy, float &loz, float &ldx, float &ldy, float &ldz) {
	float4 r0, r1, r2, t;
	ldg8(&I->r0, r0, r1);
	ldg8(&I->r2, r2, t);
	ldx = dx * r0.x + dy * r1.x + dz * r2.x;  // ITransformVec
	ldy = dx * r0.y + dy * r1.y + dz * r2.y;
	ldz = dx * r0.z + dy * r1.z + dz * r2.z;
	const float px = ox - t.x, py = oy - t.y, pz = oz - t.z;  // ITransformPoint: p -= translation first
	lox = px * r0.x + py * r1.x + pz * r2.x;
	loy = px * r0.y + py * r1.y + pz * r2.y;
	loz = px * r0.z + py * r1.z + pz * r2.z;
}

// DBVH::TraversePrimary0 per ray. SHARED: the rays of the launch share their origin (the reference then uses the
// sharedOrigin triangle arithmetic inside the instances).
template <bool SHARED>
__device__ __forceinline__ void traverse_nearest_inst(const SceneDev &S, float ox, float oy, float oz, float dx, float dy,
													  float dz, float ix, float iy, float iz, Hit &h, TravCount &tc) {
	int stack[kTopStack];
	int sp = 0;
	stack[sp++] = kSentinel;
	int cur = 0;
	const float bx = ox * ix, by = oy * iy, bz = oz * iz;
	while(cur != kSentinel) {
		if(cur >= 0) {
			bool h0, h1;
			float lmin0, lmin1;
			int l0, l1;
			node_test(S.topNodes + cur, bx, by, bz, ix, iy, iz, h.dist, h0, h1, lmin0, lmin1, l0, l1);
			const bool swap = lmin1 < lmin0;
			if(h0 && h1) {
				stack[sp++] = swap ? l0 : l1;
				cur = swap ? l1 : l0;
			}
			else if(h0 || h1)
				cur = h0 ? l0 : l1;
			else
				cur = stack[--sp];
		}
		else {
			const int lf = ~cur;
			const int first = lf & kLeafFirstMask, count = (lf >> kLeafFirstBits) + 1;
			for(int i = 0; i < count; i++) {
				float lox, loy, loz, ldx, ldy, ldz;
				inst_ray(S.instances + first + i, ox, oy, oz, dx, dy, dz, lox, loy, loz, ldx, ldy, ldz);
				const float before = h.dist;
				traverse_nearest<false, SHARED>(S.nodes, S.gtris, lox, loy, loz, ldx, ldy, ldz, safe_inv(ldx), safe_inv(ldy),
												safe_inv(ldz), h, tc);
				if(h.dist < before)
					h.inst = first + i;
			}
			cur = stack[--sp];
		}
	}
}

// DBVH::TraverseShadow0 per ray: any hit inside any instance; the shadow rays of a light share their origin.
__device__ __forceinline__ bool traverse_shadow_inst(const SceneDev &S, float ox, float oy, float oz, float dx, float dy, float dz,
													 float ix, float iy, float iz, float dist, TravCount &tc) {
	int stack[kTopStack];
	int sp = 0;
	stack[sp++] = kSentinel;
	int cur = 0;
	const float bx = ox * ix, by = oy * iy, bz = oz * iz;
	while(cur != kSentinel) {
		if(cur >= 0) {
			bool h0, h1;
			float lmin0, lmin1;
			int l0, l1;
			node_test(S.topNodes + cur, bx, by, bz, ix, iy, iz, dist, h0, h1, lmin0, lmin1, l0, l1);
			if(h0 && h1) {
				stack[sp++] = l1;
				cur = l0;
			}
			else if(h0 || h1)
				cur = h0 ? l0 : l1;
			else
				cur = stack[--sp];
		}
		else {
			const int lf = ~cur;
			const int first = lf & kLeafFirstMask, count = (lf >> kLeafFirstBits) + 1;
			for(int i = 0; i < count; i++) {
				float lox, loy, loz, ldx, ldy, ldz;
				inst_ray(S.instances + first + i, ox, oy, oz, dx, dy, dz, lox, loy, loz, ldx, ldy, ldz);
				if(traverse_shadow<true>(S.nodes, S.gtris, lox, loy, loz, ldx, ldy, ldz, safe_inv(ldx), safe_inv(ldy),
										 safe_inv(ldz), dist, tc))
					return true;
			}
			cur = stack[--sp];
		}
	}
	return false;
}

// RayGenerator::Generate (src/ray_generator.cpp:23-47) for thread (tx, ty) of the 16x16 packet at (x0, y0).
__device__ __forceinline__ void gen_primary(const RayGenParams &g, int x0, int y0, int tx, int ty, float &dx, float &dy,
											float &dz) {
	const int k = tx & 3, j = tx >> 2;
	const float xoff = float(x0 + (k & 2));        // floatq xoff(x + 0, x + 0, x + 2, x + 2)
	const float yoff = float(y0 - (k >> 1));       // floatq yoff(y + 0, y + 0, y - 1, y - 1)
	const float tposx = float(4 * j) + xoff;
	const float tposy = float(ty) + yoff;
	const float tpx = g.tup[0] * tposy + g.txyz[0][k];
	const float tpy = g.tup[1] * tposy + g.txyz[1][k];
	const float tpz = g.tup[2] * tposy + g.txyz[2][k];
	const float px = g.tright[0] * tposx + tpx;
	const float py = g.tright[1] * tposx + tpy;
	const float pz = g.tright[2] * tposx + tpz;
	const float r = 1.0f / sqrtf(px * px + py * py + pz * pz);
	dx = px * r;
	dy = py * r;
	dz = pz * r;
}

struct Sample {
	float px, py, pz;  // position
	float nx, ny, nz;  // normal (not normalised when interpolated, src/scene_trace.cpp:204,220,249)
	float dr, dg, db;  // diffuse
	float sr, sg, sb;  // specular
	float opacity;     // set by Uber / Transparent materials only (src/shading.h:23)
};

__device__ __forceinline__ float quad_min(float v) {
	v = fminf(v, __shfl_xor_sync(SNL_FULL, v, 1));
	return fminf(v, __shfl_xor_sync(SNL_FULL, v, 2));
}
__device__ __forceinline__ float quad_max(float v) {
	v = fmaxf(v, __shfl_xor_sync(SNL_FULL, v, 1));
	return fmaxf(v, __shfl_xor_sync(SNL_FULL, v, 2));
}

// PointSampler::Sample (src/sampling/point_sampler.cpp:126-210): bilinear RGB8 fetch, vertical flip,
// mip picked from the quad's texDiff.
__device__ __forceinline__ void tex_sample(const DevTexture &T, float tu, float tv, float tdx, float tdy, float &r, float &g,
										   float &b) {
	// ClampTexCoord (point_sampler.cpp:43-51)
	float ux = tu - float(int(tu)), uy = tv - float(int(tv));
	ux = ux < 0.0f ? ux + 1.0f : ux;
	uy = uy < 0.0f ? uy + 1.0f : uy;
	const float posx = ux * T.wMul, posy = uy * T.hMul;

	unsigned mip = 0;
	{
		const float mn = fminf(tdx * T.wMul, tdy * T.hMul);
		unsigned pixels = unsigned(mn * 0.6f);
		while(pixels) {
			mip++;
			pixels >>= 1;
		}
		mip = min(mip, unsigned(T.mips - 1));
	}
	const unsigned char *data = T.mip[mip];
	const int pitch = T.pitch[mip];

	int x1 = int(posx), y1 = int(posy);
	int x2 = x1 + 1, y2 = y1 + 1;
	const float fx = posx - float(x1), fy = posy - float(y1);
	y1 = T.height - y1;
	y2 = T.height - y2;
	const int wm = (T.width - 1) >> mip, hm = (T.height - 1) >> mip;
	x1 = ((x1 >> mip) & wm) * 3;
	x2 = ((x2 >> mip) & wm) * 3;
	y1 = ((y1 >> mip) & hm) * pitch;
	y2 = ((y2 >> mip) & hm) * pitch;
	const unsigned char *p0 = data + x1 + y1, *p1 = data + x2 + y1, *p2 = data + x1 + y2, *p3 = data + x2 + y2;
	const float r0 = float(__ldg(p0 + 0)), g0 = float(__ldg(p0 + 1)), b0 = float(__ldg(p0 + 2));
	const float r1 = float(__ldg(p1 + 0)), g1 = float(__ldg(p1 + 1)), b1 = float(__ldg(p1 + 2));
	const float r2 = float(__ldg(p2 + 0)), g2 = float(__ldg(p2 + 1)), b2 = float(__ldg(p2 + 2));
	const float r3 = float(__ldg(p3 + 0)), g3 = float(__ldg(p3 + 1)), b3 = float(__ldg(p3 + 2));
	// Lerp(a, b, x) = a + (b - a) * x   (veclib/vecbase.h:80)
	const float ra = r0 + (r1 - r0) * fx, rb = r2 + (r3 - r2) * fx;
	const float ga = g0 + (g1 - g0) * fx, gb = g2 + (g3 - g2) * fx;
	const float ba = b0 + (b1 - b0) * fx, bb = b2 + (b3 - b2) * fx;
	const float k = 1.0f / 255.0f;
	r = (ra + (rb - ra) * fy) * k;
	g = (ga + (gb - ga) * fy) * k;
	b = (ba + (bb - ba) * fy) * k;
}

// mat->Shade (src/shading/*.h Shade_) for one lane. matId == -1 (~0): Scene::defaultMat = white SimpleMaterial<true>.
__device__ __forceinline__ void material_shade(const SceneDev &S, int matId, bool masked, float dx, float dy, float dz,
											   float tu, float tv, float tdx, float tdy, Sample &s) {
	const float dot = dx * s.nx + dy * s.ny + dz * s.nz;
	if(matId < 0) {
		const float a = fabsf(dot);
		s.dr = s.dg = s.db = a;  // Vec3q(1,1,1) * Abs(dir | normal)
		s.sr = s.sg = s.sb = a;
		return;
	}
	const DevMaterial &m = S.mats[matId];
	if(m.kind == 3 /* SNL_MAT_TRANSPARENT */) {
		// TransparentMaterial<1>::Shade_ (src/shading/transparent_material.h:15-41): colour texture * (dir | normal),
		// opacity = x channel of the second texture, both sampled with the same texCoord / texDiff
		float r, g, b, tr, tg, tb;
		tex_sample(S.texs[m.texture], tu, tv, tdx, tdy, r, g, b);
		tex_sample(S.texs[m.texture2], tu, tv, tdx, tdy, tr, tg, tb);
		s.dr = r * dot;
		s.dg = g * dot;
		s.db = b * dot;
		s.sr = s.dr;
		s.sg = s.dg;
		s.sb = s.db;
		s.opacity = tr;
	}
	else if(m.kind == 2 /* SNL_MAT_TEX */) {
		float r, g, b;
		tex_sample(S.texs[m.texture], tu, tv, tdx, tdy, r, g, b);
		s.dr = r * dot;
		s.dg = g * dot;
		s.db = b * dot;
		s.sr = s.dr;
		s.sg = s.dg;
		s.sb = s.db;
	}
	else {
		const float a = fabsf(dot);
		s.dr = m.color[0] * a;
		s.dg = m.color[1] * a;
		s.db = m.color[2] * a;
		if(m.kind == 1 /* SNL_MAT_UBER */)
			s.opacity = m.opacity;  // dissFactor, src/shading/uber_material.h:23,28
		if(m.kind == 1 /* SNL_MAT_UBER */ && masked) {
			// UberMaterial::Shade_ hasMask branch stores the material's specular colour, the other branch the
			// diffuse term (src/shading/uber_material.h:21-29)
			s.sr = m.specular[0];
			s.sg = m.specular[1];
			s.sb = m.specular[2];
		}
		else {
			s.sr = s.dr;
			s.sg = s.dg;
			s.sb = s.db;
		}
	}
}

// Sample construction + material evaluation of Scene::RayTrace (src/scene_trace.cpp:139-452) for the calling lane;
// the 16 lanes of a half-warp are one 4-quad block. Must be called by all 32 lanes of the warp.
// TRANSP: also evaluates transSel for the lane (`trans`): the material has fTransparency and, in a block shaded per
// material, it is the LAST such material in order of first appearance - `transSel[b4 + q] = mask[q]` overwrites the
// whole block for every transparent material in turn (src/scene_trace.cpp:190,306,347-349).
// INST: instanced scene. DBVH::HasShadingData() is 0 (src/dbvh/tree.h:204-206), so shading is always the simple path,
// with the geometric normal rotated into world space (ObjectInstance::GetNormal, :171-174).
template <bool TRANSP, bool INST = false>
__device__ __forceinline__ void shade_block(const SceneDev &S, bool sel, const Hit &h, float ox, float oy, float oz, float dx,
											float dy, float dz, bool packetHasMask, Sample &s, bool &trans) {
	const int lane = threadIdx.x & 31;
	const int half = lane >> 4;
	const int obj = sel ? h.obj : 0;

	s.px = dx * h.dist + ox;
	s.py = dy * h.dist + oy;
	s.pz = dz * h.dist + oz;
	s.dr = s.dg = s.db = s.sr = s.sg = s.sb = 0.0f;
	s.nx = s.ny = s.nz = 0.0f;
	s.opacity = 1.0f;
	trans = false;

	const unsigned selB = (__ballot_sync(SNL_FULL, sel) >> (half * 16)) & 0xffffu;
	const int obj0 = __shfl_sync(SNL_FULL, obj, half * 16);
	const unsigned sameB = (__ballot_sync(SNL_FULL, obj == obj0) >> (half * 16)) & 0xffffu;
	const bool blockFull = selB == 0xffffu;
	const bool fast = blockFull && sameB == 0xffffu;

	if(!INST && S.gvFull && S.shtris) {
		int matId = -1;
		float tu = 0.0f, tv = 0.0f;
		bool texCoords = false;
		float4 w0, w1, w2, w3;
		if(sel) {
			const float4 *sp = reinterpret_cast<const float4 *>(S.shtris + obj);
			ldg8(sp, w0, w1);      // uv0.x uv0.y uv1.x uv1.y | uv2.x uv2.y n0.x n0.y
			ldg8(sp + 2, w2, w3);  // n0.z n1.x n1.y n1.z     | n2.x n2.y n2.z mat
			const int smat = __float_as_int(w3.w);
			const int mi = smat & 0x7fffffff;
			matId = mi < S.nTriMat ? S.triMat[mi] : -1;
			if(matId >= S.nMats)
				matId = -1;
			const bool flat = smat < 0;
			texCoords = matId >= 0 && (S.mats[matId].flags & 1);
			const float bx = h.bu, by = h.bv;
			if(fast && !texCoords) {
				if(flat) {
					s.nx = w1.z;
					s.ny = w1.w;
					s.nz = w2.x;
				}
				else {  // nrm0 + (nrm1 * bar.x + nrm2 * bar.y)   src/scene_trace.cpp:204
					s.nx = w1.z + (w2.y * bx + w3.x * by);
					s.ny = w1.w + (w2.z * bx + w3.y * by);
					s.nz = w2.x + (w2.w * bx + w3.z * by);
				}
			}
			else {  // (nrm0 + nrm1 * bar.x) + nrm2 * bar.y   src/scene_trace.cpp:220,249,275
				tu = w0.x + w0.z * bx + w1.x * by;
				tv = w0.y + w0.w * bx + w1.y * by;
				s.nx = w1.z + w2.y * bx + w3.x * by;
				s.ny = w1.w + w2.z * bx + w3.y * by;
				s.nz = w2.x + w2.w * bx + w3.z * by;
			}
		}
		// texDiff: per-quad extent of uv on the single-triangle path, 0 otherwise (src/scene_trace.cpp:219,293-297)
		float tdx = quad_max(tu) - quad_min(tu), tdy = quad_max(tv) - quad_min(tv);
		if(!fast) {
			tdx = 0.0f;
			tdy = 0.0f;
		}
		// which Shade overload runs: fast path keeps the packet's hasMask; the per-lane path is unmasked only when
		// all 16 lanes hit and share one material (src/scene_trace.cpp:224,302-311,343-345)
		const int mat0 = __shfl_sync(SNL_FULL, matId, half * 16);
		const unsigned sameMatB = (__ballot_sync(SNL_FULL, matId == mat0) >> (half * 16)) & 0xffffu;
		const bool masked = fast ? packetHasMask : !(blockFull && sameMatB == 0xffffu);
		if(sel)
			material_shade(S, matId, masked, dx, dy, dz, tu, tv, tdx, tdy, s);
		if(TRANSP) {
			const bool isT = sel && matId >= 0 && (S.mats[matId].flags & 8 /* fTransparency */);
			// lowest lane of the block shaded with my material = its place in the reference's matIds[] list
			const unsigned peers = __match_any_sync(SNL_FULL, sel ? matId : -2 - lane) & (selB << (half * 16));
			int key = isT ? __ffs(peers) : 0;
			int best = key;
#pragma unroll
			for(int o = 8; o; o >>= 1)
				best = max(best, __shfl_xor_sync(SNL_FULL, best, o));
			trans = isT && key == best;
		}
	}
	else if(sel) {
		// simple shading: geometric normal, default material (src/scene_trace.cpp:359-452)
		const float4 tn = ldg4(&S.gtris[obj].n);
		if(INST) {  // TransformVec: rotation rows dotted with the normal
			const DevInstance *I = S.instances + h.inst;
			const float4 r0 = ldg4(&I->r0), r1 = ldg4(&I->r1), r2 = ldg4(&I->r2);
			s.nx = tn.x * r0.x + tn.y * r0.y + tn.z * r0.z;
			s.ny = tn.x * r1.x + tn.y * r1.y + tn.z * r1.z;
			s.nz = tn.x * r2.x + tn.y * r2.y + tn.z * r2.z;
		}
		else {
			s.nx = tn.x;
			s.ny = tn.y;
			s.nz = tn.z;
		}
		material_shade(S, -1, false, dx, dy, dz, 0.0f, 0.0f, 0.0f, 0.0f, s);
	}
}

// Scene::TraceLight (src/scene_trace.cpp:523-601) in two halves so that the shadow ray between them can be traced
// either inline (trace_light) or by the queued traversal kernel of the wavefront pipeline.
// light_setup: direction / distance / N.L of the shadow ray; returns true when the reference casts one (N.L > 0).
__device__ __forceinline__ bool light_setup(const DevLight &L, float px, float py, float pz, float nx, float ny, float nz,
											float &fx, float &fy, float &fz, float &dot, float &distance) {
	float lx = px - L.pos[0], ly = py - L.pos[1], lz = pz - L.pos[2];
	if(lx * lx + ly * ly + lz * lz < 0.0001f) {
		lx = 0.0f;
		ly = 1.0f;
		lz = 0.0f;
	}
	distance = sqrtf(lx * lx + ly * ly + lz * lz);
	const float inv = 1.0f / distance;
	fx = lx * inv;
	fy = ly * inv;
	fz = lz * inv;
	dot = nx * fx + ny * fy + nz * fz;
	return dot > 0.0f;
}

// light_accumulate: attenuation + diffuse / specular sums of an unoccluded light (:575-598).
__device__ __forceinline__ float light_attenuation(const DevLight &L, float distance) {
	const float atten = distance * L.iRadius;
	const float x = ((1.0f - atten) * 0.2f + 1.0f / (16.0f * atten * atten)) - 0.0625f;
	return 0.0f > x ? 0.0f : x;  // Max(f32x4(0), x)
}
__device__ __forceinline__ void light_accumulate(const DevLight &L, float dot, float atten, float &ldr, float &ldg,
												 float &ldb, float &lsr, float &lsg, float &lsb) {
	const float diffMul = dot * atten;
	float specMul = dot;
	specMul *= specMul;
	specMul *= specMul;
	specMul *= specMul;
	specMul *= specMul;
	specMul *= atten;
	ldr += L.color[0] * diffMul;
	ldg += L.color[1] * diffMul;
	ldb += L.color[2] * diffMul;
	lsr += L.color[0] * specMul;
	lsg += L.color[1] * specMul;
	lsb += L.color[2] * specMul;
}

// Returns true when a shadow ray was cast.
template <bool INST = false>
__device__ __forceinline__ bool trace_light(const SceneDev &S, int li, bool sel, const Sample &s, float &ldr, float &ldg,
											float &ldb, float &lsr, float &lsg, float &lsb, TravCount &tc) {
	if(!sel)
		return false;
	const DevLight &L = S.lights[li];
	float fx, fy, fz, dot, distance;
	if(!light_setup(L, s.px, s.py, s.pz, s.nx, s.ny, s.nz, fx, fy, fz, dot, distance))
		return false;
	const float tmax = distance * 0.9999f;
	// Beyond the light's radius the attenuation is exactly 0 and the ray's verdict cannot change the sums (x + 0 = x):
	// the ray still counts as traced - the reference traces it unless the whole packet is out of range
	// (src/scene_trace.cpp:498-499) - but the walk is skipped.
	const float atten = light_attenuation(L, distance);
	if(!(atten > 0.0f))
		return true;
	const bool occluded =
		INST ? traverse_shadow_inst(S, L.pos[0], L.pos[1], L.pos[2], fx, fy, fz, safe_inv(fx), safe_inv(fy), safe_inv(fz), tmax, tc)
			 : traverse_shadow(S.nodes, S.lightPre[li].tris, L.pos[0], L.pos[1], L.pos[2], fx,
							   fy, fz, safe_inv(fx), safe_inv(fy), safe_inv(fz), tmax, tc);
	if(!occluded && tmax > 0.0f)
		light_accumulate(L, dot, atten, ldr, ldg, ldb, lsr, lsg, lsb);
	return true;
}

// Scene::TraceReflection's ray (src/scene_trace.cpp:603-614): r = d - 2 (N.d) N, origin = P + r * 0.001
__device__ __forceinline__ void reflection_ray(const Sample &s, float dx, float dy, float dz, float &rx, float &ry, float &rz,
											   float &rox, float &roy, float &roz) {
	const float dot = s.nx * dx + s.ny * dy + s.nz * dz;
	const float dd = dot + dot;
	rx = dx - s.nx * dd;
	ry = dy - s.ny * dd;
	rz = dz - s.nz * dd;
	rox = s.px + rx * 0.001f;
	roy = s.py + ry * 0.001f;
	roz = s.pz + rz * 0.001f;
}

struct PacketCounts {
	unsigned primary, shadow, reflection, reflection_shadow, transparency;
	unsigned truncated;  // transparency chains cut at kMaxTransp
	int voteGen;  // packet votes taken so far (PacketVote), uniform over the CTA
	int boxGen;   // packet hit boxes posted so far (PacketBox), uniform over the CTA
};

// selector.All() of a packet without a CTA barrier. The verdict is only consumed after the NEXT traversal (it picks
// RayGroup<0,0> or <0,1> for the reflection pass, which only UberMaterial::Shade_ can tell apart), so a warp posts its
// share when its primary rays are done and goes on; the reader spins only if another warp has not posted yet - by
// then it practically always has. Generation-stamped, so it needs no reset between votes: a warp can be at most one
// vote ahead of the slowest reader. packet_vote_init() once per kernel (shared memory starts undefined).
struct PacketVote {
	int arrived;
	int notAllGen[4];
};
__device__ __forceinline__ PacketVote &packet_vote_state() {
	__shared__ PacketVote v;
	return v;
}
__device__ __forceinline__ void packet_vote_init() {
	PacketVote &v = packet_vote_state();
	if(threadIdx.x == 0) {
		v.arrived = 0;
		v.notAllGen[0] = v.notAllGen[1] = v.notAllGen[2] = v.notAllGen[3] = 0;
	}
	__syncthreads();
}
// All 32 lanes of every warp of the CTA call it, the same number of times. Returns the generation to ask about.
__device__ __forceinline__ int packet_vote_post(bool pred, PacketCounts &pc) {
	PacketVote &v = packet_vote_state();
	const int gen = ++pc.voteGen;
	const bool warpAll = __all_sync(SNL_FULL, pred);
	if((threadIdx.x & 31) == 0) {
		if(!warpAll)
			*(volatile int *)&v.notAllGen[gen & 3] = gen;
		__threadfence_block();
		atomicAdd(&v.arrived, 1);
	}
	return gen;
}
__device__ __forceinline__ bool packet_vote_all(int gen) {
	PacketVote &v = packet_vote_state();
	const int need = gen * (int)(blockDim.x >> 5);
	while(*(volatile int *)&v.arrived < need) {
	}
	__threadfence_block();
	return *(volatile int *)&v.notAllGen[gen & 3] != gen;
}

// What one Scene::RayTrace level leaves for the final colour of a lane:
//   out = diffuse * lDiffuse + specular * lSpecular      (src/scene_trace.cpp:504-511)
// with diffuse possibly replaced by VLerp(transColor, diffuse, opacity) once the transparency pass behind the hit has
// returned (:468-483).
// Warp-wide float min / max in one instruction: redux.sync.{min,max}.f32 is new with sm_100a (CREDUX.MIN/MAX.F32).
__device__ __forceinline__ float warp_min_f32(float v) {
	float r;
	asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
	return r;
}
__device__ __forceinline__ float warp_max_f32(float v) {
	float r;
	asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
	return r;
}

// The box of a packet's hit points (minPos / maxPos of Scene::RayTrace, src/scene_trace.cpp:139-140,164-165,378-379)
// for the per-packet light cull (:494-499). Only used when a light can be out of range somewhere (SceneDev::cullMask).
// Every warp posts its own box after the walk; the readers meet at one barrier right before the light loop. Four
// slots: a level-0 box stays posted across the whole reflection pass, which posts and reads its own.
struct PacketBox {
	float v[4][kPacket / 32][6];
};
__device__ __forceinline__ PacketBox &packet_box_state() {
	__shared__ PacketBox b;
	return b;
}
__device__ __forceinline__ int packet_box_post(bool sel, float px, float py, float pz, PacketCounts &pc) {
	float lo[3] = {sel ? px : CUDART_INF_F, sel ? py : CUDART_INF_F, sel ? pz : CUDART_INF_F};
	float hi[3] = {sel ? px : -CUDART_INF_F, sel ? py : -CUDART_INF_F, sel ? pz : -CUDART_INF_F};
#pragma unroll
	for(int a = 0; a < 3; a++) {
		lo[a] = warp_min_f32(lo[a]);
		hi[a] = warp_max_f32(hi[a]);
	}
	const int slot = pc.boxGen++ & 3;
	if((threadIdx.x & 31) == 0) {
		float *d = packet_box_state().v[slot][threadIdx.x >> 5];
		d[0] = lo[0], d[1] = lo[1], d[2] = lo[2], d[3] = hi[0], d[4] = hi[1], d[5] = hi[2];
	}
	return slot;
}
__device__ __forceinline__ void packet_box_read(int slot, float lo[3], float hi[3]) {
	__syncthreads();
	const PacketBox &b = packet_box_state();
	lo[0] = lo[1] = lo[2] = CUDART_INF_F;
	hi[0] = hi[1] = hi[2] = -CUDART_INF_F;
	for(int w = 0; w < (int)(blockDim.x >> 5); w++)
#pragma unroll
		for(int a = 0; a < 3; a++) {
			lo[a] = fminf(lo[a], b.v[slot][w][a]);
			hi[a] = fmaxf(hi[a], b.v[slot][w][3 + a]);
		}
}
// BoxPointDistanceSq (src/funcs.cpp:8-50).
__device__ __forceinline__ float box_point_distance_sq(const float lo[3], const float hi[3], const float p[3]) {
	float sq = 0.0f;
#pragma unroll
	for(int a = 0; a < 3; a++) {
		if(p[a] < lo[a]) {
			const float d = p[a] - lo[a];
			sq += d * d;
		}
		else if(p[a] > hi[a]) {
			const float d = p[a] - hi[a];
			sq += d * d;
		}
	}
	return sq;
}

// Deferred form of the cull for the plain k_render (FLAGS bit 4): no barrier inside the frame. Lanes skip the walk on
// their own where the attenuation is 0 (trace_light), shadow rays of cullable lights are counted per (level, light)
// in shared memory, every warp leaves its level boxes there, and cull_resolve() at the end of the kernel keeps the
// counts of the (level, light) pairs the reference would have traced. cull_state_init() + barrier at kernel start.
enum { kCullLevels = 3 };  // primary rays, first bounce, second bounce (FLAGS bit 8)
struct CullState {
	float box[kCullLevels][kPacket / 32][6];
	unsigned cnt[kCullLevels][kMaxLights];
};
__device__ __forceinline__ CullState &cull_state() {
	__shared__ CullState c;
	return c;
}
__device__ __forceinline__ void cull_state_init() {
	if(threadIdx.x < kCullLevels * kMaxLights)
		(&cull_state().cnt[0][0])[threadIdx.x] = 0;
}
__device__ __forceinline__ void cull_box_store(int level, bool sel, float px, float py, float pz) {
	float lo[3] = {sel ? px : CUDART_INF_F, sel ? py : CUDART_INF_F, sel ? pz : CUDART_INF_F};
	float hi[3] = {sel ? px : -CUDART_INF_F, sel ? py : -CUDART_INF_F, sel ? pz : -CUDART_INF_F};
#pragma unroll
	for(int a = 0; a < 3; a++) {
		lo[a] = warp_min_f32(lo[a]);
		hi[a] = warp_max_f32(hi[a]);
	}
	if((threadIdx.x & 31) == 0) {
		float *d = cull_state().box[level][threadIdx.x >> 5];
		d[0] = lo[0], d[1] = lo[1], d[2] = lo[2], d[3] = hi[0], d[4] = hi[1], d[5] = hi[2];
	}
}
__device__ __forceinline__ void cull_resolve(const SceneDev &S, PacketCounts &pc) {
	__syncthreads();
	if((int)threadIdx.x >= kCullLevels * S.nLights)
		return;
	const int level = (int)threadIdx.x / S.nLights, li = (int)threadIdx.x % S.nLights;
	const CullState &c = cull_state();
	const unsigned n = c.cnt[level][li];
	if(!n)
		return;
	float lo[3] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F}, hi[3] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
	for(int w = 0; w < (int)(blockDim.x >> 5); w++)
#pragma unroll
		for(int a = 0; a < 3; a++) {
			lo[a] = fminf(lo[a], c.box[level][w][a]);
			hi[a] = fmaxf(hi[a], c.box[level][w][3 + a]);
		}
	if(box_point_distance_sq(lo, hi, S.lights[li].pos) > S.lights[li].radSq)
		return;
	if(level == 0)
		pc.shadow += n;
	else
		pc.reflection_shadow += n;
}

struct LevelOut {
	float ar, ag, ab;  // samples.diffuse after the reflection blend
	float lr, lg, lb;  // lDiffuse
	float sr, sg, sb;  // specular * lSpecular
	float opacity;
};

__device__ __forceinline__ void level_compose(const SceneDev &S, const LevelOut &f, float dr, float dg, float db, float &outr,
											  float &outg, float &outb) {
	if(S.nLights) {
		outr = dr * f.lr + f.sr;
		outg = dg * f.lg + f.sg;
		outb = db * f.lb + f.sb;
	}
	else {
		outr = dr;
		outg = dg;
		outb = db;
	}
}

enum { kMaxTransp = 8 };  // transparent surfaces followed along one ray (the reference recursion has no cap)

template <int LEVEL, bool SHARED, int FLAGS>
__device__ __forceinline__ void raytrace_lane(const SceneDev &S, const PreSet originPre, bool active, int maskGen,
											  float ox, float oy, float oz, float dx, float dy, float dz, float &outr,
											  float &outg, float &outb, PacketCounts &pc, TravCount &tc);

// One level of Scene::RayTrace for the calling lane (src/scene_trace.cpp:92-520) up to, not including, the
// transparency pass: nearest hit, Sample + material, reflection bounce (LEVEL 0), lights. Returns true when the lane
// is in transSel with opacity < 1 (:468-472); `hitDist` is its tDistance. DEPTH0: the rays are the level's own
// (else: a transparency continuation, counted as such). All 256 threads of the CTA must call it.
template <int LEVEL, bool SHARED, int FLAGS, bool DEPTH0>
__device__ __forceinline__ bool raytrace_level(const SceneDev &S, const PreSet originPre, bool active, int maskGen,
											   float ox, float oy, float oz, float dx, float dy, float dz, LevelOut &f,
											   float &hitDist, bool &depthOnly, PacketCounts &pc, TravCount &tc) {
	// FLAGS: 1 = the scene can start transparency passes, 2 = instanced scene (DBVH), 4 = deferred light-cull accounting,
	// 8 = TWO reflection bounces (snl_set_reflection_depth(2): the reference's own cap is `cache.reflections < 1`,
	// src/scene_trace.cpp:454; BASELINE config 5 asks for depth 2, which is that condition relaxed to `< 2`)
	constexpr bool TRANSP = (FLAGS & 1) != 0, INST = (FLAGS & 2) != 0;
	constexpr int MAXREFL = (FLAGS & 8) ? 2 : 1;
	Hit h;
	h.dist = CUDART_INF_F;
	h.obj = 0;
	h.inst = 0;
	h.bu = h.bv = 0.0f;
	if(active) {
		const float ix = safe_inv(dx), iy = safe_inv(dy), iz = safe_inv(dz);
		if(INST)
			traverse_nearest_inst<SHARED>(S, ox, oy, oz, dx, dy, dz, ix, iy, iz, h, tc);
		else if(SHARED)
			traverse_nearest<true>(S.nodes, originPre.tris, ox, oy, oz, dx, dy, dz, ix, iy, iz, h, tc);
		else
			traverse_nearest<false>(S.nodes, S.gtris, ox, oy, oz, dx, dy, dz, ix, iy, iz, h, tc);
		if(!DEPTH0)
			pc.transparency++;
		else if(LEVEL == 0)
			pc.primary++;
		else
			pc.reflection++;
	}
	__syncwarp();
	hitDist = h.dist;
	f.opacity = 1.0f;
	f.lr = f.lg = f.lb = f.sr = f.sg = f.sb = 0.0f;

	depthOnly = S.gvDepth != 0;
	if(S.gvDepth) {  // src/scene_trace.cpp:128-137
		const float dist = active ? (h.dist > CUDART_INF_F ? 0.0f : 1.0f / h.dist) : 0.0f;
		f.ar = dist * 20.0f;
		f.ag = dist * 250.0f;
		f.ab = dist * 2.0f;
		return false;
	}

	const bool sel = active && h.dist < CUDART_INF_F;
	Sample s;
	bool trans;
	// maskGen > 0: "ask packet vote maskGen" (posted by the level above before this level's traversal)
	const bool packetHasMask = maskGen > 0 ? !packet_vote_all(maskGen) : maskGen != 0;
	shade_block<TRANSP, INST>(S, sel, h, ox, oy, oz, dx, dy, dz, packetHasMask, s, trans);
	constexpr bool DEFER = (FLAGS & 4) != 0;  // light-cull accounting at the end of the kernel (cull_resolve)
	int boxSlot = 0;
	if(S.cullMask) {
		if(DEFER)
			cull_box_store(LEVEL, sel, s.px, s.py, s.pz);
		else
			boxSlot = packet_box_post(sel, s.px, s.py, s.pz, pc);
	}

	if constexpr(LEVEL < MAXREFL) {
	if(S.gvRefl) {  // src/scene_trace.cpp:454-466, TraceReflection :603-618
		// selector.All() of the packet (src/scene_trace.cpp:615-617) picks RayGroup<0,0> or <0,1> for the bounce; only
		// UberMaterial::Shade_ looks at it, so the vote is skipped for scenes without one
		const int refMask = S.hasUber ? packet_vote_post(sel, pc) : 0;
		float rx, ry, rz, rox, roy, roz;
		reflection_ray(s, dx, dy, dz, rx, ry, rz, rox, roy, roz);
		float cr = 0.0f, cg = 0.0f, cb = 0.0f;
		raytrace_lane<LEVEL + 1, false, FLAGS>(S, PreSet{nullptr}, sel, refMask, rox, roy, roz, rx, ry, rz, cr, cg, cb, pc, tc);
		if(sel) {
			s.dr = s.dr + (cr - s.dr) * 0.3f;
			s.dg = s.dg + (cg - s.dg) * 0.3f;
			s.db = s.db + (cb - s.db) * 0.3f;
		}
	}
	}
	f.ar = s.dr;
	f.ag = s.dg;
	f.ab = s.db;
	f.opacity = s.opacity;

	if(S.nLights) {
		float ldr = S.ambient[0], ldg = S.ambient[1], ldb = S.ambient[2];
		float lsr = 0.0f, lsg = 0.0f, lsb = 0.0f;
		float boxLo[3], boxHi[3];
		if(S.cullMask && !DEFER)
			packet_box_read(boxSlot, boxLo, boxHi);
		for(int li = 0; li < S.nLights; li++) {
			const bool cullable = (S.cullMask >> li & 1) != 0;
			// the packet's hit points are all farther from this light than its radius: src/scene_trace.cpp:498-499
			if(!DEFER && cullable && box_point_distance_sq(boxLo, boxHi, S.lights[li].pos) > S.lights[li].radSq)
				continue;
			const bool cast = trace_light<INST>(S, li, sel, s, ldr, ldg, ldb, lsr, lsg, lsb, tc);
			if(DEFER && cullable) {
				const unsigned m = __ballot_sync(SNL_FULL, cast);
				if((threadIdx.x & 31) == 0 && m)
					atomicAdd(&cull_state().cnt[LEVEL][li], (unsigned)__popc(m));
			}
			else if(cast) {
				if(LEVEL == 0)
					pc.shadow++;
				else
					pc.reflection_shadow++;
			}
			__syncwarp();
		}
		f.lr = ldr;
		f.lg = ldg;
		f.lb = ldb;
		f.sr = s.sr * lsr;
		f.sg = s.sg * lsg;
		f.sb = s.sb * lsb;
	}
	return TRANSP && trans && s.opacity < 1.0f;
}

// Scene::RayTrace for the calling lane. LEVEL 0 = the caller's rays, LEVEL 1 = the reflection pass
// (cache.reflections == 1: no further bounce). TRANSP adds the transparency pass (TraceTransparency,
// src/scene_trace.cpp:468-483,620-634): the reference recurses with the transSel lanes continued from
// dir * (t + 0.001) + origin and lerps the result into diffuse BEFORE the lights are applied; lights and the bounce do
// not depend on that result, so the recursion is unrolled into a loop that keeps one LevelOut per surface and composes
// the colours back to front afterwards with the reference's expression order. All 256 threads must call it.
template <int LEVEL, bool SHARED, int FLAGS>
__device__ __forceinline__ void raytrace_lane(const SceneDev &S, const PreSet originPre, bool active, int maskGen,
											  float ox, float oy, float oz, float dx, float dy, float dz, float &outr,
											  float &outg, float &outb, PacketCounts &pc, TravCount &tc) {
	LevelOut f;
	float hitDist;
	bool depthOnly;
	constexpr bool TRANSP = (FLAGS & 1) != 0;
	bool cont = raytrace_level<LEVEL, SHARED, FLAGS, true>(S, originPre, active, maskGen, ox, oy, oz, dx, dy, dz, f,
															hitDist, depthOnly, pc, tc);
	if(depthOnly) {
		outr = f.ar;
		outg = f.ag;
		outb = f.ab;
		return;
	}
	if(TRANSP) {
		LevelOut behind[kMaxTransp];
		int depth = 0;
		for(;;) {
			if(cont && depth >= kMaxTransp) {  // the reference would keep recursing: counted, reported in snl_stats
				pc.truncated++;
				cont = false;
			}
			// transSel.Any() / transSel.All() of the packet (src/scene_trace.cpp:474, :630-633)
			const int n = __syncthreads_count(cont);
			if(n == 0)
				break;
			if(cont) {
				behind[depth++] = f;
				const float t = hitDist + 0.001f;
				ox = dx * t + ox;
				oy = dy * t + oy;
				oz = dz * t + oz;
			}
			// every lane takes part (shading blocks are warp-wide); only continued lanes take the new level's result
			LevelOut nf;
			float nDist;
			const bool more = raytrace_level<LEVEL, false, FLAGS, false>(S, PreSet{nullptr}, cont, n != kPacket ? -1 : 0, ox,
																		  oy, oz, dx, dy, dz, nf, nDist, depthOnly, pc, tc);
			if(cont) {
				f = nf;
				hitDist = nDist;
			}
			cont = cont && more;
		}
		level_compose(S, f, f.ar, f.ag, f.ab, outr, outg, outb);
		while(depth > 0) {
			const LevelOut &b = behind[--depth];
			// VLerp(transColor, diffuse, opacity) = transColor + (diffuse - transColor) * opacity
			const float r = outr + (b.ar - outr) * b.opacity;
			const float g = outg + (b.ag - outg) * b.opacity;
			const float bl = outb + (b.ab - outb) * b.opacity;
			level_compose(S, b, r, g, bl, outr, outg, outb);
		}
	}
	else
		level_compose(S, f, f.ar, f.ag, f.ab, outr, outg, outb);
}

// ---- cross-GPU signalling (snl_comm.cu): system-scope waits with a time-out so that a missing peer ends in an error
// status instead of a hung device ------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
	unsigned v;
	asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
	asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
	unsigned long long t;
	asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
	return t;
}
enum : unsigned long long { kWaitTimeoutNs = 8000000000ull };
// waits until *p >= v (ge) or *p == v; false = timed out (and *err set)
template <bool GE> __device__ __forceinline__ bool wait_sys(const unsigned *p, unsigned v, unsigned *err) {
	const unsigned long long t0 = global_timer_ns();
	for(;;) {
		const unsigned cur = ld_acquire_sys(p);
		if(GE ? cur >= v : cur == v)
			return true;
		if(global_timer_ns() - t0 > kWaitTimeoutNs) {
			if(err)
				st_release_sys(err, 1u);
			return false;
		}
		__nanosleep(100);
	}
}

// CTA-level aggregation of the per-lane ray counters: warp sums go to `sh` (shared, kCounters entries, zeroed by the
// caller before a barrier) ...
__device__ __forceinline__ void counts_to_shared(const PacketCounts &pc, const TravCount &tc, unsigned *sh) {
	unsigned v[kCounters] = {pc.primary, pc.shadow, pc.reflection, pc.reflection_shadow, tc.nodes, tc.tris, tc.gtris, pc.transparency,
							 pc.truncated};
#pragma unroll
	for(int i = 0; i < kCounters; i++) {
		unsigned x = v[i];
#pragma unroll
		for(int o = 16; o; o >>= 1)
			x += __shfl_xor_sync(SNL_FULL, x, o);
		if((threadIdx.x & 31) == 0 && x)
			atomicAdd(&sh[i], x);
	}
}
// ... and from there (after a barrier) into the scene's global counters.
__device__ __forceinline__ void counts_to_global(const SceneDev &S, const unsigned *sh) {
	if(threadIdx.x < kCounters && sh[threadIdx.x])
		atomicAdd(reinterpret_cast<unsigned long long *>(S.counters) + threadIdx.x, (unsigned long long)sh[threadIdx.x]);
}
__device__ __forceinline__ void flush_counts(const SceneDev &S, const PacketCounts &pc, const TravCount &tc) {
	__shared__ unsigned sh[kCounters];
	if(threadIdx.x < kCounters)
		sh[threadIdx.x] = 0;
	__syncthreads();
	counts_to_shared(pc, tc, sh);
	__syncthreads();
	counts_to_global(S, sh);
}

}  // namespace snl
