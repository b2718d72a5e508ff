/* TEST INFRASTRUCTURE — plain-C restatement of the hot path of nadult/Snail, one ray at a time.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may link or call this file; the product
 * (snail_b200/) never does. It restates the reference's ALGORITHM per ray (the reference walks packets of 256 rays,
 * which only adds conservative culling: src/bounding_box.cpp:208-236, src/triangle.cpp:110-167): the result of a ray
 * is the nearest two-sided hit with t > 0 (strict <, first tested wins on exact ties), shadow rays are one-sided
 * any-hit, shading follows Scene::RayTrace including its 4-quad block structure. All arithmetic is IEEE f32 in the
 * reference's operation order (compile with -ffp-contract=off), i.e. it matches the reference built on veclib's scalar
 * back-end (veclib/scalar/f32.h:102-111: Inv = 1/x, RSqrt = 1/sqrt(x)).
 *
 * Parity pinned: tests/test_oracle_c.py checks this file against (a) the committed goldens generated from the
 * reference itself (tests/golden, tools/make_golden.py) and (b) oracle/_ref (the reference compiled in place) when
 * present. Works on the REFERENCE record layouts (BVH::Node 32 B, Triangle 64 B, ShTriangle 64 B).
 *
 * It also carries the instrumented single-ray walker SURVEY.md 8(d) asks for (node visits / triangle tests per ray).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { /* BVH::Node, src/bvh/tree.h:60-72 */
	float bmin[3], bmax[3];
	int32_t sub; /* subNode, or first | 0x80000000 */
	union {
		struct { int16_t axis, firstNode; } s;
		int32_t count;
	} u;
} ONode;

typedef struct { /* Triangle, src/triangle.h:133-135 */
	float a[3], ba[3], ca[3];
	float t0, it0;
	int32_t pad;
	float plane[4];
} OTri;

typedef struct { /* ShTriangle, src/triangle.h:181-230 */
	float uv[3][2];
	float nrm[3][3];
	int32_t mat; /* matId | flat << 31 */
} OShTri;

typedef struct { /* = snl_material (include/snail_b200.h) */
	int32_t kind, flags;
	float color[3], specular[3], opacity;
	int32_t texture, mask_texture;
} OMat;

typedef struct { /* = snl_texture */
	const unsigned char *data;
	size_t size;
	int32_t width, height, mips;
} OTex;

typedef struct { float pos[3], color[3], radius, radSq, iRadius; } OLight; /* src/light.h:5-17 */

typedef struct oracle_scene {
	const ONode *nodes;
	const OTri *tris;
	const OShTri *shtris;
	const int32_t *matIds;
	const OMat *mats;
	const OTex *texs;
	int nNodes, nTris, nMatIds, nMats, nTex;
	OLight lights[16];
	int nLights;
	float ambient[3];
	int gVals[16];
	int maxRefl; /* bounces followed: 1 = the reference (`cache.reflections < 1`, src/scene_trace.cpp:454); 2 = that condition
	              * relaxed to `< 2`, the comparison for BASELINE config 5 (oracle/_ref/libsnailref_depth2.so) */
	/* instrumentation of the last call */
	unsigned char *undefLanes; /* when set: packet_raytrace reports the lanes the reference shades from uninitialised memory */
	uint64_t nodeVisits, triTests, rays[5]; /* primary, shadow, reflection, reflection shadow, transparency */
} oracle_scene;

typedef struct { float dist; int obj; float bu, bv; } OHit;

#define OINF (__builtin_inff())

static float fmin2(float a, float b) { return a < b ? a : b; } /* veclib/scalar/f32.h Min/Max */
static float fmax2(float a, float b) { return a > b ? a : b; }

/* BBox::Test for one ray (src/bounding_box.cpp:61-97): lmax >= 0 && lmin <= min(lmax, dist) */
static int box_test(const ONode *n, const float o[3], const float id[3], float dist) {
	float l1 = id[0] * (n->bmin[0] - o[0]), l2 = id[0] * (n->bmax[0] - o[0]);
	float lmin = fmin2(l1, l2), lmax = fmax2(l1, l2);
	l1 = id[1] * (n->bmin[1] - o[1]);
	l2 = id[1] * (n->bmax[1] - o[1]);
	lmin = fmax2(lmin, fmin2(l1, l2));
	lmax = fmin2(lmax, fmax2(l1, l2));
	l1 = id[2] * (n->bmin[2] - o[2]);
	l2 = id[2] * (n->bmax[2] - o[2]);
	lmin = fmax2(lmin, fmin2(l1, l2));
	lmax = fmin2(lmax, fmax2(l1, l2));
	return lmax >= 0.0f && lmin <= fmin2(lmax, dist);
}

/* Triangle::Collide(Context&) for one ray (src/triangle.cpp:3-63). The sharedOrigin and per-ray-origin forms differ
 * in where it0 is multiplied in (:13-18 vs :29-37), which changes rounding, so both are restated. */
static void tri_nearest(const OTri *t, int idx, const float o[3], const float d[3], int sharedOrigin, OHit *h) {
	const float *n = t->plane;
	const float det = d[0] * n[0] + d[1] * n[1] + d[2] * n[2];
	const float tx = o[0] - t->a[0], ty = o[1] - t->a[1], tz = o[2] - t->a[2];
	/* tvec0 = ba ^ tvec, tvec1 = tvec ^ ca (veclib/vec3.h:101-106) */
	float t0x = t->ba[1] * tz - t->ba[2] * ty, t0y = t->ba[2] * tx - t->ba[0] * tz, t0z = t->ba[0] * ty - t->ba[1] * tx;
	float t1x = ty * t->ca[2] - tz * t->ca[1], t1y = tz * t->ca[0] - tx * t->ca[2], t1z = tx * t->ca[1] - ty * t->ca[0];
	const float tmul = -(tx * n[0] + ty * n[1] + tz * n[2]);
	float u, v;
	if(sharedOrigin) {
		t0x *= t->it0; t0y *= t->it0; t0z *= t->it0;
		t1x *= t->it0; t1y *= t->it0; t1z *= t->it0;
		v = d[0] * t0x + d[1] * t0y + d[2] * t0z;
		u = d[0] * t1x + d[1] * t1y + d[2] * t1z;
	}
	else {
		v = (d[0] * t0x + d[1] * t0y + d[2] * t0z) * t->it0;
		u = (d[0] * t1x + d[1] * t1y + d[2] * t1z) * t->it0;
	}
	const float w = det - u - v;
	const float mn = fmin2(u, fmin2(v, w)), mx = fmax2(u, fmax2(v, w));
	if(mx <= 0.0f || mn >= 0.0f) {
		const float idet = 1.0f / det;
		const float dist = idet * tmul;
		if(dist < h->dist && dist > 0.0f) {
			h->dist = dist;
			h->obj = idx;
			h->bu = u * idet;
			h->bv = v * idet;
		}
	}
}

/* Triangle::Collide(ShadowContext&) for one ray (src/triangle.cpp:65-102): one-sided, no division */
static int tri_shadow(const OTri *t, const float o[3], const float d[3], float dist) {
	const float *n = t->plane;
	const float det = d[0] * n[0] + d[1] * n[1] + d[2] * n[2];
	const float tx = o[0] - t->a[0], ty = o[1] - t->a[1], tz = o[2] - t->a[2];
	const float t0x = (t->ba[1] * tz - t->ba[2] * ty) * t->it0, t0y = (t->ba[2] * tx - t->ba[0] * tz) * t->it0,
				t0z = (t->ba[0] * ty - t->ba[1] * tx) * t->it0;
	const float t1x = (ty * t->ca[2] - tz * t->ca[1]) * t->it0, t1y = (tz * t->ca[0] - tx * t->ca[2]) * t->it0,
				t1z = (tx * t->ca[1] - ty * t->ca[0]) * t->it0;
	const float tmul = -(tx * n[0] + ty * n[1] + tz * n[2]);
	const float v = d[0] * t0x + d[1] * t0y + d[2] * t0z;
	const float u = d[0] * t1x + d[1] * t1y + d[2] * t1z;
	return fmin2(u, v) >= 0.0f && u + v <= det && tmul > 0.0f && tmul < dist * det;
}

/* BVH::TraversePrimaryN for one ray (src/bvh/traverse.cpp:14-80): explicit stack, near child =
 * firstNode ^ sign[axis] with the sign of THIS ray's direction. */
static void walk_nearest(oracle_scene *s, const float o[3], const float d[3], const float id[3], int sharedOrigin, OHit *h) {
	int stack[128], sp = 0;
	const int sign[3] = {d[0] < 0.0f, d[1] < 0.0f, d[2] < 0.0f};
	stack[sp++] = 0;
	while(sp) {
		int nn = stack[--sp];
		for(;;) {
			const ONode *n = s->nodes + nn;
			s->nodeVisits++;
			if(!box_test(n, o, id, h->dist))
				break;
			if(n->sub & 0x80000000) {
				const int first = n->sub & 0x7fffffff;
				for(int i = 0; i < n->u.count; i++) {
					s->triTests++;
					tri_nearest(s->tris + first + i, first + i, o, d, sharedOrigin, h);
				}
				break;
			}
			const int fn = n->u.s.firstNode ^ sign[n->u.s.axis];
			stack[sp++] = n->sub + (fn ^ 1);
			nn = n->sub + fn;
		}
	}
}

/* BVH::TraverseShadow for one ray (src/bvh/traverse.cpp:82-149) */
static int walk_shadow(oracle_scene *s, const float o[3], const float d[3], const float id[3], float dist) {
	int stack[128], sp = 0;
	const int sign[3] = {d[0] < 0.0f, d[1] < 0.0f, d[2] < 0.0f};
	stack[sp++] = 0;
	while(sp) {
		int nn = stack[--sp];
		for(;;) {
			const ONode *n = s->nodes + nn;
			s->nodeVisits++;
			if(!box_test(n, o, id, dist))
				break;
			if(n->sub & 0x80000000) {
				const int first = n->sub & 0x7fffffff;
				for(int i = 0; i < n->u.count; i++) {
					s->triTests++;
					if(tri_shadow(s->tris + first + i, o, d, dist))
						return 1;
				}
				break;
			}
			const int fn = n->u.s.firstNode ^ sign[n->u.s.axis];
			stack[sp++] = n->sub + (fn ^ 1);
			nn = n->sub + fn;
		}
	}
	return 0;
}

static float safe_inv(float v) { return 1.0f / (v + 0.00000001f); } /* SafeInv, src/rtbase.h:117-119 */

/* ---- RayGenerator (src/ray_generator.cpp:4-47): directions of the 16x16 packet at (x0, y0), lane i = ty*16 + tx --- */
typedef struct { float origin[3], tright[3], tup[3], txyz[3][4]; } ORayGen;

static void raygen_init(const float cam[13], int W, int H, ORayGen *g) {
	const float *pos = cam, *right = cam + 3, *up = cam + 6, *front = cam + 9;
	const float planeDist = cam[12];
	float invW = 1.0f / (float)W, invH = 1.0f / (float)H;
	invW *= (float)W / (float)H;
	const float addx[4] = {0.0f, 1.0f, 0.0f, 1.0f}, addy[4] = {0.0f, 0.0f, 1.0f, 1.0f};
	for(int a = 0; a < 3; a++) {
		g->origin[a] = pos[a];
		g->tright[a] = right[a] * invW;
		g->tup[a] = up[a] * invH;
	}
	for(int k = 0; k < 4; k++) {
		const float taddx = addx[k] - (float)(W * 0.5f), taddy = addy[k] - (float)(H * 0.5f);
		for(int a = 0; a < 3; a++)
			g->txyz[a][k] = (g->tright[a] * taddx + g->tup[a] * taddy) + front[a] * planeDist;
	}
}

static void raygen_dir(const ORayGen *g, int x0, int y0, int tx, int ty, float d[3]) {
	const int k = tx & 3, j = tx >> 2;
	const float xoff = (float)(x0 + (k & 2)), yoff = (float)(y0 - (k >> 1));
	const float tposx = (float)(4 * j) + xoff, tposy = (float)ty + yoff;
	float p[3];
	for(int a = 0; a < 3; a++)
		p[a] = g->tright[a] * tposx + (g->tup[a] * tposy + g->txyz[a][k]);
	const float r = 1.0f / sqrtf(p[0] * p[0] + p[1] * p[1] + p[2] * p[2]);
	for(int a = 0; a < 3; a++)
		d[a] = p[a] * r;
}

/* ---- PointSampler::Sample (src/sampling/point_sampler.cpp:126-210) ---------------------------------------------- */
static void tex_sample(const OTex *T, float tu, float tv, float tdx, float tdy, float rgb[3]) {
	const float wMul = (float)(T->width - 1), hMul = (float)(T->height - 1);
	float ux = tu - (float)(int)tu, uy = tv - (float)(int)tv; /* ClampTexCoord :43-51 */
	ux = ux < 0.0f ? ux + 1.0f : ux;
	uy = uy < 0.0f ? uy + 1.0f : uy;
	const float posx = ux * wMul, posy = uy * hMul;
	unsigned mip = 0, pixels = (unsigned)(fmin2(tdx * wMul, tdy * hMul) * 0.6f);
	while(pixels) {
		mip++;
		pixels >>= 1;
	}
	if(mip > (unsigned)(T->mips - 1))
		mip = T->mips - 1;
	const unsigned char *data = T->data;
	for(unsigned m = 0; m < mip; m++) {
		const int w = (T->width >> m) > 1 ? (T->width >> m) : 1, h = (T->height >> m) > 1 ? (T->height >> m) : 1;
		data += (size_t)w * h * 3;
	}
	const int mw = (T->width >> mip) > 1 ? (T->width >> mip) : 1;
	const int pitch = mw * 3;
	int x1 = (int)posx, y1 = (int)posy;
	int x2 = x1 + 1, y2 = y1 + 1;
	const float fx = posx - (float)x1, fy = posy - (float)y1;
	y1 = T->height - y1; /* vertical flip :153-154 */
	y2 = T->height - y2;
	const int wm = (T->width - 1) >> mip, hm = (T->height - 1) >> mip;
	x1 = ((x1 >> mip) & wm) * 3;
	x2 = ((x2 >> mip) & wm) * 3;
	y1 = ((y1 >> mip) & hm) * pitch;
	y2 = ((y2 >> mip) & hm) * pitch;
	const unsigned char *p0 = data + x1 + y1, *p1 = data + x2 + y1, *p2 = data + x1 + y2, *p3 = data + x2 + y2;
	for(int c = 0; c < 3; c++) {
		const float a = (float)p0[c] + ((float)p1[c] - (float)p0[c]) * fx; /* Lerp, veclib/vecbase.h:80 */
		const float b = (float)p2[c] + ((float)p3[c] - (float)p2[c]) * fx;
		rgb[c] = (a + (b - a) * fy) * (1.0f / 255.0f);
	}
}

typedef struct { float p[3], n[3], diff[3], spec[3], opacity; } OSample;

/* mat->Shade (src/shading/{simple,uber,tex}_material.h Shade_) for one lane; matId -1 = Scene::defaultMat */
static void material_shade(const oracle_scene *s, int matId, int masked, const float d[3], float tu, float tv, float tdx,
						   float tdy, OSample *sm) {
	const float dot = d[0] * sm->n[0] + d[1] * sm->n[1] + d[2] * sm->n[2];
	if(matId < 0) {
		for(int c = 0; c < 3; c++)
			sm->diff[c] = sm->spec[c] = fabsf(dot);
		return;
	}
	const OMat *m = s->mats + matId;
	if(m->kind == 3) { /* TransparentMaterial<1>::Shade_, src/shading/transparent_material.h:15-41 */
		float rgb[3], mask[3];
		tex_sample(s->texs + m->texture, tu, tv, tdx, tdy, rgb);
		tex_sample(s->texs + m->mask_texture, tu, tv, tdx, tdy, mask);
		for(int c = 0; c < 3; c++)
			sm->diff[c] = sm->spec[c] = rgb[c] * dot;
		sm->opacity = mask[0];
	}
	else if(m->kind == 2) {
		float rgb[3];
		tex_sample(s->texs + m->texture, tu, tv, tdx, tdy, rgb);
		for(int c = 0; c < 3; c++)
			sm->diff[c] = sm->spec[c] = rgb[c] * dot;
	}
	else {
		for(int c = 0; c < 3; c++) {
			sm->diff[c] = m->color[c] * fabsf(dot);
			sm->spec[c] = (m->kind == 1 && masked) ? m->specular[c] : sm->diff[c]; /* uber_material.h:21-29 */
		}
		if(m->kind == 1)
			sm->opacity = m->opacity; /* dissFactor, uber_material.h:23,28 */
	}
}

/* ---- Scene::RayTrace on one packet of up to 256 lanes (src/scene_trace.cpp:92-520) ------------------------------- */
typedef struct {
	int n; /* lanes, multiple of 16 */
	float o[256][3], d[256][3];
	unsigned char active[256];
} OPacket;

static void packet_raytrace(oracle_scene *s, const OPacket *pk, int level, int sharedOrigin, int packetHasMask,
							float out[256][3]);

/* Scene::TraceLight for one lane (src/scene_trace.cpp:523-601); returns 1 when a shadow ray was cast */
static int trace_light(oracle_scene *s, const OLight *L, const OSample *sm, float ld[3], float ls[3]) {
	float l[3] = {sm->p[0] - L->pos[0], sm->p[1] - L->pos[1], sm->p[2] - L->pos[2]};
	if(l[0] * l[0] + l[1] * l[1] + l[2] * l[2] < 0.0001f) {
		l[0] = 0.0f;
		l[1] = 1.0f;
		l[2] = 0.0f;
	}
	const float distance = sqrtf(l[0] * l[0] + l[1] * l[1] + l[2] * l[2]);
	const float inv = 1.0f / distance;
	const float f[3] = {l[0] * inv, l[1] * inv, l[2] * inv};
	const float dot = sm->n[0] * f[0] + sm->n[1] * f[1] + sm->n[2] * f[2];
	if(!(dot > 0.0f))
		return 0;
	const float tmax = distance * 0.9999f;
	const float id[3] = {safe_inv(f[0]), safe_inv(f[1]), safe_inv(f[2])};
	const int occluded = walk_shadow(s, L->pos, f, id, tmax);
	if(!occluded && tmax > 0.0f) {
		float atten = distance * L->iRadius;
		const float x = ((1.0f - atten) * 0.2f + 1.0f / (16.0f * atten * atten)) - 0.0625f;
		atten = 0.0f > x ? 0.0f : x;
		const float diffMul = dot * atten;
		float specMul = dot;
		specMul *= specMul;
		specMul *= specMul;
		specMul *= specMul;
		specMul *= specMul;
		specMul *= atten;
		for(int c = 0; c < 3; c++) {
			ld[c] += L->color[c] * diffMul;
			ls[c] += L->color[c] * specMul;
		}
	}
	return 1;
}

/* Sample construction + material evaluation for one 16-lane block (4 quads), src/scene_trace.cpp:139-452 */
static void shade_block(const oracle_scene *s, const OPacket *pk, int b0, const OHit *hit, const unsigned char *sel,
						int packetHasMask, OSample *sm, unsigned char *trans, unsigned char *undef) {
	int nsel = 0, same = 1, sameMat = 1;
	int matId[16];
	float tu[16], tv[16];
	const int obj0 = sel[b0] ? hit[b0].obj : 0;
	for(int l = 0; l < 16; l++) {
		const int i = b0 + l;
		OSample *x = sm + i;
		for(int c = 0; c < 3; c++) {
			x->p[c] = pk->d[i][c] * hit[i].dist + pk->o[i][c]; /* :163 */
			x->n[c] = x->diff[c] = x->spec[c] = 0.0f;
		}
		x->opacity = 1.0f;
		trans[i] = 0;
		nsel += sel[i];
		same &= (sel[i] ? hit[i].obj : 0) == obj0;
		matId[l] = -1;
		tu[l] = tv[l] = 0.0f;
	}
	const int blockFull = nsel == 16, fast = blockFull && same;
	/* Reference quirk (src/scene_trace.cpp:240-290 full shading, :410-423 simple shading): lane 0 of a quad supplies
	 * the quad's triangle; when it is inactive or missed its object reads 0, and the per-lane fix-up skips lanes whose
	 * own triangle is 0 as "same as lane 0" -> those lanes are shaded from an uninitialised Sample. We shade them
	 * properly and report them as undefined so that fixtures can exclude them. */
	int poisoned = 0;
	for(int q = 0; q < 4; q++)
		if(!sel[b0 + q * 4])
			for(int k = 1; k < 4; k++)
				poisoned |= sel[b0 + q * 4 + k] && hit[b0 + q * 4 + k].obj == 0;
	/* such a lane keeps matId ~0 and a garbage normal, which also steers block-level decisions (one material for the
	 * block or not -> masked / unmasked Shade, the list of materials) -> the whole block counts as undefined */
	if(poisoned)
		for(int l = 0; l < 16; l++)
			undef[b0 + l] |= sel[b0 + l];
	if(s->gVals[6] && s->shtris) {
		for(int l = 0; l < 16; l++) {
			const int i = b0 + l;
			if(!sel[i])
				continue;
			const OShTri *st = s->shtris + hit[i].obj;
			const int mi = st->mat & 0x7fffffff, flat = st->mat < 0;
			int m = mi < s->nMatIds ? s->matIds[mi] : -1;
			if(m >= s->nMats)
				m = -1;
			matId[l] = m;
			const int texCoords = m >= 0 && (s->mats[m].flags & 1);
			const float bx = hit[i].bu, by = hit[i].bv;
			OSample *x = sm + i;
			if(fast && !texCoords) {
				for(int c = 0; c < 3; c++) /* nrm0 + (nrm1*bx + nrm2*by), :204 */
					x->n[c] = flat ? st->nrm[0][c] : st->nrm[0][c] + (st->nrm[1][c] * bx + st->nrm[2][c] * by);
			}
			else { /* (nrm0 + nrm1*bx) + nrm2*by, :220,249,275 */
				tu[l] = st->uv[0][0] + st->uv[1][0] * bx + st->uv[2][0] * by;
				tv[l] = st->uv[0][1] + st->uv[1][1] * bx + st->uv[2][1] * by;
				for(int c = 0; c < 3; c++)
					x->n[c] = st->nrm[0][c] + st->nrm[1][c] * bx + st->nrm[2][c] * by;
			}
		}
		for(int l = 0; l < 16; l++)
			sameMat &= matId[l] == matId[0];
		const int masked = fast ? packetHasMask : !(blockFull && sameMat); /* :224,302-311,343-345 */
		for(int q = 0; q < 4; q++) {
			float tdx = 0.0f, tdy = 0.0f;
			if(fast) { /* texDiff = per-quad extent of uv on the single-triangle path, :219 */
				float mnu = tu[q * 4], mxu = mnu, mnv = tv[q * 4], mxv = mnv;
				for(int k = 1; k < 4; k++) {
					mnu = fmin2(mnu, tu[q * 4 + k]);
					mxu = fmax2(mxu, tu[q * 4 + k]);
					mnv = fmin2(mnv, tv[q * 4 + k]);
					mxv = fmax2(mxv, tv[q * 4 + k]);
				}
				tdx = mxu - mnu;
				tdy = mxv - mnv;
			}
			for(int k = 0; k < 4; k++) {
				const int l = q * 4 + k, i = b0 + l;
				if(sel[i])
					material_shade(s, matId[l], masked, pk->d[i], tu[l], tv[l], tdx, tdy, sm + i);
			}
		}
		/* transSel (:190,306,347-349): one material for the block -> its fTransparency flag for every lane; else each
		 * transparent material, in order of first appearance among the selected lanes, OVERWRITES the block's mask
		 * (`transSel[b4 + q] = mask[q]`), so only the last one keeps its lanes */
		int last = -1;
		for(int l = 0; l < 16; l++) {
			if(!sel[b0 + l] || matId[l] < 0 || !(s->mats[matId[l]].flags & 8))
				continue;
			int seen = 0;
			for(int k = 0; k < l; k++)
				seen |= sel[b0 + k] && matId[k] == matId[l];
			if(!seen)
				last = matId[l];
		}
		for(int l = 0; l < 16; l++)
			trans[b0 + l] = last >= 0 && sel[b0 + l] && matId[l] == last;
	}
	else { /* simple shading: geometric normal, default material (:359-452) */
		for(int l = 0; l < 16; l++) {
			const int i = b0 + l;
			if(!sel[i])
				continue;
			for(int c = 0; c < 3; c++)
				sm[i].n[c] = s->tris[hit[i].obj].plane[c];
			material_shade(s, -1, 0, pk->d[i], 0.0f, 0.0f, 0.0f, 0.0f, sm + i);
		}
	}
}

/* level = cache.reflections (1 inside the reflection pass: no further bounce); transp = inside a transparency pass
 * (only changes which counter the rays go to). Recursive like the reference; buffers live on the stack. */
static void packet_raytrace_ex(oracle_scene *s, const OPacket *pk, int level, int transp, int sharedOrigin,
							   int packetHasMask, float out[256][3], unsigned char undef[256]) {
	OHit hit[256];
	OSample sm[256];
	unsigned char sel[256], trans[256];
	for(int i = 0; i < pk->n; i++) {
		hit[i].dist = OINF;
		hit[i].obj = 0;
		hit[i].bu = hit[i].bv = 0.0f;
		out[i][0] = out[i][1] = out[i][2] = 0.0f;
		if(pk->active[i]) {
			const float id[3] = {safe_inv(pk->d[i][0]), safe_inv(pk->d[i][1]), safe_inv(pk->d[i][2])};
			walk_nearest(s, pk->o[i], pk->d[i], id, sharedOrigin, &hit[i]);
			s->rays[transp ? 4 : level ? 2 : 0]++;
		}
	}
	if(s->gVals[1]) { /* depth shading, :128-137 */
		for(int i = 0; i < pk->n; i++) {
			const float dist = pk->active[i] ? (hit[i].dist > OINF ? 0.0f : 1.0f / hit[i].dist) : 0.0f;
			out[i][0] = dist * 20.0f;
			out[i][1] = dist * 250.0f;
			out[i][2] = dist * 2.0f;
		}
		return;
	}
	int all = 1;
	for(int i = 0; i < pk->n; i++) {
		sel[i] = pk->active[i] && hit[i].dist < OINF;
		all &= sel[i];
	}
	for(int b0 = 0; b0 < pk->n; b0 += 16)
		shade_block(s, pk, b0, hit, sel, packetHasMask, sm, trans, undef);

	if(level < s->maxRefl && s->gVals[7]) { /* TraceReflection :603-618, blend :454-466 */
		OPacket rp;
		float rc[256][3];
		rp.n = pk->n;
		for(int i = 0; i < pk->n; i++) {
			const float *d = pk->d[i], *n = sm[i].n;
			const float dot = n[0] * d[0] + n[1] * d[1] + n[2] * d[2];
			const float dd = dot + dot;
			for(int c = 0; c < 3; c++) {
				rp.d[i][c] = d[c] - n[c] * dd;
				rp.o[i][c] = sm[i].p[c] + rp.d[i][c] * 0.001f;
			}
			rp.active[i] = sel[i];
		}
		unsigned char ru[256] = {0};
		packet_raytrace_ex(s, &rp, level + 1, 0, 0, !all, rc, ru);
		for(int i = 0; i < pk->n; i++)
			undef[i] |= sel[i] && ru[i];
		for(int i = 0; i < pk->n; i++)
			if(sel[i])
				for(int c = 0; c < 3; c++)
					sm[i].diff[c] = sm[i].diff[c] + (rc[i][c] - sm[i].diff[c]) * 0.3f;
	}
	{ /* transparency pass :468-483, TraceTransparency :620-634 (cache.reflections is NOT bumped: the continued rays
	   * bounce again on level 0) */
		int any = 0, allT = 1;
		for(int i = 0; i < pk->n; i++) {
			trans[i] = trans[i] && sm[i].opacity < 1.0f;
			any |= trans[i];
			allT &= trans[i];
		}
		if(any) {
			OPacket tp;
			float tcol[256][3];
			tp.n = pk->n;
			for(int i = 0; i < pk->n; i++) {
				const float t = hit[i].dist + 0.001f;
				for(int c = 0; c < 3; c++) {
					tp.d[i][c] = pk->d[i][c];
					tp.o[i][c] = pk->d[i][c] * t + pk->o[i][c];
				}
				tp.active[i] = trans[i];
			}
			unsigned char tu[256] = {0};
			packet_raytrace_ex(s, &tp, level, 1, 0, !allT, tcol, tu);
			for(int i = 0; i < pk->n; i++)
				undef[i] |= trans[i] && tu[i];
			for(int i = 0; i < pk->n; i++)
				if(trans[i])
					for(int c = 0; c < 3; c++) /* VLerp(a, b, x) = a + (b - a) * x, veclib/vecbase.h:81 */
						sm[i].diff[c] = tcol[i][c] + (sm[i].diff[c] - tcol[i][c]) * sm[i].opacity;
		}
	}
	/* per-packet light cull (src/scene_trace.cpp:139-140,164-165,494-499): the box of the packet's hit points against each
	   light's radius, BoxPointDistanceSq of src/funcs.cpp:8-50 */
	unsigned char lightOn[64];
	{
		float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
		for(int i = 0; i < pk->n; i++)
			if(sel[i])
				for(int c = 0; c < 3; c++) {
					lo[c] = sm[i].p[c] < lo[c] ? sm[i].p[c] : lo[c];
					hi[c] = sm[i].p[c] > hi[c] ? sm[i].p[c] : hi[c];
				}
		for(int li = 0; li < s->nLights && li < 64; li++) {
			const float *p = s->lights[li].pos;
			float sq = 0.0f;
			for(int c = 0; c < 3; c++) {
				if(p[c] < lo[c]) {
					const float d = p[c] - lo[c];
					sq += d * d;
				}
				else if(p[c] > hi[c]) {
					const float d = p[c] - hi[c];
					sq += d * d;
				}
			}
			lightOn[li] = !(sq > s->lights[li].radSq);
		}
	}
	for(int i = 0; i < pk->n; i++) {
		if(s->nLights) {
			float ld[3] = {s->ambient[0], s->ambient[1], s->ambient[2]}, ls[3] = {0.0f, 0.0f, 0.0f};
			for(int li = 0; li < s->nLights; li++)
				if(lightOn[li & 63] && sel[i] && trace_light(s, &s->lights[li], &sm[i], ld, ls))
					s->rays[level ? 3 : 1]++;
			for(int c = 0; c < 3; c++)
				out[i][c] = sm[i].diff[c] * ld[c] + sm[i].spec[c] * ls[c];
		}
		else
			for(int c = 0; c < 3; c++)
				out[i][c] = sm[i].diff[c];
	}
}

static void packet_raytrace(oracle_scene *s, const OPacket *pk, int level, int sharedOrigin, int packetHasMask,
							float out[256][3]) {
	unsigned char undef[256] = {0};
	packet_raytrace_ex(s, pk, level, 0, sharedOrigin, packetHasMask, out, undef);
	if(s->undefLanes)
		memcpy(s->undefLanes, undef, 256);
}

/* ---- exported API (bound by oracle/snail_oracle.py) --------------------------------------------------------------- */

oracle_scene *oracle_create(const void *nodes, int nNodes, const void *tris, const void *shtris, int nTris,
							const int32_t *matIds, int nMatIds, const void *mats, int nMats, const void *texs, int nTex) {
	oracle_scene *s = (oracle_scene *)calloc(1, sizeof(oracle_scene));
	s->maxRefl = 1;
	s->nodes = (const ONode *)nodes;
	s->tris = (const OTri *)tris;
	s->shtris = (const OShTri *)shtris;
	s->matIds = matIds;
	s->mats = (const OMat *)mats;
	s->texs = (const OTex *)texs;
	s->nNodes = nNodes;
	s->nTris = nTris;
	s->nMatIds = nMatIds;
	s->nMats = nMats;
	s->nTex = nTex;
	s->ambient[0] = s->ambient[1] = s->ambient[2] = 0.1f; /* src/scene.cpp:9 */
	return s;
}
void oracle_destroy(oracle_scene *s) { free(s); }
void oracle_set_gvals(oracle_scene *s, const int *g) { memcpy(s->gVals, g, sizeof(s->gVals)); }
void oracle_set_reflection_depth(oracle_scene *s, int depth) { s->maxRefl = depth < 1 ? 1 : depth > 2 ? 2 : depth; }
void oracle_set_lights(oracle_scene *s, const float *l7, int n) { /* pos, color, radius */
	s->nLights = n;
	for(int i = 0; i < n; i++) {
		memcpy(s->lights[i].pos, l7 + i * 7, 12);
		memcpy(s->lights[i].color, l7 + i * 7 + 3, 12);
		s->lights[i].radius = l7[i * 7 + 6];
		s->lights[i].radSq = s->lights[i].radius * s->lights[i].radius;
		s->lights[i].iRadius = 1.0f / s->lights[i].radius;
	}
}
static void reset_counts(oracle_scene *s) {
	s->nodeVisits = s->triTests = 0;
	memset(s->rays, 0, sizeof(s->rays));
}
/* counts: nodeVisits, triTests, primary, shadow, reflection, reflection shadow, transparency */
void oracle_counts(const oracle_scene *s, uint64_t out[7]) {
	out[0] = s->nodeVisits;
	out[1] = s->triTests;
	memcpy(out + 2, s->rays, sizeof(s->rays));
}

/* BVH::TraversePrimary on the generated primary rays of the W x H image (row-major outputs) */
void oracle_trace_primary(oracle_scene *s, const float cam[13], int W, int H, int *triId, float *t, float *bary) {
	ORayGen g;
	raygen_init(cam, W, H, &g);
	reset_counts(s);
	for(int y = 0; y < H; y++)
		for(int x = 0; x < W; x++) {
			float d[3];
			raygen_dir(&g, x & ~15, y & ~15, x & 15, y & 15, d);
			const float id[3] = {safe_inv(d[0]), safe_inv(d[1]), safe_inv(d[2])};
			OHit h = {OINF, 0, 0.0f, 0.0f};
			walk_nearest(s, g.origin, d, id, 1, &h);
			const size_t o = (size_t)y * W + x;
			const int hit = h.dist < OINF;
			triId[o] = hit ? h.obj : -1;
			t[o] = h.dist;
			bary[o * 2] = hit ? h.bu : 0.0f;
			bary[o * 2 + 1] = hit ? h.bv : 0.0f;
		}
}

void oracle_trace_rays(oracle_scene *s, const float *o3, const float *d3, const unsigned char *mask, int n,
					   int sharedOrigin, int *triId, float *t, float *bary) {
	reset_counts(s);
	for(int i = 0; i < n; i++) {
		OHit h = {OINF, 0, 0.0f, 0.0f};
		if(!mask || mask[i]) {
			const float *o = sharedOrigin ? o3 : o3 + (size_t)i * 3, *d = d3 + (size_t)i * 3;
			const float id[3] = {safe_inv(d[0]), safe_inv(d[1]), safe_inv(d[2])};
			walk_nearest(s, o, d, id, sharedOrigin, &h);
		}
		const int hit = h.dist < OINF;
		triId[i] = hit ? h.obj : -1;
		t[i] = h.dist;
		bary[i * 2] = hit ? h.bu : 0.0f;
		bary[i * 2 + 1] = hit ? h.bv : 0.0f;
	}
}

void oracle_trace_shadow(oracle_scene *s, const float origin[3], const float *d3, float *tmax, int n) {
	reset_counts(s);
	for(int i = 0; i < n; i++) {
		if(tmax[i] < 0.0f)
			continue;
		const float *d = d3 + (size_t)i * 3;
		const float id[3] = {safe_inv(d[0]), safe_inv(d[1]), safe_inv(d[2])};
		if(walk_shadow(s, origin, d, id, tmax[i]))
			tmax[i] = -OINF;
	}
}

/* ConvColor (src/render.cpp:11-17) */
static unsigned char conv(float c) {
	float v = c * 255.0f;
	v = v > 0.0f ? v : 0.0f;
	v = v < 255.0f ? v : 255.0f;
	return (unsigned char)(int)v;
}

/* Render() (src/render.cpp:47-211, 214-243): 16x16 packets, BGR8 rows of 3*W bytes. gVals[9]: each packet is traced
 * as four packets of the 2W x 2H image and box-filtered with the reference's add order (:71-110). */
void oracle_render_ex(oracle_scene *s, const float cam[13], int W, int H, unsigned char *out, unsigned char *undefOut);
void oracle_render(oracle_scene *s, const float cam[13], int W, int H, unsigned char *out) {
	oracle_render_ex(s, cam, W, H, out, NULL);
}
/* undefOut (W*H bytes, may be NULL): 1 where the reference's pixel depends on an uninitialised Sample (see shade_block) */
void oracle_render_ex(oracle_scene *s, const float cam[13], int W, int H, unsigned char *out, unsigned char *undefOut) {
	static _Thread_local OPacket pk;
	unsigned char lanes[256], ulanes[256];
	s->undefLanes = undefOut ? lanes : NULL;
	static _Thread_local float col[256][3], sub[256][3];
	const int ss = s->gVals[9] != 0;
	ORayGen g;
	raygen_init(cam, ss ? 2 * W : W, ss ? 2 * H : H, &g);
	reset_counts(s);
	for(int y0 = 0; y0 < H; y0 += 16)
		for(int x0 = 0; x0 < W; x0 += 16) {
			pk.n = 256;
			for(int k = 0; k < (ss ? 4 : 1); k++) {
				const int sx = ss ? x0 * 2 + (k & 1) * 16 : x0, sy = ss ? y0 * 2 + (k >> 1) * 16 : y0;
				for(int i = 0; i < 256; i++) {
					raygen_dir(&g, sx, sy, i & 15, i >> 4, pk.d[i]);
					memcpy(pk.o[i], g.origin, 12);
					pk.active[i] = 1; /* overhanging lanes are traced too, src/render.cpp:112-115 */
				}
				packet_raytrace(s, &pk, 0, 1, 0, ss ? sub : col);
				if(undefOut && !ss)
					memcpy(ulanes, lanes, 256);
				if(undefOut && ss)
					for(int oy = 0; oy < 8; oy++)
						for(int ox = 0; ox < 8; ox++) {
							const int i0 = (2 * oy) * 16 + 2 * ox;
							ulanes[((k >> 1) * 8 + oy) * 16 + (k & 1) * 8 + ox] = lanes[i0] | lanes[i0 + 1] | lanes[i0 + 16] | lanes[i0 + 17];
						}
				if(ss)
					for(int oy = 0; oy < 8; oy++)
						for(int ox = 0; ox < 8; ox++)
							for(int c = 0; c < 3; c++) {
								const float *t0 = &sub[(2 * oy) * 16 + 2 * ox][c], *t1 = &sub[(2 * oy + 1) * 16 + 2 * ox][c];
								/* (tcolors[q] + tcolors[q + line]) * 0.25, then lane 0 + lane 1 */
								const float l = (t0[0] + t1[0]) * 0.25f, r = (t0[3] + t1[3]) * 0.25f;
								col[((k >> 1) * 8 + oy) * 16 + (k & 1) * 8 + ox][c] = l + r;
							}
			}
			for(int i = 0; i < 256; i++) {
				const int x = x0 + (i & 15), y = y0 + (i >> 4);
				if(x < W && y < H) {
					unsigned char *dst = out + ((size_t)y * W + x) * 3;
					dst[0] = conv(col[i][2]);
					dst[1] = conv(col[i][1]);
					dst[2] = conv(col[i][0]);
					if(undefOut)
						undefOut[(size_t)y * W + x] = ulanes[i];
				}
			}
		}
	s->undefLanes = NULL;
}
