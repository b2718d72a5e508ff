// TEST INFRASTRUCTURE — not part of the product.
//
// C API over the *unmodified* nadult/Snail reference, compiled in place from
// /root/reference by oracle/build_ref.sh into oracle/_ref/libsnailref_{sse,scalar}.so.
// It exists so that tests/, bench.py's cpu_baseline / --impl reference legs and
// tools/make_golden.py can drive the reference's own code for the hot path:
//   BaseScene::LoadWavefrontObj      src/formats/wavefront_obj.cpp:66-200
//   BVH::Construct                   src/bvh/tree.cpp:293-328
//   Render()                         src/render.cpp:276-280
//   Scene<BVH>::RayTrace             src/scene_trace.cpp:92-520
//   BVH::TraversePrimary/Shadow      src/bvh/traverse.cpp:14-149
// The scene set-up mirrors src/rtracer.cpp:505-586 minus the GL window.
// Nothing here is linked into libsnail_b200.so.

#include "rtbase.h"
#include "base_scene.h"
#include "bvh/tree.h"
#include "dbvh/tree.h"
#include "camera.h"
#include "mipmap_texture.h"
#include "ray_generator.h"
#include "render.h"
#include "scene.h"
#include "shading/material.h"

#include <chrono>
#include <cstdio>
#include <cstring>
#include <map>

namespace {

struct RefScene {
	Scene<BVH> scene;
	vector<shading::MaterialDesc> matDescs;
	// instanced form (src/rtracer.cpp:146-178,359-363): a DBVH over ObjectInstances of `scene.geometry`
	Scene<DBVH> dscene;
	bool instanced = false;
};

Camera MakeCamera(const float *c) {
	return Camera(Vec3f(c[0], c[1], c[2]), Vec3f(c[3], c[4], c[5]), Vec3f(c[6], c[7], c[8]),
				  Vec3f(c[9], c[10], c[11]), c[12]);
}

void CopyStats(const TreeStats &st, unsigned *out) {
	if(!out)
		return;
	out[0] = st.GetIntersects();
	out[1] = st.GetLoopIters();
	out[2] = st.GetRays();
	out[3] = st.GetSkips();
}

} // namespace

extern "C" {

// flipNormalsFlag / buildFlags have the meaning of conf.flipNormals / conf.buildFlags in
// src/rtracer.cpp:192,482-487 (flag 0 => normals ARE flipped, the program default).
void *ref_scene_load(const char *objPath, const char *mtlPath, const char *texDir, int flipNormalsFlag,
					 int buildFlags) {
	RefScene *rs = new RefScene;
	BaseScene base;
	base.LoadWavefrontObj(objPath);
	if(!flipNormalsFlag)
		base.FlipNormals();
	base.GenNormals();
	rs->scene.geometry.Construct(base, buildFlags);

	if(mtlPath && mtlPath[0])
		rs->matDescs = shading::LoadMaterialDescs(mtlPath);
	rs->scene.texDict = shading::LoadTextures(rs->matDescs, texDir ? texDir : "");
	rs->scene.matDict = shading::MakeMaterials(rs->matDescs, rs->scene.texDict);
	rs->scene.UpdateMaterials();
	rs->scene.geometry.UpdateMaterialIds(rs->scene.GetMatIdMap());
	return rs;
}

void ref_scene_free(void *h) { delete(RefScene *)h; }

void ref_scene_counts(void *h, int *nNodes, int *nTris, int *nShTris, int *nMatIds, int *depth) {
	const BVH &bvh = ((RefScene *)h)->scene.geometry;
	*nNodes = (int)bvh.nodes.size();
	*nTris = (int)bvh.tris.size();
	*nShTris = (int)bvh.shTris.size();
	*nMatIds = (int)bvh.materials.size();
	*depth = bvh.depth;
}

int ref_sizeof_node() { return (int)sizeof(BVH::Node); }
int ref_sizeof_triangle() { return (int)sizeof(Triangle); }
int ref_sizeof_shtriangle() { return (int)sizeof(ShTriangle); }

const void *ref_scene_nodes(void *h) { return ((RefScene *)h)->scene.geometry.nodes.data(); }
const void *ref_scene_tris(void *h) { return ((RefScene *)h)->scene.geometry.tris.data(); }
const void *ref_scene_shtris(void *h) { return ((RefScene *)h)->scene.geometry.shTris.data(); }

// BVH::materials[i].id (src/bvh/tree.h:74-84): scene material index or ~0 for the default material.
void ref_scene_matids(void *h, int *out) {
	const BVH &bvh = ((RefScene *)h)->scene.geometry;
	for(size_t n = 0; n < bvh.materials.size(); n++)
		out[n] = bvh.materials[n].id;
}

int ref_scene_num_materials(void *h) { return (int)((RefScene *)h)->scene.materials.size(); }

// Per scene material (Scene::materials order, src/scene.cpp:12-21):
//   flags (shading::Material::f*), has-texture, texture w,h,mips.
void ref_scene_material_info(void *h, int idx, int *flags, int *texW, int *texH, int *texMips) {
	const auto &mat = ((RefScene *)h)->scene.materials[idx];
	*flags = mat->flags;
	const MipmapTexture *tex = mat->GetTexture();
	*texW = tex ? (int)tex->Width() : 0;
	*texH = tex ? (int)tex->Height() : 0;
	*texMips = tex ? (int)tex->Mips() : 0;
}

// Copies the whole RGB8 mip chain (MipmapTexture layout, src/mipmap_texture.cpp:299-309) of a
// material's diffuse texture; returns the byte count (call with out = 0 to size the buffer).
long ref_scene_material_texture(void *h, int idx, unsigned char *out) {
	const MipmapTexture *tex = ((RefScene *)h)->scene.materials[idx]->GetTexture();
	if(!tex)
		return 0;
	long total = 0;
	for(size_t m = 0; m < tex->Mips(); m++) {
		if(out)
			memcpy(out + total, tex->DataPointer(m), tex->Size(m));
		total += (long)tex->Size(m);
	}
	return total;
}

void ref_scene_bbox(void *h, float *out) {
	BBox box = ((RefScene *)h)->scene.geometry.GetBBox();
	out[0] = box.min.x; out[1] = box.min.y; out[2] = box.min.z;
	out[3] = box.max.x; out[4] = box.max.y; out[5] = box.max.z;
}

void ref_set_gvals(const int *g) {
	for(int n = 0; n < 16; n++)
		gVals[n] = g[n];
}

// lights: n x {pos.xyz, color.rgb, radius}  (Light ctor, src/light.h:8-13)
void ref_set_lights(void *h, const float *l, int n) {
	auto &lights = ((RefScene *)h)->scene.lights;
	lights.clear();
	for(int i = 0; i < n; i++, l += 7)
		lights.push_back(Light(Vec3f(l[0], l[1], l[2]), Vec3f(l[3], l[4], l[5]), l[6]));
	((RefScene *)h)->dscene.lights = lights;
}

// cam: pos.xyz right.xyz up.xyz front.xyz plane_dist. out: H rows of 3*W bytes, byte order of
// ConvColor (src/render.cpp:11-17). Returns wall milliseconds of the Render() call.
double ref_render(void *h, const float *cam, int W, int H, int threads, unsigned char *out, unsigned *stats) {
	RefScene *rs = (RefScene *)h;
	MipmapTexture image(W, H, fwk::GlFormat::rgb8, 1);
	memset(image.DataPointer(), 0, image.Size());
	auto t0 = std::chrono::steady_clock::now();
	TreeStats st = rs->instanced ? Render(rs->dscene, MakeCamera(cam), image, Options(), (uint)threads)
								 : Render(rs->scene, MakeCamera(cam), image, Options(), (uint)threads);
	auto t1 = std::chrono::steady_clock::now();
	if(out)
		for(int y = 0; y < H; y++)
			memcpy(out + (size_t)y * W * 3, image.DataPointer() + image.Pitch() * y, (size_t)W * 3);
	CopyStats(st, stats);
	return std::chrono::duration<double, std::milli>(t1 - t0).count();
}

// The farm overload (src/render.h:20-23, src/render.cpp:243-265): tiles coords[4i..] written at data + offsets[i] in
// the planar delta format of RenderTask::Work with compress == 1.
double ref_render_tiles(void *h, const float *cam, int resx, int resy, unsigned char *data, const int *coords,
						const int *offsets, int nTiles, int rank, int threads, unsigned *stats) {
	RefScene *rs = (RefScene *)h;
	vector<int> c(coords, coords + (size_t)nTiles * 4), o(offsets, offsets + nTiles);
	auto t0 = std::chrono::steady_clock::now();
	TreeStats st = rs->instanced ? Render(rs->dscene, MakeCamera(cam), (uint)resx, (uint)resy, data, c, o, Options(), (uint)rank, (uint)threads)
								 : Render(rs->scene, MakeCamera(cam), (uint)resx, (uint)resy, data, c, o, Options(), (uint)rank, (uint)threads);
	auto t1 = std::chrono::steady_clock::now();
	CopyStats(st, stats);
	return std::chrono::duration<double, std::milli>(t1 - t0).count();
}

// Primary packets exactly as RenderTask::Work builds them (src/render.cpp:61-62,112-115) but
// stopping after geometry.TraversePrimary (src/scene_trace.cpp:119-120): per-pixel nearest hit.
// Misses: triId = -1, t = +inf.
void ref_trace_primary(void *h, const float *cam, int W, int H, int *triId, float *t, float *bary,
					   unsigned *stats) {
	RefScene *rs = (RefScene *)h;
	Camera camera = MakeCamera(cam);
	RayGenerator rayGen(3, W, H, camera.plane_dist, camera.right, camera.up, camera.front);
	Vec3q origin;
	Broadcast(camera.pos, origin);
	TreeStats total;

	for(int y = 0; y < H; y += 16)
		for(int x = 0; x < W; x += 16) {
			Vec3q dir[64], idir[64];
			rayGen.Generate(16, 16, x, y, dir);
			for(int n = 0; n < 64; n++)
				idir[n] = SafeInv(dir[n]);

			floatq dist[64];
			i32x4 object[64], element[64];
			Vec2q bar[64];
			for(int q = 0; q < 64; q++) {
				dist[q] = floatq(constant::inf);
				object[q] = i32x4(0);
				bar[q] = Vec2q(floatq(0.0f), floatq(0.0f));
			}
			TreeStats st;
			RayGroup<1, 0> rays(&origin, dir, idir, 64);
			Context<1, 0> ctx(rays, dist, object, element, bar, &st);
			rs->scene.geometry.TraversePrimary(ctx);
			total += st;

			for(int q = 0; q < 64; q++) {
				int py = y + q / 4;
				if(py >= H)
					continue;
				for(int k = 0; k < 4; k++) {
					int px = x + (q & 3) * 4 + k;
					if(px >= W)
						continue;
					size_t o = (size_t)py * W + px;
					bool hit = dist[q][k] < constant::inf;
					triId[o] = hit ? object[q][k] : -1;
					t[o] = dist[q][k];
					bary[o * 2 + 0] = hit ? bar[q].x[k] : 0.0f;
					bary[o * 2 + 1] = hit ? bar[q].y[k] : 0.0f;
				}
			}
		}
	CopyStats(total, stats);
}

namespace {
// Packs up to 256 AoS rays [first, first+count) into 64 quads; tail lanes replicate the last ray
// and are masked off.
struct Packet {
	Vec3q origin[64], dir[64], idir[64];
	char mask[64 + 4];
	int count;

	void Load(const float *o3, const float *d3, const unsigned char *rmask, int first, int count_,
			  bool sharedOrigin) {
		count = count_;
		for(int q = 0; q < 64; q++) {
			mask[q] = 0;
			for(int k = 0; k < 4; k++) {
				int i = q * 4 + k, src = first + (i < count ? i : count - 1);
				const float *o = sharedOrigin ? o3 : o3 + (size_t)src * 3;
				const float *d = d3 + (size_t)src * 3;
				origin[q].x[k] = o[0]; origin[q].y[k] = o[1]; origin[q].z[k] = o[2];
				dir[q].x[k] = d[0]; dir[q].y[k] = d[1]; dir[q].z[k] = d[2];
				if(i < count && (!rmask || rmask[src]))
					mask[q] |= 1 << k;
			}
			idir[q] = SafeInv(dir[q]);
		}
	}
	bool All() const {
		for(int q = 0; q < 64; q++)
			if(mask[q] != 15)
				return false;
		return true;
	}
};
} // namespace

// Generic nearest-hit query = BVH::TraversePrimary(Context<so,hm>) (src/bvh/traverse.cpp:151-157)
// on consecutive 256-ray packets of caller rays (AoS xyz). mask may be null (all active).
void ref_trace_rays(void *h, const float *o3, const float *d3, const unsigned char *rmask, int n,
					int sharedOrigin, int *triId, float *t, float *bary) {
	RefScene *rs = (RefScene *)h;
	for(int first = 0; first < n; first += 256) {
		Packet p;
		p.Load(o3, d3, rmask, first, std::min(256, n - first), sharedOrigin);
		floatq dist[64];
		i32x4 object[64], element[64];
		Vec2q bar[64];
		for(int q = 0; q < 64; q++) {
			dist[q] = Condition(GetSSEMaskFromBits(p.mask[q]), floatq(constant::inf), floatq(-constant::inf));
			object[q] = i32x4(0);
			bar[q] = Vec2q(floatq(0.0f), floatq(0.0f));
		}
		if(sharedOrigin) {
			RayGroup<1, 0> rays(p.origin, p.dir, p.idir, 64);
			Context<1, 0> ctx(rays, dist, object, element, bar, 0);
			rs->scene.geometry.TraversePrimary(ctx);
		}
		else if(p.All()) {
			RayGroup<0, 0> rays(p.origin, p.dir, p.idir, 64);
			Context<0, 0> ctx(rays, dist, object, element, bar, 0);
			rs->scene.geometry.TraversePrimary(ctx);
		}
		else {
			RayGroup<0, 1> rays(p.origin, p.dir, p.idir, 64, p.mask);
			Context<0, 1> ctx(rays, dist, object, element, bar, 0);
			rs->scene.geometry.TraversePrimary(ctx);
		}
		for(int i = 0; i < p.count; i++) {
			int q = i >> 2, k = i & 3;
			bool hit = dist[q][k] < constant::inf && dist[q][k] > -constant::inf;
			triId[first + i] = hit ? object[q][k] : -1;
			t[first + i] = hit ? dist[q][k] : constant::inf;
			bary[(first + i) * 2 + 0] = hit ? bar[q].x[k] : 0.0f;
			bary[(first + i) * 2 + 1] = hit ? bar[q].y[k] : 0.0f;
		}
	}
}

// Any-hit query = BVH::TraverseShadow(ShadowContext) (src/bvh/traverse.cpp:82-149): tmax[i] < 0 on
// input = inactive ray; on output occluded rays hold -inf, others keep their tmax.
void ref_trace_shadow(void *h, const float *origin3, const float *d3, float *tmax, int n) {
	RefScene *rs = (RefScene *)h;
	for(int first = 0; first < n; first += 256) {
		Packet p;
		p.Load(origin3, d3, 0, first, std::min(256, n - first), true);
		floatq dist[64];
		for(int q = 0; q < 64; q++)
			for(int k = 0; k < 4; k++) {
				int i = q * 4 + k;
				dist[q][k] = i < p.count ? tmax[first + i] : -constant::inf;
			}
		ShadowContext ctx(RayGroup<1, 0>(p.origin, p.dir, p.idir, 64), dist, 0);
		rs->scene.geometry.TraverseShadow(ctx);
		for(int i = 0; i < p.count; i++)
			tmax[first + i] = dist[i >> 2][i & 3];
	}
}

// Scene<BVH>::RayTrace (src/scene_trace.cpp:92-520) on consecutive 256-ray packets; rgb3 receives
// outColor.{x,y,z} per ray.
void ref_raytrace(void *h, const float *o3, const float *d3, const unsigned char *rmask, int n,
				  int sharedOrigin, float *rgb3, unsigned *stats) {
	RefScene *rs = (RefScene *)h;
	TreeStats total;
	for(int first = 0; first < n; first += 256) {
		Packet p;
		p.Load(o3, d3, rmask, first, std::min(256, n - first), sharedOrigin);
		Vec3q colors[64];
		Cache cache;
		if(sharedOrigin)
			total += rs->scene.RayTrace(RayGroup<1, 0>(p.origin, p.dir, p.idir, 64), cache, colors);
		else if(p.All())
			total += rs->scene.RayTrace(RayGroup<0, 0>(p.origin, p.dir, p.idir, 64), cache, colors);
		else
			total += rs->scene.RayTrace(RayGroup<0, 1>(p.origin, p.dir, p.idir, 64, p.mask), cache, colors);
		for(int i = 0; i < p.count; i++) {
			int q = i >> 2, k = i & 3;
			rgb3[(first + i) * 3 + 0] = colors[q].x[k];
			rgb3[(first + i) * 3 + 1] = colors[q].y[k];
			rgb3[(first + i) * 3 + 2] = colors[q].z[k];
		}
	}
	CopyStats(total, stats);
}

// Primary ray directions as RayGenerator::Generate lays them out (src/ray_generator.cpp:23-47),
// scattered to pixel order; for checking the device ray generator.
void ref_primary_dirs(const float *cam, int W, int H, float *d3) {
	Camera camera = MakeCamera(cam);
	RayGenerator rayGen(3, W, H, camera.plane_dist, camera.right, camera.up, camera.front);
	for(int y = 0; y < H; y += 16)
		for(int x = 0; x < W; x += 16) {
			Vec3q dir[64];
			rayGen.Generate(16, 16, x, y, dir);
			for(int q = 0; q < 64; q++)
				for(int k = 0; k < 4; k++) {
					int px = x + (q & 3) * 4 + k, py = y + q / 4;
					if(px >= W || py >= H)
						continue;
					float *o = d3 + ((size_t)py * W + px) * 3;
					o[0] = dir[q].x[k]; o[1] = dir[q].y[k]; o[2] = dir[q].z[k];
				}
		}
}

const char *ref_backend() {
#ifdef VECLIB_SSE
	return "sse";
#else
	return "scalar";
#endif
}

// ---- instanced two-level BVH (DBVH, src/dbvh/tree.h) ---------------------------------------------------------------
// inst: n x {rotation[0..2] (3 x Vec3f), translation}: 12 floats each, the data members of ObjectInstance
// (src/dbvh/tree.h:177-178). Builds DBVH(instances) over this scene's BVH exactly as MakeDBVH does
// (ComputeBBox per instance, then the DBVH constructor) and routes ref_render / ref_trace_primary_inst to Scene<DBVH>.
void ref_scene_set_instances(void *h, const float *inst, int n) {
	RefScene *rs = (RefScene *)h;
	rs->instanced = n > 0;
	if(!n)
		return;
	vector<ObjectInstance> instances;
	for(int i = 0; i < n; i++, inst += 12) {
		ObjectInstance oi;
		oi.tree = &rs->scene.geometry;
		oi.rotation[0] = Vec3f(inst[0], inst[1], inst[2]);
		oi.rotation[1] = Vec3f(inst[3], inst[4], inst[5]);
		oi.rotation[2] = Vec3f(inst[6], inst[7], inst[8]);
		oi.translation = Vec3f(inst[9], inst[10], inst[11]);
		oi.ComputeBBox();
		instances.push_back(oi);
	}
	rs->dscene.geometry = DBVH(instances);
	rs->dscene.materials = rs->scene.materials;
	rs->dscene.lights = rs->scene.lights;
}

int ref_dbvh_counts(void *h, int *nNodes, int *nInstances) {
	RefScene *rs = (RefScene *)h;
	*nNodes = (int)rs->dscene.geometry.nodes.size();
	*nInstances = (int)rs->dscene.geometry.elements.size();
	return (int)sizeof(DBVH::Node);
}
const void *ref_dbvh_nodes(void *h) { return ((RefScene *)h)->dscene.geometry.nodes.data(); }
// elements in the DBVH's own order (std::partition moved them): out = n x {rotation 9, translation 3, bbox min 3, max 3}
void ref_dbvh_elements(void *h, float *out) {
	for(const ObjectInstance &e : ((RefScene *)h)->dscene.geometry.elements) {
		for(int r = 0; r < 3; r++) {
			*out++ = e.rotation[r].x;
			*out++ = e.rotation[r].y;
			*out++ = e.rotation[r].z;
		}
		*out++ = e.translation.x;
		*out++ = e.translation.y;
		*out++ = e.translation.z;
		*out++ = e.bbox.min.x;
		*out++ = e.bbox.min.y;
		*out++ = e.bbox.min.z;
		*out++ = e.bbox.max.x;
		*out++ = e.bbox.max.y;
		*out++ = e.bbox.max.z;
	}
}

// ref_trace_primary on the instanced scene: instId = Context::object (index into DBVH::elements), triId = element.
void ref_trace_primary_inst(void *h, const float *cam, int W, int H, int *instId, int *triId, float *t, float *bary) {
	RefScene *rs = (RefScene *)h;
	Camera camera = MakeCamera(cam);
	RayGenerator rayGen(3, W, H, camera.plane_dist, camera.right, camera.up, camera.front);
	Vec3q origin;
	Broadcast(camera.pos, origin);
	for(int y = 0; y < H; y += 16)
		for(int x = 0; x < W; x += 16) {
			Vec3q dir[64], idir[64];
			rayGen.Generate(16, 16, x, y, dir);
			for(int n = 0; n < 64; n++)
				idir[n] = SafeInv(dir[n]);
			floatq dist[64];
			i32x4 object[64], element[64];
			Vec2q bar[64];
			for(int q = 0; q < 64; q++) {
				dist[q] = floatq(constant::inf);
				object[q] = i32x4(0);
				element[q] = i32x4(0);
				bar[q] = Vec2q(floatq(0.0f), floatq(0.0f));
			}
			TreeStats st;
			RayGroup<1, 0> rays(&origin, dir, idir, 64);
			Context<1, 0> ctx(rays, dist, object, element, bar, &st);
			rs->dscene.geometry.TraversePrimary(ctx);
			for(int q = 0; q < 64; q++) {
				int py = y + q / 4;
				if(py >= H)
					continue;
				for(int k = 0; k < 4; k++) {
					int px = x + (q & 3) * 4 + k;
					if(px >= W)
						continue;
					size_t o = (size_t)py * W + px;
					bool hit = dist[q][k] < constant::inf;
					instId[o] = hit ? object[q][k] : -1;
					triId[o] = hit ? element[q][k] : -1;
					t[o] = dist[q][k];
					bary[o * 2 + 0] = hit ? bar[q].x[k] : 0.0f;
					bary[o * 2 + 1] = hit ? bar[q].y[k] : 0.0f;
				}
			}
		}
}

} // extern "C"
