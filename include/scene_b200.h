// scene_b200.h — the reference-side adapter: a `Scene<>` whose RayTrace runs on the B200 through libsnail_b200.so.
//
// Header-only C++; compile it INSIDE the reference (nadult/Snail) with the reference's own headers on the include path
// and link libsnail_b200.so. It is the seam SURVEY.md §8(b) names:
//
//     template <bool sharedOrigin, bool hasMask>
//     const TreeStats Scene<AccStruct>::RayTrace(const RayGroup<sharedOrigin, hasMask>&, Cache&, Vec3q *outColor) const
//                                                                                     (src/scene.h:44-46)
//
// `Scene<B200Acc>` is an explicit specialisation of the reference's Scene template, so everything written against
// `Scene<AccStruct>` - RenderTask::Work's packet loop (src/render.cpp:47-211), both Render() overloads
// (src/render.cpp:214-280) - compiles unchanged with `AccStruct = B200Acc` once render.cpp's templates are instantiated
// for it (oracle/ref_adapter.cpp does exactly that and tests/test_gpu_adapter.py compares the frames).
//
// Construction mirrors the set-up of src/rtracer.cpp:505-586: it takes the host `Scene<BVH>` (BVH::nodes / tris / shTris
// / materials, src/bvh/tree.h:60-92; lights; ambient) plus the MaterialDescs its materials were made from
// (shading::MakeMaterials, src/shading/material.cpp:157-183 - the material classes keep their parameters private, the
// descs are the public form) and uploads all of it once. RayTrace converts the packet's SoA Vec3q arrays to the ABI's
// AoS rays, calls snl_raytrace and converts the colours back; the global gVals[] and the scene's lights are re-sent on
// every call, like the reference reads them on every call. Calls are serialised per adapter (one device stream);
// batch entry points (RenderFrame = snl_render) are what a GPU wants - RayTrace exists for drop-in fidelity.
#pragma once

#include <cstring>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "bvh/tree.h"
#include "camera.h"
#include "mipmap_texture.h"
#include "scene.h"
#include "shading/material.h"

#include "snail_b200.h"

struct B200Acc {};  // tag: Scene<B200Acc> is the device-backed scene

template <> class Scene<B200Acc> {
  public:
	Scene(const Scene<BVH> &host, const vector<shading::MaterialDesc> &descs, int device = 0)
		: ambientLight(host.ambientLight), lights(host.lights) {
		handle = snl_scene_create(device);
		if(!handle)
			Fail("snl_scene_create");
		const BVH &bvh = host.geometry;
		static_assert(sizeof(BVH::Node) == 32 && sizeof(Triangle) == 64 && sizeof(ShTriangle) == 64, "reference record sizes");
		vector<int> matIds(bvh.materials.size());
		for(size_t n = 0; n < bvh.materials.size(); n++)
			matIds[n] = bvh.materials[n].id;
		Check(snl_scene_set_geometry(handle, bvh.nodes.data(), (int)bvh.nodes.size(), bvh.tris.data(),
									 bvh.shTris.empty() ? nullptr : bvh.shTris.data(), (int)bvh.tris.size(),
									 matIds.empty() ? nullptr : matIds.data(), (int)matIds.size()),
			  "snl_scene_set_geometry");

		// textures: MipmapTexture chains (src/mipmap_texture.h:34-57), one table entry per texDict name
		std::map<string, int> texIndex;
		vector<vector<unsigned char>> chains;
		vector<snl_texture> texs;
		for(const auto &kv : host.texDict) {
			const MipmapTexture &t = *kv.second;
			vector<unsigned char> chain;
			for(size_t m = 0; m < t.Mips(); m++)
				chain.insert(chain.end(), t.DataPointer(m), t.DataPointer(m) + t.Size(m));
			texIndex[kv.first] = (int)texs.size();
			chains.push_back(std::move(chain));
			texs.push_back(snl_texture{nullptr, chains.back().size(), (int)t.Width(), (int)t.Height(), (int)t.Mips()});
		}
		for(size_t n = 0; n < texs.size(); n++)
			texs[n].data = chains[n].data();

		// materials in Scene::materials order (UpdateMaterials, src/scene.cpp:12-21: matDict iteration = id), each described
		// by the LAST desc of its name, as MakeMaterials' `out[mat.name] = newMat` leaves it
		vector<snl_material> mats(host.materials.size());
		for(const auto &kv : host.matDict) {
			const shading::MaterialDesc *desc = nullptr;
			for(const auto &d : descs)
				if(d.name == kv.first)
					desc = &d;
			const int id = kv.second->id;
			if(id < 0 || id >= (int)mats.size() || !desc)
				Fail("material table does not match the MaterialDescs");
			snl_material &m = mats[id];
			memset(&m, 0, sizeof(m));
			m.flags = kv.second->flags;
			m.texture = m.mask_texture = -1;
			const auto diff = texIndex.find(desc->diffuseMap), trans = texIndex.find(desc->dissolveMap);
			const bool hasDiff = desc->diffuseMap != "" && diff != texIndex.end();
			const bool hasTrans = desc->dissolveMap != "" && trans != texIndex.end();
			if(hasDiff && hasTrans) {  // TransparentMaterial<1> (src/shading/transparent_material.h:15-41)
				m.kind = SNL_MAT_TRANSPARENT;
				m.texture = diff->second;
				m.mask_texture = trans->second;
			}
			else if(hasDiff) {  // NewMaterial(diff) = TexMaterial<true> (src/shading/tex_material.h:14-34)
				m.kind = SNL_MAT_TEX;
				m.texture = diff->second;
			}
			else {  // UberMaterial (src/shading/uber_material.cpp:5-12: diffuse x/z swapped)
				m.kind = SNL_MAT_UBER;
				m.color[0] = desc->diffuse.z;
				m.color[1] = desc->diffuse.y;
				m.color[2] = desc->diffuse.x;
				m.specular[0] = desc->specular.x;
				m.specular[1] = desc->specular.y;
				m.specular[2] = desc->specular.z;
				m.opacity = desc->dissolveFactor;
			}
		}
		Check(snl_scene_set_materials(handle, mats.empty() ? nullptr : mats.data(), (int)mats.size(),
									  texs.empty() ? nullptr : texs.data(), (int)texs.size()),
			  "snl_scene_set_materials");
	}
	~Scene() { snl_scene_destroy(handle); }
	Scene(const Scene &) = delete;
	void operator=(const Scene &) = delete;

	Vec3f ambientLight;
	vector<Light> lights;

	// Scene<AccStruct>::RayTrace (src/scene.h:44-46, src/scene_trace.cpp:92-520) for one packet of up to 64 quads.
	// A mask that happens to be full is traced like the unmasked group (the reference only builds RayGroup<0,1> when
	// !selector.All(), src/scene_trace.cpp:615-617, 630-633).
	template <bool sharedOrigin, bool hasMask>
	const TreeStats RayTrace(const RayGroup<sharedOrigin, hasMask> &rays, Cache &, Vec3q *outColor) const {
		const int quads = rays.Size(), n = quads * 4;
		std::vector<float> o(sharedOrigin ? 3 : (size_t)n * 3), d((size_t)n * 3), rgb((size_t)n * 3);
		std::vector<unsigned char> mask(hasMask ? n : 0);
		for(int q = 0; q < quads; q++)
			for(int k = 0; k < 4; k++) {
				const int i = q * 4 + k;
				const Vec3q &dir = rays.Dir(q);
				d[i * 3 + 0] = dir.x[k];
				d[i * 3 + 1] = dir.y[k];
				d[i * 3 + 2] = dir.z[k];
				if(!sharedOrigin) {
					const Vec3q &org = rays.Origin(q);
					o[i * 3 + 0] = org.x[k];
					o[i * 3 + 1] = org.y[k];
					o[i * 3 + 2] = org.z[k];
				}
				if(hasMask)
					mask[i] = (rays.Mask(q) >> k) & 1;
			}
		if(sharedOrigin) {
			const Vec3q &org = rays.Origin(0);
			o[0] = org.x[0];
			o[1] = org.y[0];
			o[2] = org.z[0];
		}
		snl_stats st;
		{
			std::lock_guard<std::mutex> lock(mutex);
			SyncState();
			Check(snl_raytrace(handle, o.data(), d.data(), hasMask ? mask.data() : nullptr, n, sharedOrigin ? 1 : 0, rgb.data(), &st),
				  "snl_raytrace");
		}
		for(int q = 0; q < quads; q++)
			for(int k = 0; k < 4; k++) {
				const int i = q * 4 + k;
				outColor[q].x[k] = rgb[i * 3 + 0];
				outColor[q].y[k] = rgb[i * 3 + 1];
				outColor[q].z[k] = rgb[i * 3 + 2];
			}
		TreeStats stats;
		stats.TracingRays((uint)st.rays);
		return stats;
	}

	// Render(scene, camera, image, options, threads) (src/render.h:16-23) as ONE device call: the batch entry point.
	TreeStats RenderFrame(const Camera &camera, MipmapTexture &image, uint rank = 0) const {
		snl_camera cam;
		memcpy(cam.pos, &camera.pos, 12);
		memcpy(cam.right, &camera.right, 12);
		memcpy(cam.up, &camera.up, 12);
		memcpy(cam.front, &camera.front, 12);
		cam.plane_dist = camera.plane_dist;
		snl_stats st;
		std::lock_guard<std::mutex> lock(mutex);
		SyncState();
		Check(snl_render(handle, &cam, (int)image.Width(), (int)image.Height(), nullptr, 0, image.DataPointer(), image.Pitch(),
						 (int)rank, &st),
			  "snl_render");
		TreeStats stats;
		stats.TracingRays((uint)st.rays);
		return stats;
	}

	snl_scene *Handle() const { return handle; }

  private:
	void SyncState() const {  // the reference reads the global gVals[] and Scene::lights on every call
		Check(snl_set_options(handle, gVals), "snl_set_options");
		static_assert(sizeof(Light) == sizeof(snl_light), "Light is 9 floats (src/light.h:5-17)");
		const float amb[3] = {ambientLight.x, ambientLight.y, ambientLight.z};
		Check(snl_scene_set_lights(handle, reinterpret_cast<const snl_light *>(lights.data()), (int)lights.size(), amb),
			  "snl_scene_set_lights");
	}
	static void Fail(const char *what) { throw std::runtime_error(std::string("SceneB200: ") + what + ": " + snl_last_error()); }
	static void Check(int code, const char *what) {
		if(code != SNL_OK)
			Fail(what);
	}

	snl_scene *handle = nullptr;
	mutable std::mutex mutex;
};

typedef Scene<B200Acc> SceneB200;
