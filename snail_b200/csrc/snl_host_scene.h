// snail_b200 — host side of the path: scene ingest + SAH BVH builder. Produces the reference's own layouts
// (BVH::nodes / tris / shTris / materials, src/bvh/tree.h:86-92), which the flatten step then repacks for the GPU.
//   load_obj        BaseScene::LoadWavefrontObj + Object ctor + Repair   src/formats/wavefront_obj.cpp:66-200,
//                                                                        src/base_scene.cpp:114-184
//   flip/gen normals BaseScene::FlipNormals / GenNormals                 src/base_scene.cpp:326-335,517-529
//   build_bvh       BVH::Construct + FindSplitSweep                      src/bvh/tree.cpp:51-159,293-328
//   load_materials  LoadMaterialDescs / LoadTextures / MakeMaterials     src/shading/material.cpp:59-192
#pragma once

#include <string>
#include <vector>

#include "../../include/snail_b200.h"
#include "snl_types.h"

namespace snl {

struct HostBVH {
	std::vector<RefNode> nodes;
	std::vector<RefTriangle> tris;
	std::vector<RefShTri> shTris;       // empty with noShadingData
	std::vector<int> triMatId;          // BVH::materials[i].id
	std::vector<std::string> matNames;  // BVH::materials[i].name
	int depth = 0;
};

struct HostTexture {
	std::string name;
	std::vector<unsigned char> data;  // RGB8 mip chain, MipmapTexture layout
	int width = 0, height = 0, mips = 0;
};

struct HostMaterial {
	std::string name;
	snl_material m;
};

struct HostScene {
	HostBVH bvh;
	std::vector<HostMaterial> mats;  // Scene::materials order (std::map order of the names, src/scene.cpp:12-21)
	std::vector<HostTexture> texs;
};

enum { kNoShadingData = 1, kFastBuild = 2, kUseSah = 4 };  // src/bvh/tree.h:21-25
enum { kDeviceBuild = 8 };  // ours: leave HostBVH::nodes empty, the tree is built on the GPU (snl_build.cu)

// ObjectInstance (src/dbvh/tree.h:7-187): rotation rows, translation, world box.
struct HostInstance {
	float rot[3][3];
	float tr[3];
	float bmin[3], bmax[3];
};
// ObjectInstance::ComputeBBox + DBVH::Construct (src/dbvh/tree.cpp:4-21,159-173), snl_host_dbvh.cpp. build_dbvh reorders
// `elems` exactly as the reference's std::partition does and emits DBVH::Node records (same 32-byte layout as BVH::Node).
void instance_compute_bbox(HostInstance &inst, const float treeMin[3], const float treeMax[3]);
bool build_dbvh(std::vector<HostInstance> &elems, std::vector<RefNode> &nodes, int &depth, std::string &err);

// Returns false and fills err on failure.
bool load_obj_scene(const char *objPath, const char *mtlPath, const char *texDir, int flipNormalsFlag, int buildFlags,
					HostScene &out, std::string &err);

}  // namespace snl
