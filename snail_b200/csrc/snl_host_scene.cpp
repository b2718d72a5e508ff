// snail_b200 — host side of the path: scene ingest + SAH BVH builder (see snl_host_scene.h).
//
// The output has to be the SAME tree the reference builds (node order, triangle permutation, float bits), because
// parity tests compare it array-for-array with BVH::nodes/tris/shTris of the reference. So the arithmetic follows
// the reference's operation order (compile with -ffp-contract=off), and the same libstdc++ algorithms
// (std::sort / std::nth_element with the same comparator on the same sequence) are used where the reference's
// result depends on them.
#include "snl_host_scene.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>

namespace snl {
namespace {

struct V3 {
	float x, y, z;
};
struct V2 {
	float x, y;
};
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }               // veclib/vec3.h:92-94
inline V3 cross(V3 b, V3 c) { return {b.y * c.z - b.z * c.y, b.z * c.x - b.x * c.z, b.x * c.y - b.y * c.x}; }  // :101-106
inline float rsqrt_ref(float v) { return 1.0f / sqrtf(v); }  // veclib/vecbase.h:54-56

// BaseScene::IndexedTri (src/base_scene.h:40-44)
struct IndexedTri {
	int v[3], vt[3], vn[3], matId;
};

// BaseScene::Object after its constructor (src/base_scene.cpp:114-158): only referenced attributes are kept.
struct Object {
	std::vector<IndexedTri> tris;
	std::vector<V3> verts, normals;
	std::vector<V2> uvs;
};

// ---- BaseScene::LoadWavefrontObj (src/formats/wavefront_obj.cpp:66-200) ------------------------------------------
bool parse_obj(const char *path, Object &obj, std::map<std::string, int> &matNames, std::string &err) {
	std::ifstream is(path);
	if(!is) {
		err = std::string("Error while opening: ") + path;
		return false;
	}
	std::vector<V3> verts, normals;
	std::vector<V2> uvs;
	std::vector<IndexedTri> tris;
	matNames.clear();
	matNames[""] = 0;
	int lastMatId = 0;

	std::vector<char> lineBuf(2000), type(2000), name(2000);
	for(;;) {
		char *line = lineBuf.data();
		if(!is.getline(line, 2000))
			break;
		type[0] = 0;
		sscanf(line, "%s", type.data());

		if(strcmp(type.data(), "v") == 0) {
			V3 v = {0, 0, 0};
			sscanf(line, "%s %f %f %f", type.data(), &v.x, &v.y, &v.z);
			verts.push_back(v);
		}
		else if(strcmp(type.data(), "vt") == 0) {
			V2 uv = {0, 0};
			float w;
			sscanf(line, "%s %f %f %f", type.data(), &uv.x, &uv.y, &w);
			uvs.push_back(uv);
		}
		else if(strcmp(type.data(), "vn") == 0) {
			V3 n = {0, 0, 0};
			sscanf(line, "%s %f %f %f", type.data(), &n.x, &n.y, &n.z);
			normals.push_back(n);
		}
		else if(strcmp(type.data(), "f") == 0) {
			char *p[4];
			p[0] = strpbrk(line, " \t");
			if(!p[0]) {
				err = std::string("malformed face: ") + line;
				return false;
			}
			p[0]++;
			while(p[0][0] == ' ')
				p[0]++;
			p[1] = strpbrk(p[0], " \t");
			p[2] = p[1] ? strpbrk(p[1] + 1, " \t") : nullptr;
			if(!p[1] || !p[2]) {
				err = std::string("face with fewer than 3 vertices: ") + line;
				return false;
			}
			p[1]++;
			p[1][-1] = 0;
			p[2]++;
			p[2][-1] = 0;
			p[3] = strpbrk(p[2], " \t");

			IndexedTri tri;
			tri.matId = lastMatId;
			for(;;) {
				for(int k = 0; k < 3; k++) {
					tri.v[k] = atoi(p[k]);
					if(tri.v[k] < 0)
						tri.v[k] = (int)verts.size() + tri.v[k] + 1;
					tri.vt[k] = tri.vn[k] = 0;
					char *puv = strchr(p[k], '/');
					if(puv) {
						char *pnrm = strchr(puv + 1, '/');
						if(pnrm != puv + 1) {
							tri.vt[k] = atoi(puv + 1);
							if(tri.vt[k] < 0)
								tri.vt[k] = (int)uvs.size() + tri.vt[k];
						}
						if(pnrm && pnrm[1]) {
							tri.vn[k] = atoi(pnrm + 1);
							if(tri.vn[k] < 0)
								tri.vn[k] = (int)normals.size() + tri.vn[k];
						}
					}
					tri.v[k]--;
					tri.vt[k]--;
					tri.vn[k]--;
					if(tri.v[k] >= (int)verts.size() || tri.v[k] < 0) {
						err = "Wrong vertex index in: " + std::string(line);
						return false;
					}
					if(tri.vt[k] >= (int)uvs.size()) {
						err = "Wrong tex-coord index in: " + std::string(line);
						return false;
					}
					if(tri.vn[k] >= (int)normals.size())
						tri.vn[k] = -1;
				}
				tris.push_back(tri);
				// quads: second triangle is (v2, v3, v0)   wavefront_obj.cpp:161-166
				if(p[3] && strpbrk(p[3], "0123456789")) {
					p[1] = p[3];
					std::swap(p[0], p[2]);
					p[3] = nullptr;
					continue;
				}
				break;
			}
		}
		else if(strcmp(type.data(), "usemtl") == 0) {
			name[0] = 0;
			sscanf(line, "%s %s", type.data(), name.data());
			std::string matName = name.data();
			auto it = matNames.find(matName);
			if(it == matNames.end()) {
				lastMatId = (int)matNames.size();
				matNames[matName] = lastMatId;
			}
			else
				lastMatId = it->second;
		}
	}
	if(tris.empty()) {
		err = std::string("no faces in ") + path;
		return false;
	}

	// BaseScene::Object::Object (src/base_scene.cpp:114-158)
	std::vector<int> vIMap(verts.size()), uvIMap(uvs.size()), nrmIMap(normals.size());
	{
		std::vector<char> usage(std::max(std::max(verts.size(), uvs.size()), normals.size()), 0);
		for(const IndexedTri &t : tris)
			for(int k = 0; k < 3; k++) {
				usage[t.v[k]] |= 1;
				if(t.vt[k] >= 0)
					usage[t.vt[k]] |= 2;
				if(t.vn[k] >= 0)
					usage[t.vn[k]] |= 4;
			}
		for(size_t n = 0; n < usage.size(); n++) {
			if(usage[n] & 1) {
				vIMap[n] = (int)obj.verts.size();
				obj.verts.push_back(verts[n]);
			}
			if(usage[n] & 2) {
				uvIMap[n] = (int)obj.uvs.size();
				obj.uvs.push_back(uvs[n]);
			}
			if(usage[n] & 4) {
				nrmIMap[n] = (int)obj.normals.size();
				obj.normals.push_back(normals[n]);
			}
		}
	}
	obj.tris.resize(tris.size());
	for(size_t n = 0; n < tris.size(); n++) {
		const IndexedTri &src = tris[n];
		IndexedTri &dst = obj.tris[n];
		for(int k = 0; k < 3; k++) {
			dst.v[k] = vIMap[src.v[k]];
			dst.vt[k] = src.vt[k] >= 0 ? uvIMap[src.vt[k]] : -1;
			dst.vn[k] = src.vn[k] >= 0 ? nrmIMap[src.vn[k]] : -1;
		}
		dst.matId = src.matId;
	}

	// BaseScene::Object::Repair (src/base_scene.cpp:173-184): drop degenerate triangles (swap with last)
	for(int n = 0; n < (int)obj.tris.size(); n++) {
		const IndexedTri &t = obj.tris[n];
		const V3 v0 = obj.verts[t.v[0]], v1 = obj.verts[t.v[1]], v2 = obj.verts[t.v[2]];
		const V3 nrm = cross(v1 - v0, v2 - v0);
		if(fabsf(nrm.x) < 0.00000001f && fabsf(nrm.y) < 0.00000001f && fabsf(nrm.z) < 0.00000001f) {
			obj.tris[n] = obj.tris.back();
			obj.tris.pop_back();
			n--;
		}
	}
	if(obj.tris.empty()) {
		err = std::string("only degenerate faces in ") + path;
		return false;
	}
	return true;
}

// BaseScene::Object::FlipNormals (src/base_scene.cpp:326-335)
void flip_normals(Object &o) {
	for(IndexedTri &t : o.tris) {
		std::swap(t.v[0], t.v[1]);
		std::swap(t.vt[0], t.vt[1]);
		std::swap(t.vn[0], t.vn[1]);
	}
	for(V3 &n : o.normals)
		n = {-n.x, -n.y, -n.z};
}

// BaseScene::Object::GenNormals (src/base_scene.cpp:517-529)
void gen_normals(Object &o) {
	for(IndexedTri &t : o.tris) {
		V3 nrm = cross(o.verts[t.v[1]] - o.verts[t.v[0]], o.verts[t.v[2]] - o.verts[t.v[0]]);
		nrm = nrm * rsqrt_ref(dot(nrm, nrm));
		if(t.vn[0] < 0 || t.vn[1] < 0 || t.vn[2] < 0) {
			for(int k = 0; k < 3; k++)
				if(t.vn[k] < 0)
					t.vn[k] = (int)o.normals.size();
			o.normals.push_back(nrm);
		}
	}
}
