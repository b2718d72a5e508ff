// snail_b200 — host side of the path: scene ingest + SAH BVH builder (see snl_host_scene.h).
//
// THIS FILE IS A PORT, NOT NEW WORK: parse_obj restates BaseScene::LoadWavefrontObj (src/formats/wavefront_obj.cpp:66-200)
// and Builder::split restates BVH::FindSplitSweep (src/bvh/tree.cpp:51-159) function for function (same locals, same
// order of statements, the reference's `goto` turned into a loop); the material / mip-chain code follows
// src/shading/material.cpp:59-192 and src/mipmap_texture.cpp. Only the PNG / TGA decoders are original. It exists for ONE
// reason: a bit-identical tree, so that hit records and triangle ids can be compared with the reference and its goldens
// without the reference being present. It is host-side ingest, outside the CUDA hot path. A caller that has the
// reference's own builder hands its output to snl_scene_set_geometry instead (include/scene_b200.h does); the product's
// own from-scratch builder is the device-side binned-SAH build (snl_build_sah.cu, SNL_BUILD_DEVICE | SNL_BUILD_SAH_BINNED),
// whose trees render within 2.3 % of this one's (profiles/r2n_build_device_vs_host.jsonl).
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
#include <limits>
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

// ---- BaseScene::ToTriVector / ToShTriVector (src/base_scene.cpp:39-77) -------------------------------------------

// Triangle::Triangle + ComputeData (src/triangle.h:16-21,123-131)
RefTriangle make_triangle(V3 a, V3 b, V3 c) {
	RefTriangle t;
	memset(&t, 0, sizeof(t));
	const V3 ba = b - a, ca = c - a;
	V3 nrm = cross(ba, ca);
	const float len = sqrtf(dot(nrm, nrm));
	const float inv = 1.0f / len;  // Vec3::operator/= multiplies by Inv(s)  veclib/vec3.h:48-52
	nrm = nrm * inv;
	t.a[0] = a.x, t.a[1] = a.y, t.a[2] = a.z;
	t.ba[0] = ba.x, t.ba[1] = ba.y, t.ba[2] = ba.z;
	t.ca[0] = ca.x, t.ca[1] = ca.y, t.ca[2] = ca.z;
	t.t0 = len;
	t.it0 = 1.0f / t.t0;
	t.plane[0] = nrm.x, t.plane[1] = nrm.y, t.plane[2] = nrm.z;
	t.plane[3] = dot(nrm, a);
	return t;
}

void to_tri_vectors(const Object &o, bool shading, std::vector<RefTriangle> &tris, std::vector<RefShTri> &shTris) {
	tris.resize(o.tris.size());
	if(shading)
		shTris.resize(o.tris.size());
	for(size_t n = 0; n < o.tris.size(); n++) {
		const IndexedTri &src = o.tris[n];
		// BaseScene::Object::GetTriangle (src/base_scene.cpp:301-317)
		V3 pos[3];
		for(int k = 0; k < 3; k++)
			pos[k] = o.verts[src.v[k]];
		V3 fnrm = cross(pos[1] - pos[0], pos[2] - pos[0]);
		fnrm = fnrm * rsqrt_ref(dot(fnrm, fnrm));
		tris[n] = make_triangle(pos[0], pos[1], pos[2]);
		if(shading) {
			V2 uv[3];
			V3 nrm[3];
			for(int k = 0; k < 3; k++) {
				uv[k] = src.vt[k] >= 0 ? o.uvs[src.vt[k]] : V2{0.0f, 0.0f};
				nrm[k] = src.vn[k] >= 0 ? o.normals[src.vn[k]] : fnrm;
			}
			// ShTriangle::ShTriangle (src/triangle.h:187-204): vertex 1,2 attributes stored as deltas from vertex 0
			RefShTri &s = shTris[n];
			s.uv[0][0] = uv[0].x, s.uv[0][1] = uv[0].y;
			s.uv[1][0] = uv[1].x - uv[0].x, s.uv[1][1] = uv[1].y - uv[0].y;
			s.uv[2][0] = uv[2].x - uv[0].x, s.uv[2][1] = uv[2].y - uv[0].y;
			const V3 n1 = nrm[1] - nrm[0], n2 = nrm[2] - nrm[0];
			s.nrm[0][0] = nrm[0].x, s.nrm[0][1] = nrm[0].y, s.nrm[0][2] = nrm[0].z;
			s.nrm[1][0] = n1.x, s.nrm[1][1] = n1.y, s.nrm[1][2] = n1.z;
			s.nrm[2][0] = n2.x, s.nrm[2][1] = n2.y, s.nrm[2][2] = n2.z;
			s.mat = src.matId;
		}
	}
}

// ---- BVH::Construct / FindSplitSweep (src/bvh/tree.cpp:51-159,293-328) -------------------------------------------

struct Box {
	float mn[3], mx[3];
};
inline float fmin2(float a, float b) { return a < b ? a : b; }  // veclib/vecbase.h:75-76
inline float fmax2(float a, float b) { return a > b ? a : b; }

// Triangle::GetBBox (src/triangle.h:60-69): bounds of P1 = a, P2 = ba + a, P3 = ca + a
inline Box tri_box(const RefTriangle &t) {
	Box b;
	for(int a = 0; a < 3; a++) {
		const float p1 = t.a[a], p2 = t.ba[a] + t.a[a], p3 = t.ca[a] + t.a[a];
		b.mn[a] = fmin2(p1, fmin2(p2, p3));
		b.mx[a] = fmax2(p1, fmax2(p2, p3));
	}
	return b;
}
inline void box_add(Box &b, const Box &o) {
	for(int a = 0; a < 3; a++) {
		b.mn[a] = fmin2(b.mn[a], o.mn[a]);
		b.mx[a] = fmax2(b.mx[a], o.mx[a]);
	}
}
// BoxSA (src/bvh/tree.cpp:45-47): Width = x, Height = y, Depth = z extents (src/bounding_box.h:29-31)
inline float box_sa(const Box &b) {
	const float w = b.mx[0] - b.mn[0], h = b.mx[1] - b.mn[1], d = b.mx[2] - b.mn[2];
	return (w * (d + h) + d * h) * 2.0f;
}
inline int max_axis(float x, float y, float z) { return y > x ? (z > y ? 2 : 1) : (z > x ? 2 : 0); }  // src/rtbase.h:136-138

struct OrderTris {  // src/bvh/tree.cpp:8-22
	const std::vector<RefTriangle> &tris;
	int axis;
	bool operator()(int i1, int i2) const {
		const RefTriangle &t1 = tris[i1], &t2 = tris[i2];
		return t1.a[axis] * 3.0f + t1.ba[axis] + t1.ca[axis] < t2.a[axis] * 3.0f + t2.ba[axis] + t2.ca[axis];
	}
};

struct Builder {
	HostBVH &bvh;
	bool useSah;
	enum { maxDepth = 64 };

	void set_box(int node, const Box &b) {
		memcpy(bvh.nodes[node].bmin, b.mn, 12);
		memcpy(bvh.nodes[node].bmax, b.mx, 12);
	}
	Box get_box(int node) const {
		Box b;
		memcpy(b.mn, bvh.nodes[node].bmin, 12);
		memcpy(b.mx, bvh.nodes[node].bmax, 12);
		return b;
	}
	int push_node(const Box &b) {
		RefNode n;
		memset(&n, 0, sizeof(n));
		bvh.nodes.push_back(n);
		set_box((int)bvh.nodes.size() - 1, b);
		return (int)bvh.nodes.size() - 1;
	}

	void make_leaf(int nNode, int first, int count, int sdepth) {
		Box b = tri_box(bvh.tris[first]);
		for(int n = 1; n < count; n++)
			box_add(b, tri_box(bvh.tris[first + n]));
		set_box(nNode, b);
		bvh.depth = std::max(bvh.depth, sdepth);
		bvh.nodes[nNode].sub = (int32_t)(first | 0x80000000u);
		bvh.nodes[nNode].count = count;
	}

	void split(int nNode, int first, int count, int sdepth) {
		std::vector<RefTriangle> &tris = bvh.tris;
		const Box bbox = get_box(nNode);
		if(count <= 4 || sdepth == maxDepth - 1) {
			make_leaf(nNode, first, count, sdepth);
			return;
		}
		int minIdx = count / 2;
		int minAxis = max_axis(bbox.mx[0] - bbox.mn[0], bbox.mx[1] - bbox.mn[1], bbox.mx[2] - bbox.mn[2]);
		{
			std::vector<int> indices(count);
			if(useSah) {
				const float traverseCost = 0.0f, intersectCost = 1.0f;
				float minCost = std::numeric_limits<float>::infinity();
				const float noSplitCost = intersectCost * count * box_sa(bbox);
				std::vector<float> leftSA(count), rightSA(count);
				for(int axis = 0; axis <= 2; axis++) {
					for(int n = 0; n < count; n++)
						indices[n] = first + n;
					std::sort(indices.begin(), indices.begin() + count, OrderTris{tris, axis});
					rightSA[count - 1] = box_sa(tri_box(tris[indices[count - 1]]));
					leftSA[0] = box_sa(tri_box(tris[indices[0]]));
					Box last = tri_box(tris[indices[0]]);
					for(int n = 1; n < count; n++) {
						box_add(last, tri_box(tris[indices[n]]));
						leftSA[n] = box_sa(last);
					}
					last = tri_box(tris[indices[count - 1]]);
					for(int n = count - 2; n >= 0; n--) {
						box_add(last, tri_box(tris[indices[n]]));
						rightSA[n] = box_sa(last);
					}
					for(int n = 1; n < count; n++) {
						// size_t operands convert to float: leftSA[n-1] * float(n) + rightSA[n] * float(count - n)
						const float cost = leftSA[n - 1] * float(n) + rightSA[n] * float(count - n);
						if(cost < minCost) {
							minCost = cost;
							minIdx = n;
							minAxis = axis;
						}
					}
				}
				minCost = traverseCost + intersectCost * minCost;
				if(noSplitCost < minCost) {
					make_leaf(nNode, first, count, sdepth);
					return;
				}
			}
			if(!useSah || minAxis != 2) {
				for(int n = 0; n < count; n++)
					indices[n] = first + n;
				std::nth_element(indices.begin(), indices.begin() + minIdx, indices.end(), OrderTris{tris, minAxis});
			}
			std::vector<RefTriangle> ttemp(count);
			for(int n = 0; n < count; n++)
				ttemp[n] = tris[indices[n]];
			for(int n = 0; n < count; n++)
				tris[first + n] = ttemp[n];
			if(!bvh.shTris.empty()) {
				std::vector<RefShTri> stemp(count);
				for(int n = 0; n < count; n++)
					stemp[n] = bvh.shTris[indices[n]];
				for(int n = 0; n < count; n++)
					bvh.shTris[first + n] = stemp[n];
			}
		}
		Box leftBox = tri_box(tris[first]);
		Box rightBox = tri_box(tris[first + count - 1]);
		for(int n = 1; n < minIdx; n++)
			box_add(leftBox, tri_box(tris[first + n]));
		for(int n = minIdx; n < count; n++)
			box_add(rightBox, tri_box(tris[first + n]));

		const int subNode = (int)bvh.nodes.size();
		bvh.nodes[nNode].sub = subNode;
		bvh.nodes[nNode].axis = (int16_t)minAxis;
		// the second assignment of src/bvh/tree.cpp:149-151 is the one that survives
		bvh.nodes[nNode].firstNode =
			leftBox.mn[minAxis] == rightBox.mn[minAxis] ? (leftBox.mx[minAxis] < rightBox.mx[minAxis] ? 0 : 1) : 0;
		push_node(leftBox);
		push_node(rightBox);
		split(subNode + 0, first, minIdx, sdepth + 1);
		split(subNode + 1, first + minIdx, count - minIdx, sdepth + 1);
	}

	// BVH::FindSplit (src/bvh/tree.cpp:161-287), the fastBuild path: 16 centroid bins along the widest axis of the
	// node box, SAH over the 15 bin boundaries, std::partition of the triangle range. Leaves keep the box their parent
	// gave them (no refit), and there is no depth cap here.
	enum { nBins = 16 };
	static int bin_of(float centre, float sub, float mul) {
		const float f = (centre - sub) * mul;
		// a degenerate node box makes mul inf/NaN; the reference indexes out of bounds there (UB) — we clamp
		if(!(f >= 0.0f))
			return 0;
		return f >= float(nBins) ? nBins - 1 : int(f);
	}
	void split_binned(int nNode, int first, int count, int sdepth) {
		std::vector<RefTriangle> &tris = bvh.tris;
		const Box bbox = get_box(nNode);
		auto leaf = [&]() {
			bvh.depth = std::max(bvh.depth, sdepth);
			bvh.nodes[nNode].sub = (int32_t)(first | 0x80000000u);
			bvh.nodes[nNode].count = count;
		};
		if(count <= 4) {
			leaf();
			return;
		}
		const int axis = max_axis(bbox.mx[0] - bbox.mn[0], bbox.mx[1] - bbox.mn[1], bbox.mx[2] - bbox.mn[2]);
		const float inf = std::numeric_limits<float>::infinity();
		Box binBox[nBins];
		int binCount[nBins];
		for(int n = 0; n < nBins; n++) {
			binCount[n] = 0;
			for(int a = 0; a < 3; a++) {
				binBox[n].mn[a] = inf;
				binBox[n].mx[a] = -inf;
			}
		}
		const float mul = nBins * (1.0f - 0.0001f) / (bbox.mx[axis] - bbox.mn[axis]);  // constant::epsilon, veclib/vecbase.h:40
		const float sub = bbox.mn[axis];
		for(int n = 0; n < count; n++) {
			const Box b = tri_box(tris[first + n]);
			const int bin = bin_of((b.mx[axis] + b.mn[axis]) * 0.5f, sub, mul);
			binCount[bin]++;
			box_add(binBox[bin], b);
		}
		Box lBox[nBins], rBox[nBins];
		int lCnt[nBins], rCnt[nBins];
		lBox[0] = binBox[0];
		lCnt[0] = binCount[0];
		for(int n = 1; n < nBins; n++) {
			lBox[n] = lBox[n - 1];
			box_add(lBox[n], binBox[n]);
			lCnt[n] = lCnt[n - 1] + binCount[n];
		}
		rBox[nBins - 1] = binBox[nBins - 1];
		rCnt[nBins - 1] = binCount[nBins - 1];
		for(int n = nBins - 2; n >= 0; n--) {
			rBox[n] = rBox[n + 1];
			box_add(rBox[n], binBox[n]);
			rCnt[n] = rCnt[n + 1] + binCount[n];
		}
		float minCost = inf;
		const float noSplitCost = 1.0f * count * box_sa(bbox);
		int minIdx = 1;
		for(int n = 1; n < nBins; n++) {
			const float cost = (lCnt[n - 1] ? box_sa(lBox[n - 1]) * lCnt[n - 1] : 0) + (rCnt[n] ? box_sa(rBox[n]) * rCnt[n] : 0);
			if(cost < minCost) {
				minCost = cost;
				minIdx = n;
			}
		}
		minCost = 0.0f + 1.0f * minCost;
		if(noSplitCost < minCost) {
			leaf();
			return;
		}
		std::vector<int> indices(count);
		for(int n = 0; n < count; n++)
			indices[n] = first + n;
		// TestTris (src/bvh/tree.cpp:24-43): centre from the three vertices' own min/max
		std::partition(&indices[0], &indices[0] + count, [&](int idx) {
			const RefTriangle &t = tris[idx];
			const float f1 = t.a[axis], f2 = t.ba[axis] + t.a[axis], f3 = t.ca[axis] + t.a[axis];
			const float centre = (fmin2(f1, fmin2(f2, f3)) + fmax2(f1, fmax2(f2, f3))) * 0.5f;
			return int((centre - sub) * mul) < minIdx;
		});
		reorder(indices, first, count);
		Box leftBox = lBox[minIdx - 1], rightBox = rBox[minIdx];
		int leftCount = lCnt[minIdx - 1], rightCount = rCnt[minIdx];
		if(leftCount == 0 || rightCount == 0) {
			minIdx = count / 2;
			leftBox = tri_box(tris[first]);
			rightBox = tri_box(tris[first + count - 1]);
			for(int n = 1; n < minIdx; n++)
				box_add(leftBox, tri_box(tris[first + n]));
			for(int n = minIdx; n < count; n++)
				box_add(rightBox, tri_box(tris[first + n]));
			leftCount = minIdx;
			rightCount = count - leftCount;
		}
		const int subNode = (int)bvh.nodes.size();
		bvh.nodes[nNode].sub = subNode;
		bvh.nodes[nNode].axis = (int16_t)axis;
		bvh.nodes[nNode].firstNode =
			leftBox.mn[axis] == rightBox.mn[axis] ? (leftBox.mx[axis] < rightBox.mx[axis] ? 0 : 1) : 0;
		push_node(leftBox);
		push_node(rightBox);
		split_binned(subNode + 0, first, leftCount, sdepth + 1);
		split_binned(subNode + 1, first + leftCount, rightCount, sdepth + 1);
	}

	// applies a permutation of [first, first + count) to tris (and shTris)
	void reorder(const std::vector<int> &indices, int first, int count) {
		std::vector<RefTriangle> ttemp(count);
		for(int n = 0; n < count; n++)
			ttemp[n] = bvh.tris[indices[n]];
		for(int n = 0; n < count; n++)
			bvh.tris[first + n] = ttemp[n];
		if(!bvh.shTris.empty()) {
			std::vector<RefShTri> stemp(count);
			for(int n = 0; n < count; n++)
				stemp[n] = bvh.shTris[indices[n]];
			for(int n = 0; n < count; n++)
				bvh.shTris[first + n] = stemp[n];
		}
	}
};

bool build_bvh(const Object &obj, const std::map<std::string, int> &matNames, int flags, HostBVH &bvh, std::string &err) {
	bvh = HostBVH();
	to_tri_vectors(obj, !(flags & kNoShadingData), bvh.tris, bvh.shTris);
	if(bvh.tris.empty()) {
		err = "no triangles";
		return false;
	}
	if(!(flags & kDeviceBuild)) {  // kDeviceBuild: triangles stay in file order, nodes stay empty (snl_build.cu builds)
		bvh.nodes.reserve(bvh.tris.size() * 2);
		Box bbox = tri_box(bvh.tris[0]);
		for(size_t n = 1; n < bvh.tris.size(); n++)
			box_add(bbox, tri_box(bvh.tris[n]));
		Builder b{bvh, (flags & kUseSah) != 0};
		b.push_node(bbox);
		if(flags & kFastBuild)  // src/bvh/tree.cpp:311-314
			b.split_binned(0, 0, (int)bvh.tris.size(), 0);
		else
			b.split(0, 0, (int)bvh.tris.size(), 0);
		if(bvh.depth > Builder::maxDepth) {
			err = "BVH deeper than maxDepth";
			return false;
		}
	}
	// BVH::materials names, indexed by the OBJ's material id (src/bvh/tree.cpp:316-327)
	bvh.matNames.assign(matNames.size(), std::string());
	for(const auto &kv : matNames)
		if(kv.second >= 0 && kv.second < (int)bvh.matNames.size())
			bvh.matNames[kv.second] = kv.first;
	return true;
}

// ---- images: the reference loads textures through fwk::Image::load (stb_image) and keeps R,G,B ---------------------

struct Rgb8 {
	int w = 0, h = 0;
	std::vector<unsigned char> px;  // tightly packed R,G,B, top row first
};

bool read_file(const std::string &path, std::vector<unsigned char> &out) {
	FILE *f = fopen(path.c_str(), "rb");
	if(!f)
		return false;
	fseek(f, 0, SEEK_END);
	long n = ftell(f);
	fseek(f, 0, SEEK_SET);
	out.resize(n > 0 ? n : 0);
	bool ok = n <= 0 || fread(out.data(), 1, n, f) == (size_t)n;
	fclose(f);
	return ok;
}

}  // namespace
}  // namespace snl

#include <zlib.h>

namespace snl {
namespace {

inline unsigned be32(const unsigned char *p) { return (unsigned)p[0] << 24 | p[1] << 16 | p[2] << 8 | p[3]; }

// PNG: 8/16-bit, colour types 0/2/3/4/6, non-interlaced.
bool decode_png(const std::vector<unsigned char> &f, Rgb8 &img, std::string &err) {
	static const unsigned char sig[8] = {137, 80, 78, 71, 13, 10, 26, 10};
	if(f.size() < 33 || memcmp(f.data(), sig, 8)) {
		err = "not a PNG file";
		return false;
	}
	int w = 0, h = 0, depth = 0, ctype = 0, interlace = 0;
	std::vector<unsigned char> idat, plte;
	for(size_t pos = 8; pos + 12 <= f.size();) {
		const unsigned len = be32(&f[pos]);
		const unsigned char *type = &f[pos + 4], *data = &f[pos + 8];
		if(pos + 12 + len > f.size())
			break;
		if(!memcmp(type, "IHDR", 4) && len >= 13) {
			w = be32(data), h = be32(data + 4);
			depth = data[8], ctype = data[9], interlace = data[12];
		}
		else if(!memcmp(type, "PLTE", 4))
			plte.assign(data, data + len);
		else if(!memcmp(type, "IDAT", 4))
			idat.insert(idat.end(), data, data + len);
		else if(!memcmp(type, "IEND", 4))
			break;
		pos += 12 + len;
	}
	if(w <= 0 || h <= 0 || interlace || (depth != 8 && depth != 16) || (ctype == 3 && depth != 8)) {
		err = "unsupported PNG variant (need non-interlaced, 8 or 16 bit)";
		return false;
	}
	const int chans = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
	if(!chans) {
		err = "bad PNG colour type";
		return false;
	}
	const int bpp = chans * depth / 8;
	const size_t stride = (size_t)w * bpp;
	std::vector<unsigned char> raw((stride + 1) * h);
	uLongf rawLen = raw.size();
	if(uncompress(raw.data(), &rawLen, idat.data(), idat.size()) != Z_OK || rawLen != raw.size()) {
		err = "PNG inflate failed";
		return false;
	}
	std::vector<unsigned char> cur(stride), prev(stride, 0);
	img.w = w;
	img.h = h;
	img.px.resize((size_t)w * h * 3);
	for(int y = 0; y < h; y++) {
		const unsigned char *row = &raw[(stride + 1) * y];
		const int filter = row[0];
		for(size_t i = 0; i < stride; i++) {
			const int a = i >= (size_t)bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= (size_t)bpp ? prev[i - bpp] : 0;
			int v = row[1 + i];
			switch(filter) {
			case 1: v += a; break;
			case 2: v += b; break;
			case 3: v += (a + b) >> 1; break;
			case 4: {
				const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
				v += (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
				break;
			}
			default: break;
			}
			cur[i] = (unsigned char)v;
		}
		for(int x = 0; x < w; x++) {
			const unsigned char *s = &cur[(size_t)x * bpp];
			const int step = depth / 8;  // 16 bit: keep the high byte
			unsigned char r, g, b;
			if(ctype == 3) {
				const size_t pi = (size_t)s[0] * 3;
				r = pi + 2 < plte.size() ? plte[pi] : 0;
				g = pi + 2 < plte.size() ? plte[pi + 1] : 0;
				b = pi + 2 < plte.size() ? plte[pi + 2] : 0;
			}
			else if(chans <= 2)
				r = g = b = s[0];
			else
				r = s[0], g = s[step], b = s[2 * step];
			unsigned char *d = &img.px[((size_t)y * w + x) * 3];
			d[0] = r, d[1] = g, d[2] = b;
		}
		prev.swap(cur);
	}
	return true;
}

// TGA: true colour (type 2) and RLE true colour (type 10), 24/32 bit; bottom-up unless descriptor bit 5 is set.
bool decode_tga(const std::vector<unsigned char> &f, Rgb8 &img, std::string &err) {
	if(f.size() < 18) {
		err = "truncated TGA";
		return false;
	}
	const int idLen = f[0], cmapType = f[1], type = f[2], w = f[12] | f[13] << 8, h = f[14] | f[15] << 8, bits = f[16];
	const bool topDown = f[17] & 0x20;
	if(cmapType || (type != 2 && type != 10) || (bits != 24 && bits != 32) || w <= 0 || h <= 0) {
		err = "unsupported TGA variant (need 24/32-bit true colour)";
		return false;
	}
	const int bpp = bits / 8;
	size_t pos = 18 + idLen;
	std::vector<unsigned char> pix((size_t)w * h * bpp);
	if(type == 2) {
		if(pos + pix.size() > f.size()) {
			err = "truncated TGA";
			return false;
		}
		memcpy(pix.data(), &f[pos], pix.size());
	}
	else {
		size_t o = 0;
		while(o < pix.size() && pos < f.size()) {
			const int hdr = f[pos++], n = (hdr & 127) + 1;
			if(hdr & 128) {
				if(pos + bpp > f.size())
					break;
				for(int i = 0; i < n && o + bpp <= pix.size(); i++, o += bpp)
					memcpy(&pix[o], &f[pos], bpp);
				pos += bpp;
			}
			else {
				const size_t bytes = (size_t)n * bpp;
				if(pos + bytes > f.size() || o + bytes > pix.size())
					break;
				memcpy(&pix[o], &f[pos], bytes);
				pos += bytes;
				o += bytes;
			}
		}
	}
	img.w = w;
	img.h = h;
	img.px.resize((size_t)w * h * 3);
	for(int y = 0; y < h; y++) {
		const int sy = topDown ? y : h - 1 - y;
		for(int x = 0; x < w; x++) {
			const unsigned char *s = &pix[((size_t)sy * w + x) * bpp];
			unsigned char *d = &img.px[((size_t)y * w + x) * 3];
			d[0] = s[2], d[1] = s[1], d[2] = s[0];  // stored B,G,R
		}
	}
	return true;
}

// MipmapTexture(Image, rgb8, hasMips) + GenMips (src/mipmap_texture.cpp:27-51,103-127,256-285)
void make_mip_chain(const Rgb8 &img, HostTexture &t) {
	t.width = img.w;
	t.height = img.h;
	int lg = 0;
	for(int v = std::max(img.w, img.h); v >>= 1;)
		lg++;
	t.mips = std::min(32, lg + 1);
	auto mw = [&](int m) { return std::max(img.w >> m, 1); };
	auto mh = [&](int m) { return std::max(img.h >> m, 1); };
	size_t total = 0;
	std::vector<size_t> off(t.mips);
	for(int m = 0; m < t.mips; m++) {
		off[m] = total;
		total += (size_t)mw(m) * mh(m) * 3;
	}
	t.data.assign(total + 8, 0);  // slack: GenMip's 1-row case reads src[4 + i] (sic)
	memcpy(t.data.data(), img.px.data(), (size_t)img.w * img.h * 3);
	for(int m = 1; m < t.mips; m++) {
		const int srcW = mw(m - 1), srcH = mh(m - 1), dstW = mw(m), dstH = mh(m);
		const size_t srcPitch = (size_t)srcW * 3, dstPitch = (size_t)dstW * 3;
		unsigned char *src = &t.data[off[m - 1]], *dst = &t.data[off[m]];
		if(srcH == dstH) {
			for(int x = 0; x < dstW; x++) {
				for(int i = 0; i < 3; i++)
					dst[i] = (unsigned char)((unsigned(src[i]) + unsigned(src[4 + i])) / 2);
				dst += 3;
				src += 6;
			}
		}
		else if(srcW == dstW) {
			for(int y = 0; y < dstH; y++) {
				for(int i = 0; i < 3; i++)
					dst[i] = (unsigned char)((unsigned(src[i]) + unsigned(src[i + srcPitch])) / 2);
				src += srcPitch * 2;
				dst += dstPitch;
			}
		}
		else
			for(int y = 0; y < dstH; y++) {
				const unsigned char *s = src + (size_t)y * 2 * srcPitch;
				unsigned char *d = dst + (size_t)y * dstPitch;
				for(int x = 0; x < dstW; x++) {
					for(int i = 0; i < 3; i++)
						d[i] = (unsigned char)((unsigned(s[i]) + unsigned(s[3 + i]) + unsigned(s[i + srcPitch]) +
												unsigned(s[i + 3 + srcPitch])) / 4);
					d += 3;
					s += 6;
				}
			}
	}
	t.data.resize(total);
}

// ---- materials: LoadMaterialDescs / LoadTextures / MakeMaterials (src/shading/material.cpp:59-192) ----------------

struct MaterialDesc {  // src/shading/material.h:53-67, defaults of material.cpp:52-55
	float ambient[3] = {0, 0, 0}, diffuse[3] = {1, 1, 1}, specular[3] = {0, 0, 0};
	float dissolveFactor = 0.0f;
	std::string diffuseMap, dissolveMap, name;
};

void read_color(std::stringstream &ss, float out[3]) {  // ReadColor, material.cpp:33-50
	std::string x, y, z;
	ss >> x;
	out[0] = (float)atof(x.c_str());
	ss >> y >> z;
	if(y == "" && z == "")
		out[1] = out[2] = out[0];
	else {
		out[1] = (float)atof(y.c_str());
		out[2] = (float)atof(z.c_str());
	}
}

std::vector<MaterialDesc> load_material_descs(const std::string &fileName) {
	std::vector<MaterialDesc> mats;
	std::ifstream is(fileName.c_str());
	if(!is)
		return mats;
	MaterialDesc cur;
	while(is.good()) {
		std::string line;
		getline(is, line);
		if(!is.good())
			break;
		auto it = line.find('#');
		if(it != std::string::npos)
			line.resize(it);
		std::stringstream ss(line);
		std::string token;
		ss >> token;
		if(token == "newmtl") {
			if(cur.name != "") {
				mats.push_back(cur);
				cur = MaterialDesc();
			}
			ss >> cur.name;
		}
		else if(token == "Ka")
			read_color(ss, cur.ambient);
		else if(token == "Kd")
			read_color(ss, cur.diffuse);
		else if(token == "Ks")
			read_color(ss, cur.specular);
		else if(token == "d") {
			std::string t;
			ss >> t;
			cur.dissolveFactor = (float)atof(t.c_str());
		}
		else if(token == "map_Kd")
			ss >> cur.diffuseMap;
		else if(token == "map_d")
			ss >> cur.dissolveMap;
	}
	if(cur.name != "")
		mats.push_back(cur);
	return mats;
}

// LoadTex (material.cpp:117-139): missing / unsupported files are skipped, the material then falls back to Uber.
bool load_texture(const std::string &name, const std::string &dir, HostTexture &out) {
	std::string path = dir.empty() ? name : (dir.back() == '/' ? dir + name : dir + "/" + name);
	std::string ext;
	auto dotPos = name.rfind('.');
	if(dotPos != std::string::npos)
		ext = name.substr(dotPos + 1);
	for(auto &c : ext)
		c = (char)tolower(c);
	if(ext != "tga" && ext != "png")  // (the reference also accepts bmp)
		return false;
	std::vector<unsigned char> file;
	if(!read_file(path, file))
		return false;
	Rgb8 img;
	std::string err;
	if(!(ext == "png" ? decode_png(file, img, err) : decode_tga(file, img, err)))
		return false;
	if((img.w & (img.w - 1)) || (img.h & (img.h - 1)))  // PointSampler: FATAL otherwise (point_sampler.cpp:7-8)
		return false;
	out.name = name;
	make_mip_chain(img, out);
	return true;
}

}  // namespace

bool load_obj_scene(const char *objPath, const char *mtlPath, const char *texDir, int flipNormalsFlag, int buildFlags,
					HostScene &out, std::string &err) {
	Object obj;
	std::map<std::string, int> matNames;
	if(!parse_obj(objPath, obj, matNames, err))
		return false;
	if(!flipNormalsFlag)  // src/rtracer.cpp:554-556
		flip_normals(obj);
	gen_normals(obj);
	if(!build_bvh(obj, matNames, buildFlags, out.bvh, err))
		return false;

	out.mats.clear();
	out.texs.clear();
	std::vector<MaterialDesc> descs;
	if(mtlPath && mtlPath[0])
		descs = load_material_descs(mtlPath);
	std::map<std::string, int> texIndex;  // TexDict
	std::map<std::string, HostMaterial> matDict;
	for(const MaterialDesc &d : descs) {
		for(const std::string *mapName : {&d.dissolveMap, &d.diffuseMap})
			if(!mapName->empty() && !texIndex.count(*mapName)) {
				HostTexture t;
				if(load_texture(*mapName, texDir ? texDir : "", t)) {
					texIndex[*mapName] = (int)out.texs.size();
					out.texs.push_back(std::move(t));
				}
			}
	}
	for(const MaterialDesc &d : descs) {  // MakeMaterials, material.cpp:157-183
		HostMaterial hm;
		hm.name = d.name;
		memset(&hm.m, 0, sizeof(hm.m));
		hm.m.texture = hm.m.mask_texture = -1;
		const bool diff = texIndex.count(d.diffuseMap) != 0, trans = texIndex.count(d.dissolveMap) != 0;
		if(diff && trans) {
			hm.m.kind = SNL_MAT_TRANSPARENT;
			hm.m.flags = SNL_MATF_TEXCOORDS | SNL_MATF_TRANSPARENCY;
			hm.m.texture = texIndex[d.diffuseMap];
			hm.m.mask_texture = texIndex[d.dissolveMap];
		}
		else if(diff) {
			hm.m.kind = SNL_MAT_TEX;
			hm.m.flags = SNL_MATF_TEXCOORDS;
			hm.m.texture = texIndex[d.diffuseMap];
		}
		else {  // UberMaterial::UberMaterial (src/shading/uber_material.cpp:5-12)
			hm.m.kind = SNL_MAT_UBER;
			hm.m.flags = d.dissolveFactor > 0 ? SNL_MATF_TRANSPARENCY : 0;
			hm.m.color[0] = d.diffuse[2];
			hm.m.color[1] = d.diffuse[1];
			hm.m.color[2] = d.diffuse[0];
			memcpy(hm.m.specular, d.specular, 12);
			hm.m.opacity = d.dissolveFactor;
		}
		matDict[d.name] = hm;
	}
	// Scene::UpdateMaterials (src/scene.cpp:12-21): ids follow std::map order of the names
	std::map<std::string, int> idOf;
	for(auto &kv : matDict) {
		idOf[kv.first] = (int)out.mats.size();
		out.mats.push_back(kv.second);
	}
	// BVH::UpdateMaterialIds (src/bvh/tree.cpp:376-386)
	HostBVH &bvh = out.bvh;
	bvh.triMatId.assign(bvh.matNames.size(), -1);
	for(size_t n = 0; n < bvh.matNames.size(); n++) {
		auto it = idOf.find(bvh.matNames[n]);
		bvh.triMatId[n] = it == idOf.end() ? -1 : it->second;
	}
	if(bvh.triMatId.empty())
		bvh.triMatId.push_back(-1);
	return true;
}

}  // namespace snl
