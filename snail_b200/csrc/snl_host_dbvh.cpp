// snail_b200 — host side of the instanced two-level BVH. A PORT, not new work:
// ObjectInstance::ComputeBBox and DBVH::Construct / FindSplit (src/dbvh/tree.cpp:4-21,40-185) are restated function for
// function (same bins, same locals, the same std::partition on the same range) so that the top-level tree AND the order the
// builder leaves the instances in (std::partition) are the reference's: Context::object is an index into that order.
// Compiled with -ffp-contract=off like the rest of the host code.
#include <algorithm>
#include <cmath>
#include <limits>

#include "snl_host_scene.h"

namespace snl {

namespace {

struct Box {
	float mn[3], mx[3];
};
inline float fmin2(float a, float b) { return a < b ? a : b; }
inline float fmax2(float a, float b) { return a > b ? a : b; }
inline void box_add(Box &b, const Box &o) {
	for(int a = 0; a < 3; a++) {
		b.mn[a] = fmin2(b.mn[a], o.mn[a]);
		b.mx[a] = fmax2(b.mx[a], o.mx[a]);
	}
}
inline float box_sa(const Box &b) {  // BoxSA, src/dbvh/tree.cpp:35-37
	const float w = b.mx[0] - b.mn[0], h = b.mx[1] - b.mn[1], d = b.mx[2] - b.mn[2];
	return (w * (d + h) + d * h) * 2.0f;
}
inline int max_axis(float x, float y, float z) { return y > x ? (z > y ? 2 : 1) : (z > x ? 2 : 0); }
inline Box inst_box(const HostInstance &i) {
	Box b;
	for(int a = 0; a < 3; a++) {
		b.mn[a] = i.bmin[a];
		b.mx[a] = i.bmax[a];
	}
	return b;
}
inline float center(const HostInstance &i, int axis) { return (i.bmin[axis] + i.bmax[axis]) * 0.5f; }  // ObjectInstance::Center

struct Builder {
	std::vector<HostInstance> &elems;
	std::vector<RefNode> &nodes;
	int depth = 0;

	void push(const Box &b) {
		RefNode n;
		for(int a = 0; a < 3; a++) {
			n.bmin[a] = b.mn[a];
			n.bmax[a] = b.mx[a];
		}
		n.sub = 0;
		n.count = 0;
		nodes.push_back(n);
	}

	void split(int nNode, int first, int count, int sdepth) {  // DBVH::FindSplit
		Box bbox;
		for(int a = 0; a < 3; a++) {
			bbox.mn[a] = nodes[nNode].bmin[a];
			bbox.mx[a] = nodes[nNode].bmax[a];
		}
		bool leaf = count <= 1;
		int minAxis = 0, minIdx = 1, nBins = 0;
		float mul = 0.0f, sub = 0.0f;
		const float inf = std::numeric_limits<float>::infinity();
		Box leftBoxes[16], rightBoxes[16];
		int leftCounts[16], rightCounts[16];
		if(!leaf) {
			minAxis = max_axis(bbox.mx[0] - bbox.mn[0], bbox.mx[1] - bbox.mn[1], bbox.mx[2] - bbox.mn[2]);
			const float traverseCost = 0.0f, intersectCost = 1.0f;
			nBins = count < 8 ? 8 : 16;
			Box bins[16];
			int binCount[16];
			for(int n = 0; n < nBins; n++) {
				binCount[n] = 0;
				for(int a = 0; a < 3; a++) {
					bins[n].mn[a] = inf;
					bins[n].mx[a] = -inf;
				}
			}
			mul = nBins * (1.0f - 0.0001f) / (bbox.mx[minAxis] - bbox.mn[minAxis]);  // constant::epsilon, veclib/vecbase.h:40
			sub = bbox.mn[minAxis];
			for(int n = 0; n < count; n++) {
				const HostInstance &e = elems[first + n];
				const float c = (e.bmax[minAxis] + e.bmin[minAxis]) * 0.5f;
				const int nBin = int((c - sub) * mul);
				binCount[nBin]++;
				box_add(bins[nBin], inst_box(e));
			}
			rightBoxes[nBins - 1] = bins[nBins - 1];
			rightCounts[nBins - 1] = binCount[nBins - 1];
			leftBoxes[0] = bins[0];
			leftCounts[0] = binCount[0];
			for(int n = 1; n < nBins; n++) {
				leftBoxes[n] = leftBoxes[n - 1];
				box_add(leftBoxes[n], bins[n]);
				leftCounts[n] = leftCounts[n - 1] + binCount[n];
			}
			for(int n = nBins - 2; n >= 0; n--) {
				rightBoxes[n] = rightBoxes[n + 1];
				box_add(rightBoxes[n], bins[n]);
				rightCounts[n] = rightCounts[n + 1] + binCount[n];
			}
			float minCost = inf;
			const float noSplitCost = intersectCost * count * box_sa(bbox);
			for(int n = 1; n < nBins; n++) {
				const float cost = (leftCounts[n - 1] ? box_sa(leftBoxes[n - 1]) * leftCounts[n - 1] : 0) +
								   (rightCounts[n] ? box_sa(rightBoxes[n]) * rightCounts[n] : 0);
				if(cost < minCost) {
					minCost = cost;
					minIdx = n;
				}
			}
			minCost = traverseCost + intersectCost * minCost;
			if(noSplitCost < minCost)
				leaf = true;
		}
		if(leaf) {
			depth = std::max(depth, sdepth);
			nodes[nNode].sub = (int32_t)(first | 0x80000000u);
			nodes[nNode].count = count;
			return;
		}
		// TestBoxes (src/dbvh/tree.cpp:24-33); pointer range like the reference's, so std::partition permutes alike
		const int pAxis = minAxis, pIdx = minIdx;
		const float pSub = sub, pMul = mul;
		std::partition(&elems[first], &elems[first] + count,
					   [=](const HostInstance &e) { return int((center(e, pAxis) - pSub) * pMul) < pIdx; });
		Box leftBox = leftBoxes[minIdx - 1], rightBox = rightBoxes[minIdx];
		int leftCount = leftCounts[minIdx - 1], rightCount = rightCounts[minIdx];
		if(leftCount == 0 || rightCount == 0) {
			minIdx = count / 2;
			leftBox = inst_box(elems[first]);
			rightBox = inst_box(elems[first + count - 1]);
			for(int n = 1; n < minIdx; n++)
				box_add(leftBox, inst_box(elems[first + n]));
			for(int n = minIdx; n < count; n++)
				box_add(rightBox, inst_box(elems[first + n]));
			leftCount = minIdx;
			rightCount = count - leftCount;
		}
		const int subNode = (int)nodes.size();
		nodes[nNode].sub = subNode;
		nodes[nNode].axis = (int16_t)minAxis;
		// (the reference assigns firstNode twice; the second assignment is the one that stays, src/dbvh/tree.cpp:150-152)
		nodes[nNode].firstNode = leftBox.mn[minAxis] == rightBox.mn[minAxis] ? (leftBox.mx[minAxis] < rightBox.mx[minAxis] ? 0 : 1) : 0;
		push(leftBox);
		push(rightBox);
		split(subNode + 0, first, leftCount, sdepth + 1);
		split(subNode + 1, first + leftCount, rightCount, sdepth + 1);
	}
};

}  // namespace

// ObjectInstance::ComputeBBox (src/dbvh/tree.cpp:4-21): the 8 corners of the tree's box, rotated, then translated.
void instance_compute_bbox(HostInstance &inst, const float treeMin[3], const float treeMax[3]) {
	const float inf = std::numeric_limits<float>::infinity();
	float mn[3] = {inf, inf, inf}, mx[3] = {-inf, -inf, -inf};
	const float ys[4] = {treeMin[1], treeMin[1], treeMax[1], treeMax[1]};
	const float zs[4] = {treeMin[2], treeMax[2], treeMin[2], treeMax[2]};
	for(int c = 0; c < 3; c++) {
		const float *r = inst.rot[c];
		for(int k = 0; k < 4; k++) {
			const float p0 = ys[k] * r[1] + zs[k] * r[2];
			const float p1 = treeMax[0] * r[0] + p0;
			const float p0b = p0 + treeMin[0] * r[0];
			const float lo = fmin2(p0b, p1), hi = fmax2(p0b, p1);
			mn[c] = fmin2(mn[c], lo);
			mx[c] = fmax2(mx[c], hi);
		}
	}
	for(int c = 0; c < 3; c++) {
		inst.bmin[c] = mn[c] + inst.tr[c];
		inst.bmax[c] = mx[c] + inst.tr[c];
	}
}

bool build_dbvh(std::vector<HostInstance> &elems, std::vector<RefNode> &nodes, int &depth, std::string &err) {
	nodes.clear();
	if(elems.empty()) {
		err = "no instances";
		return false;
	}
	nodes.reserve(elems.size() * 2);
	Box bbox = inst_box(elems[0]);
	for(size_t n = 1; n < elems.size(); n++)
		box_add(bbox, inst_box(elems[n]));
	Builder b{elems, nodes};
	b.push(bbox);
	b.split(0, 0, (int)elems.size(), 0);
	depth = b.depth;
	if(depth > 64) {  // DBVH::maxDepth
		err = "DBVH deeper than maxDepth";
		return false;
	}
	return true;
}

}  // namespace snl
