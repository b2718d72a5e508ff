// snail_b200 — device-side binned-SAH BVH build (the quality half): the reference's top-down
// surface-area-heuristic builders (src/bvh/tree.cpp:51-159 full sweep, :161-287 16 bins; costs traverse 0 / intersect 1,
// :70-71; leaves of <= 4 triangles, :54) re-designed for the GPU as a LEVEL-SYNCHRONOUS build in which every step is a
// flat parallel-for over triangles or over the nodes of the current level:
//
//   per level   clear bins      [level nodes x 3 axes x 16 bins] x {count, box}
//               bin             per triangle: 3 x (1 + 6) atomics into its node's bins (centroid -> bin per axis)
//               split           per node: 3 x 15 candidate planes from prefix / suffix bin boxes, SA(L) nL + SA(R) nR
//                               against n SA(node); leaf if <= 4 triangles or (<= 32 and no plane is cheaper)
//               scan + alloc    children of the splitting nodes get consecutive ids in the next level (deterministic)
//               assign          per triangle: side of the plane -> child id; atomics build the child's box, centroid
//                               bounds and count for the next level
//   at the end  radix sort of the triangles by leaf id (leaf = contiguous range), scan of the inner nodes, emit GpuNode
//
// There is no recursion, no per-node thread block and no partitioning of index arrays: a triangle just carries the id of
// the node it is in. Box / count atomics are min / max / add on integers, so the result does not depend on their order:
// the SAME functors run as plain loops on the host (snl_host_build_binned_sah), which is how the algorithm is pinned on a
// machine without a GPU (tests/test_build_sah_cpu.py renders the host-built tree with the oracle against the goldens),
// and the device build produces the same tree (tests/test_gpu_build.py).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <cstring>
#include <numeric>
#include <vector>

#include "../../include/snail_b200.h"
#include "snl_kernels.h"
#include "snl_scene_dev.h"

namespace snl {

namespace {

enum { kBins = 16, kBinInts = 7 };  // per bin: count, lo xyz, hi xyz (ordered ints)

#if defined(__CUDA_ARCH__)
#define SNL_ATOMIC_MIN(p, v) atomicMin((p), (v))
#define SNL_ATOMIC_MAX(p, v) atomicMax((p), (v))
#define SNL_ATOMIC_ADD(p, v) atomicAdd((p), (v))
#else
#define SNL_ATOMIC_MIN(p, v) (*(p) = *(p) < (v) ? *(p) : (v))
#define SNL_ATOMIC_MAX(p, v) (*(p) > (v) ? *(p) : (*(p) = (v)))
static inline int host_fetch_add(int *p, int v) {
	const int old = *p;
	*p = old + v;
	return old;
}
#define SNL_ATOMIC_ADD(p, v) host_fetch_add((p), (v))
#endif

#define SNL_HD __host__ __device__ __forceinline__

// float <-> int with the same ordering, for min / max on floats as integers
SNL_HD int f2o(float f) {
	int i;
	memcpy(&i, &f, 4);
	return i >= 0 ? i : i ^ 0x7fffffff;
}
SNL_HD float o2f(int i) {
	i = i >= 0 ? i : i ^ 0x7fffffff;
	float f;
	memcpy(&f, &i, 4);
	return f;
}
#define SNL_ORD_POS_INF 0x7f800000                // f2o(+inf)
#define SNL_ORD_NEG_INF ((int)0x807fffff)         // f2o(-inf) = 0xff800000 ^ 0x7fffffff

struct SahState {
	int n, cap;  // triangles, node capacity (2 n)
	const GenTri *tris;
	float *triLo, *triHi;  // [3 n] triangle boxes
	int *triNode;          // [n] node the triangle is in
	int *boxLo, *boxHi;    // [3 cap] node boxes (ordered ints)
	int *cenLo, *cenHi;    // [3 cap] bounds of the triangle-box centres in the node
	int *count;            // [cap]
	int *child;            // [cap] id of the first child (second = +1), -1 = leaf, -2 = not decided yet
	int *axis, *plane;     // [cap] split: axis 0..2 and the last bin that goes left; axis 3 = alternate (coincident centres)
	int *ticket;           // [cap] alternating assignment counter of axis-3 nodes
	int *flag, *scan;      // [level nodes] splits?, exclusive scan
	int *active, *binSlot; // [level nodes] has more than 4 triangles (is binned)?, its slot in `bins`
	int *bins;             // [binCap][3][kBins][kBinInts]: only nodes of > 4 triangles are binned (at most n / 5 per level)
	int binCap;
	int *leafFirst;        // [cap]
	int *inner, *innerIdx; // [cap] 1 = inner node, its index among the inner nodes
};

SNL_HD float sa_of(const float lo[3], const float hi[3]) {  // BoxSA, src/bvh/tree.cpp:40-42
	const float w = hi[0] - lo[0], h = hi[1] - lo[1], d = hi[2] - lo[2];
	return (w * (d + h) + d * h) * 2.0f;
}

SNL_HD int bin_of(float c, float lo, float hi) {
	const float ext = hi - lo;
	if(!(ext > 0.0f))
		return 0;
	int k = (int)((c - lo) * ((float)kBins / ext));
	return k < 0 ? 0 : k > kBins - 1 ? kBins - 1 : k;
}

SNL_HD void node_clear(const SahState &S, int node) {
	for(int a = 0; a < 3; a++) {
		S.boxLo[node * 3 + a] = S.cenLo[node * 3 + a] = SNL_ORD_POS_INF;
		S.boxHi[node * 3 + a] = S.cenHi[node * 3 + a] = SNL_ORD_NEG_INF;
	}
	S.count[node] = 0;
	S.child[node] = -2;
	S.ticket[node] = 0;
}

SNL_HD void node_add(const SahState &S, int node, int tri) {
	for(int a = 0; a < 3; a++) {
		const float lo = S.triLo[tri * 3 + a], hi = S.triHi[tri * 3 + a];
		const int c = f2o((lo + hi) * 0.5f);
		SNL_ATOMIC_MIN(&S.boxLo[node * 3 + a], f2o(lo));
		SNL_ATOMIC_MAX(&S.boxHi[node * 3 + a], f2o(hi));
		SNL_ATOMIC_MIN(&S.cenLo[node * 3 + a], c);
		SNL_ATOMIC_MAX(&S.cenHi[node * 3 + a], c);
	}
	SNL_ATOMIC_ADD(&S.count[node], 1);
}

// ---- the parallel-for bodies ----------------------------------------------------------------------------------------
struct FAllocRoot {
	SahState S;
	SNL_HD void operator()(int) const { node_clear(S, 0); }
};

struct FInitTri {  // i < n: Triangle::GetBBox (a, a + ba, a + ca), everything starts in node 0
	SahState S;
	SNL_HD void operator()(int i) const {
		const GenTri &t = S.tris[i];
		const float p[3][3] = {{t.a.x, t.a.y, t.a.z}, {t.a.x + t.ba.x, t.a.y + t.ba.y, t.a.z + t.ba.z}, {t.a.x + t.ca.x, t.a.y + t.ca.y, t.a.z + t.ca.z}};
		for(int a = 0; a < 3; a++) {
			const float lo = p[0][a] < p[1][a] ? (p[0][a] < p[2][a] ? p[0][a] : p[2][a]) : (p[1][a] < p[2][a] ? p[1][a] : p[2][a]);
			const float hi = p[0][a] > p[1][a] ? (p[0][a] > p[2][a] ? p[0][a] : p[2][a]) : (p[1][a] > p[2][a] ? p[1][a] : p[2][a]);
			S.triLo[i * 3 + a] = lo;
			S.triHi[i * 3 + a] = hi;
		}
		S.triNode[i] = 0;
		node_add(S, 0, i);
	}
};

struct FActive {  // slot < level nodes
	SahState S;
	int levelFirst;
	SNL_HD void operator()(int slot) const { S.active[slot] = S.count[levelFirst + slot] > 4; }
};

struct FClearBins {  // j < min(level nodes, binCap) * 3 * kBins
	SahState S;
	SNL_HD void operator()(int j) const {
		int *b = S.bins + (size_t)j * kBinInts;
		b[0] = 0;
		b[1] = b[2] = b[3] = SNL_ORD_POS_INF;
		b[4] = b[5] = b[6] = SNL_ORD_NEG_INF;
	}
};

struct FBin {  // i < n
	SahState S;
	int levelFirst, levelEnd;
	SNL_HD void operator()(int i) const {
		const int node = S.triNode[i];
		if(node < levelFirst || node >= levelEnd || S.count[node] <= 4)
			return;
		const float *lo = S.triLo + i * 3, *hi = S.triHi + i * 3;
		for(int a = 0; a < 3; a++) {
			const int k = bin_of((lo[a] + hi[a]) * 0.5f, o2f(S.cenLo[node * 3 + a]), o2f(S.cenHi[node * 3 + a]));
			int *b = S.bins + ((size_t)(S.binSlot[node - levelFirst] * 3 + a) * kBins + k) * kBinInts;
			SNL_ATOMIC_ADD(&b[0], 1);
			for(int c = 0; c < 3; c++) {
				SNL_ATOMIC_MIN(&b[1 + c], f2o(lo[c]));
				SNL_ATOMIC_MAX(&b[4 + c], f2o(hi[c]));
			}
		}
	}
};

struct FSplit {  // slot < level nodes
	SahState S;
	int levelFirst, levelEnd, forceLeaves;
	SNL_HD void operator()(int slot) const {
		const int node = levelFirst + slot, cnt = S.count[node];
		S.flag[slot] = 0;
		S.child[node] = -1;
		if(cnt <= 4 || (forceLeaves && cnt <= kLeafMaxCount))  // `count <= 4` rule of the reference (src/bvh/tree.cpp:54)
			return;
		float nlo[3], nhi[3];
		for(int a = 0; a < 3; a++) {
			nlo[a] = o2f(S.boxLo[node * 3 + a]);
			nhi[a] = o2f(S.boxHi[node * 3 + a]);
		}
		const float noSplit = (float)cnt * sa_of(nlo, nhi);  // intersectCost 1, traverseCost 0 (src/bvh/tree.cpp:70-71)
		float best = 3.0e38f;
		int bestAxis = -1, bestPlane = 0;
		for(int a = 0; a < 3; a++) {
			const int *bins = S.bins + (size_t)(S.binSlot[slot] * 3 + a) * kBins * kBinInts;
			// suffix boxes / counts: right side of the plane after bin k = bins k+1 .. 15
			float rsa[kBins];
			int rcnt[kBins];
			float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
			int c = 0;
			for(int k = kBins - 1; k >= 1; k--) {
				const int *b = bins + k * kBinInts;
				if(b[0]) {
					for(int d = 0; d < 3; d++) {
						const float bl = o2f(b[1 + d]), bh = o2f(b[4 + d]);
						lo[d] = bl < lo[d] ? bl : lo[d];
						hi[d] = bh > hi[d] ? bh : hi[d];
					}
					c += b[0];
				}
				rsa[k - 1] = c ? sa_of(lo, hi) : 0.0f;
				rcnt[k - 1] = c;
			}
			for(int d = 0; d < 3; d++) {
				lo[d] = 3.0e38f;
				hi[d] = -3.0e38f;
			}
			c = 0;
			for(int k = 0; k < kBins - 1; k++) {  // plane after bin k
				const int *b = bins + k * kBinInts;
				if(b[0]) {
					for(int d = 0; d < 3; d++) {
						const float bl = o2f(b[1 + d]), bh = o2f(b[4 + d]);
						lo[d] = bl < lo[d] ? bl : lo[d];
						hi[d] = bh > hi[d] ? bh : hi[d];
					}
					c += b[0];
				}
				if(c == 0 || rcnt[k] == 0)
					continue;
				const float cost = sa_of(lo, hi) * (float)c + rsa[k] * (float)rcnt[k];
				if(cost < best) {
					best = cost;
					bestAxis = a;
					bestPlane = k;
				}
			}
		}
		if(bestAxis < 0) {  // every centre in one bin on every axis (coincident triangles): cannot be separated by a plane
			if(cnt <= kLeafMaxCount)
				return;
			bestAxis = 3;  // deal them out alternately, halves the count every level
		}
		else if(cnt <= kLeafMaxCount && !(best < noSplit))
			return;
		S.axis[node] = bestAxis;
		S.plane[node] = bestPlane;
		S.flag[slot] = 1;
	}
};

struct FAlloc {  // slot < level nodes
	SahState S;
	int levelFirst, levelEnd;
	SNL_HD void operator()(int slot) const {
		if(!S.flag[slot])
			return;
		const int node = levelFirst + slot, c = levelEnd + 2 * S.scan[slot];
		S.child[node] = c;
		node_clear(S, c);
		node_clear(S, c + 1);
	}
};

struct FAssign {  // i < n
	SahState S;
	int levelFirst, levelEnd;
	SNL_HD void operator()(int i) const {
		const int node = S.triNode[i];
		if(node < levelFirst || node >= levelEnd || S.child[node] < 0)
			return;
		const int a = S.axis[node];
		int side;
		if(a == 3)
			side = SNL_ATOMIC_ADD(&S.ticket[node], 1) & 1;
		else {
			const float c = (S.triLo[i * 3 + a] + S.triHi[i * 3 + a]) * 0.5f;
			side = bin_of(c, o2f(S.cenLo[node * 3 + a]), o2f(S.cenHi[node * 3 + a])) > S.plane[node];
		}
		const int dst = S.child[node] + side;
		S.triNode[i] = dst;
		node_add(S, dst, i);
	}
};

struct FMarkInner {  // node < nNodes
	SahState S;
	SNL_HD void operator()(int node) const { S.inner[node] = S.child[node] >= 0; }
};

struct FLeafFirst {  // i < n, over the triangles sorted by leaf id
	SahState S;
	const int *sortedLeaf;
	SNL_HD void operator()(int i) const {
		if(i == 0 || sortedLeaf[i] != sortedLeaf[i - 1])
			S.leafFirst[sortedLeaf[i]] = i;
	}
};

struct FPermute {  // i < n
	const unsigned *order;
	const GenTri *inTris;
	const RefShTri *inSh;
	GenTri *outTris;
	RefShTri *outSh;
	SNL_HD void operator()(int i) const {
		const unsigned src = order[i];
		outTris[i] = inTris[src];
		if(inSh)
			outSh[i] = inSh[src];
	}
};

struct FEmit {  // node < nNodes
	SahState S;
	GpuNode *out;
	SNL_HD static int leaf_link(int first, int count) { return ~(first | ((count - 1) << kLeafFirstBits)); }
	SNL_HD void operator()(int node) const {
		const bool rootLeaf = node == 0 && S.child[0] < 0;  // the whole scene is one leaf: node 0 = two links to it
		if(S.child[node] < 0 && !rootLeaf)
			return;
		GpuNode g;
		int link[2];
		float lo[2][3], hi[2][3];
		for(int k = 0; k < 2; k++) {
			const int c = rootLeaf ? 0 : S.child[node] + k;
			for(int a = 0; a < 3; a++) {
				lo[k][a] = o2f(S.boxLo[c * 3 + a]);
				hi[k][a] = o2f(S.boxHi[c * 3 + a]);
			}
			link[k] = (!rootLeaf && S.child[c] >= 0) ? S.innerIdx[c] : leaf_link(S.leafFirst[c], S.count[c]);
		}
		g.b0 = Float4{lo[0][0], hi[0][0], lo[0][1], hi[0][1]};
		g.b1 = Float4{lo[1][0], hi[1][0], lo[1][1], hi[1][1]};
		g.bz = Float4{lo[0][2], hi[0][2], lo[1][2], hi[1][2]};
		g.link = Int4{link[0], link[1], 0, 0};
		out[rootLeaf ? 0 : S.innerIdx[node]] = g;
	}
};

// ---- the first levels on the device: privatised in shared memory ---------------------------------------------------------
// With 1 .. 32 nodes in a level, every triangle's atomics land on a few hundred addresses (measured: 6.4 of the 9.0 ms of a
// 268 K-triangle build went into the bin / assign kernels of the first six levels). A CTA therefore accumulates the bins
// (or the children's boxes) of its 2048 triangles in shared memory with the SAME integer min / max / add and flushes only
// the entries it touched: the result is bit-identical to the flat functors (which the host run and the deeper levels
// use), the traffic on the hot addresses drops ~100x.
enum { kTopNodes = 32, kTopTris = 8 };  // levels of <= 32 nodes; triangles per thread

__global__ void __launch_bounds__(256) k_sah_bin_top(SahState S, int levelFirst, int levelEnd) {
	__shared__ int sb[kTopNodes * 3 * kBins * kBinInts];
	const int nb = (levelEnd - levelFirst) * 3 * kBins;
	for(int j = threadIdx.x; j < nb; j += 256) {
		int *b = sb + j * kBinInts;
		b[0] = 0;
		b[1] = b[2] = b[3] = SNL_ORD_POS_INF;
		b[4] = b[5] = b[6] = SNL_ORD_NEG_INF;
	}
	__syncthreads();
	const int base = blockIdx.x * 256 * kTopTris;
	for(int k = 0; k < kTopTris; k++) {
		const int i = base + k * 256 + (int)threadIdx.x;
		if(i >= S.n)
			break;
		const int node = S.triNode[i];
		if(node < levelFirst || node >= levelEnd || S.count[node] <= 4)
			continue;
		const float *lo = S.triLo + i * 3, *hi = S.triHi + i * 3;
		for(int a = 0; a < 3; a++) {
			const int bin = bin_of((lo[a] + hi[a]) * 0.5f, o2f(S.cenLo[node * 3 + a]), o2f(S.cenHi[node * 3 + a]));
			int *b = sb + ((S.binSlot[node - levelFirst] * 3 + a) * kBins + bin) * kBinInts;
			atomicAdd(&b[0], 1);
			for(int c = 0; c < 3; c++) {
				atomicMin(&b[1 + c], f2o(lo[c]));
				atomicMax(&b[4 + c], f2o(hi[c]));
			}
		}
	}
	__syncthreads();
	for(int j = threadIdx.x; j < nb; j += 256) {
		const int *b = sb + j * kBinInts;
		if(!b[0])
			continue;
		int *g = S.bins + (size_t)j * kBinInts;
		atomicAdd(&g[0], b[0]);
		for(int c = 0; c < 3; c++) {
			atomicMin(&g[1 + c], b[1 + c]);
			atomicMax(&g[4 + c], b[4 + c]);
		}
	}
}

__global__ void __launch_bounds__(256) k_sah_assign_top(SahState S, int levelFirst, int levelEnd, int nChildren) {
	__shared__ int sc[2 * kTopNodes * 13];  // per child: box lo xyz, box hi xyz, centre lo xyz, centre hi xyz, count
	for(int j = threadIdx.x; j < nChildren * 13; j += 256) {
		const int f = j % 13;
		sc[j] = f == 12 ? 0 : ((f / 3) & 1) ? SNL_ORD_NEG_INF : SNL_ORD_POS_INF;
	}
	__syncthreads();
	const int base = blockIdx.x * 256 * kTopTris;
	for(int k = 0; k < kTopTris; k++) {
		const int i = base + k * 256 + (int)threadIdx.x;
		if(i >= S.n)
			break;
		const int node = S.triNode[i];
		if(node < levelFirst || node >= levelEnd || S.child[node] < 0)
			continue;
		const int a = S.axis[node];
		int side;
		if(a == 3)
			side = atomicAdd(&S.ticket[node], 1) & 1;
		else {
			const float c = (S.triLo[i * 3 + a] + S.triHi[i * 3 + a]) * 0.5f;
			side = bin_of(c, o2f(S.cenLo[node * 3 + a]), o2f(S.cenHi[node * 3 + a])) > S.plane[node];
		}
		const int dst = S.child[node] + side;
		S.triNode[i] = dst;
		int *acc = sc + (dst - levelEnd) * 13;
		for(int d = 0; d < 3; d++) {
			const float lo = S.triLo[i * 3 + d], hi = S.triHi[i * 3 + d];
			const int cen = f2o((lo + hi) * 0.5f);
			atomicMin(&acc[d], f2o(lo));
			atomicMax(&acc[3 + d], f2o(hi));
			atomicMin(&acc[6 + d], cen);
			atomicMax(&acc[9 + d], cen);
		}
		atomicAdd(&acc[12], 1);
	}
	__syncthreads();
	for(int j = threadIdx.x; j < nChildren; j += 256) {
		const int *acc = sc + j * 13;
		if(!acc[12])
			continue;
		const int node = levelEnd + j;
		for(int d = 0; d < 3; d++) {
			atomicMin(&S.boxLo[node * 3 + d], acc[d]);
			atomicMax(&S.boxHi[node * 3 + d], acc[3 + d]);
			atomicMin(&S.cenLo[node * 3 + d], acc[6 + d]);
			atomicMax(&S.cenHi[node * 3 + d], acc[9 + d]);
		}
		atomicAdd(&S.count[node], acc[12]);
	}
}

// ---- executors: the same driver runs the functors as kernels or as loops ---------------------------------------------
template <class F> __global__ void __launch_bounds__(256) k_sah_for(int n, F f) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i < n)
		f(i);
}

struct DeviceExec {
	cudaStream_t st;
	int launches = 0;
	cudaError_t err = cudaSuccess;
	void *scanTemp = nullptr, *sortTemp = nullptr;
	size_t scanBytes = 0, sortBytes = 0;
	template <class F> void run(int n, const F &f) {
		if(n <= 0 || err != cudaSuccess)
			return;
		k_sah_for<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, f);
		launches++;
		err = cudaGetLastError();
	}
	void bin(const SahState &S, int levelFirst, int levelEnd) {
		if(levelEnd - levelFirst > kTopNodes)
			return run(S.n, FBin{S, levelFirst, levelEnd});
		if(err != cudaSuccess)
			return;
		k_sah_bin_top<<<(unsigned)((S.n + 256 * kTopTris - 1) / (256 * kTopTris)), 256, 0, st>>>(S, levelFirst, levelEnd);
		launches++;
		err = cudaGetLastError();
	}
	void assign(const SahState &S, int levelFirst, int levelEnd, int splits) {
		if(levelEnd - levelFirst > kTopNodes)
			return run(S.n, FAssign{S, levelFirst, levelEnd});
		if(err != cudaSuccess)
			return;
		k_sah_assign_top<<<(unsigned)((S.n + 256 * kTopTris - 1) / (256 * kTopTris)), 256, 0, st>>>(S, levelFirst, levelEnd, 2 * splits);
		launches++;
		err = cudaGetLastError();
	}
	void exclusive_scan(const int *in, int *out, int n) {
		if(err == cudaSuccess)
			err = cub::DeviceScan::ExclusiveSum(scanTemp, scanBytes, in, out, n, st);
		launches += 2;
	}
	int read_int(const int *p) {  // value of a device int (synchronises the stream)
		int v = 0;
		if(err == cudaSuccess)
			err = cudaMemcpyAsync(&v, p, sizeof(int), cudaMemcpyDeviceToHost, st);
		if(err == cudaSuccess)
			err = cudaStreamSynchronize(st);
		return v;
	}
	void sort_pairs(int *keyIn, int *keyOut, unsigned *valIn, unsigned *valOut, int n, int bits) {
		if(err == cudaSuccess)
			err = cub::DeviceRadixSort::SortPairs(sortTemp, sortBytes, keyIn, keyOut, valIn, valOut, n, 0, bits, st);
		launches += 3 * ((bits + 7) / 8);
	}
};

struct HostExec {
	int launches = 0;
	cudaError_t err = cudaSuccess;
	template <class F> void run(int n, const F &f) {
		for(int i = 0; i < n; i++)
			f(i);
	}
	void bin(const SahState &S, int levelFirst, int levelEnd) { run(S.n, FBin{S, levelFirst, levelEnd}); }
	void assign(const SahState &S, int levelFirst, int levelEnd, int) { run(S.n, FAssign{S, levelFirst, levelEnd}); }
	void exclusive_scan(const int *in, int *out, int n) {
		int s = 0;
		for(int i = 0; i < n; i++) {
			const int v = in[i];
			out[i] = s;
			s += v;
		}
	}
	int read_int(const int *p) { return *p; }
	void sort_pairs(int *keyIn, int *keyOut, unsigned *valIn, unsigned *valOut, int n, int) {
		std::vector<unsigned> idx(n);
		std::iota(idx.begin(), idx.end(), 0u);
		std::stable_sort(idx.begin(), idx.end(), [&](unsigned a, unsigned b) { return keyIn[a] < keyIn[b]; });
		for(int i = 0; i < n; i++) {
			keyOut[i] = keyIn[idx[i]];
			valOut[i] = valIn[idx[i]];
		}
	}
};

struct FIota {
	unsigned *v;
	SNL_HD void operator()(int i) const { v[i] = (unsigned)i; }
};
struct FLastFlagPlusScan {  // total = scan[n - 1] + flag[n - 1]
	const int *flag, *scan;
	int n;
	int *out;
	SNL_HD void operator()(int) const { *out = scan[n - 1] + flag[n - 1]; }
};

// The build itself. All pointers of S and the work arrays live where Exec runs. Returns the number of inner nodes written
// to outNodes (>= 1), the height (inner-node levels) and the triangle order.
template <class Exec>
int build_binned_sah(Exec &X, SahState S, int *levelTotal /* 1 int where Exec runs */, int *keyTmp /* [n] */, unsigned *order[2] /* [n] each */,
					 const GenTri *inTris, const RefShTri *inSh, GenTri *outTris, RefShTri *outSh, GpuNode *outNodes, int *heightOut,
					 int *orderOutIdx) {
	const int n = S.n;
	X.run(1, FAllocRoot{S});
	X.run(n, FInitTri{S});
	int levelFirst = 0, levelEnd = 1, height = 0;
	while(levelEnd > levelFirst) {
		const int m = levelEnd - levelFirst;
		// deep chains cannot outgrow the traversal stack: from here on whatever fits a leaf becomes one
		const int forceLeaves = height >= kStack - 8;
		X.run(m, FActive{S, levelFirst});
		X.exclusive_scan(S.active, S.binSlot, m);
		X.run(std::min(m, S.binCap) * 3 * kBins, FClearBins{S});
		X.bin(S, levelFirst, levelEnd);
		X.run(m, FSplit{S, levelFirst, levelEnd, forceLeaves});
		X.exclusive_scan(S.flag, S.scan, m);
		X.run(1, FLastFlagPlusScan{S.flag, S.scan, m, levelTotal});
		const int splits = X.read_int(levelTotal);
		if(X.err != cudaSuccess)
			return -1;
		if(splits == 0)
			break;
		if(levelEnd + 2 * splits > S.cap)
			return -2;
		X.run(m, FAlloc{S, levelFirst, levelEnd});
		X.assign(S, levelFirst, levelEnd, splits);
		levelFirst = levelEnd;
		levelEnd += 2 * splits;
		height++;
		if(height > kStack - 2)
			return -3;
	}
	const int nNodes = levelEnd;
	// leaves = contiguous triangle ranges: sort the triangles by the id of the leaf they ended in (stable)
	int bits = 1;
	while((1 << bits) < nNodes && bits < 31)
		bits++;
	X.run(n, FIota{order[0]});
	X.sort_pairs(S.triNode, keyTmp, order[0], order[1], n, bits);
	X.run(n, FLeafFirst{S, keyTmp});
	X.run(n, FPermute{order[1], inTris, inSh, outTris, outSh});
	X.run(nNodes, FMarkInner{S});
	X.exclusive_scan(S.inner, S.innerIdx, nNodes);
	X.run(1, FLastFlagPlusScan{S.inner, S.innerIdx, nNodes, levelTotal});
	X.run(nNodes, FEmit{S, outNodes});
	const int nInner = X.read_int(levelTotal);
	if(X.err != cudaSuccess)
		return -1;
	*heightOut = std::max(height, 1);
	*orderOutIdx = 1;
	return std::max(nInner, 1);
}


// carves the SahState arrays (and the work arrays of build_binned_sah) out of one block of ints
struct SahLayout {
	size_t total = 0;
	size_t triLo, triHi, triNode, boxLo, boxHi, cenLo, cenHi, count, child, axis, plane, ticket, flag, scan, active, binSlot, bins, leafFirst,
		inner, innerIdx, levelTotal, keyTmp, order0, order1;
	int cap, binCap;
	explicit SahLayout(int n) {
		cap = 2 * n + 2;
		binCap = n / 5 + 2;
		auto take = [&](size_t ints) {
			const size_t at = total;
			total += (ints + 63) & ~(size_t)63;
			return at;
		};
		triLo = take(3 * (size_t)n), triHi = take(3 * (size_t)n), triNode = take(n);
		boxLo = take(3 * (size_t)cap), boxHi = take(3 * (size_t)cap), cenLo = take(3 * (size_t)cap), cenHi = take(3 * (size_t)cap);
		count = take(cap), child = take(cap), axis = take(cap), plane = take(cap), ticket = take(cap);
		flag = take(cap), scan = take(cap), active = take(cap), binSlot = take(cap);
		bins = take((size_t)binCap * 3 * kBins * kBinInts);
		leafFirst = take(cap), inner = take(cap), innerIdx = take(cap), levelTotal = take(1);
		keyTmp = take(n), order0 = take(n), order1 = take(n);
	}
	SahState state(int *base, const GenTri *tris, int n) const {
		SahState S;
		S.n = n, S.cap = cap, S.tris = tris;
		S.triLo = reinterpret_cast<float *>(base + triLo), S.triHi = reinterpret_cast<float *>(base + triHi);
		S.triNode = base + triNode;
		S.boxLo = base + boxLo, S.boxHi = base + boxHi, S.cenLo = base + cenLo, S.cenHi = base + cenHi;
		S.count = base + count, S.child = base + child, S.axis = base + axis, S.plane = base + plane, S.ticket = base + ticket;
		S.flag = base + flag, S.scan = base + scan, S.active = base + active, S.binSlot = base + binSlot;
		S.bins = base + bins, S.binCap = binCap;
		S.leafFirst = base + leafFirst, S.inner = base + inner, S.innerIdx = base + innerIdx;
		return S;
	}
};

}  // namespace

cudaError_t device_build_sah(const GenTri *inTris, const RefShTri *inSh, int n, GenTri *outTris, RefShTri *outSh, GpuNode *outNodes,
							 int *nNodesOut, int *heightOut, float bboxOut[6], cudaStream_t st, int *launches, float *kernelMs) {
	if(n <= 0)
		return cudaErrorInvalidValue;
	const SahLayout L(n);
	DeviceExec X;
	X.st = st;
	cudaError_t err = cub::DeviceScan::ExclusiveSum(nullptr, X.scanBytes, (const int *)nullptr, (int *)nullptr, L.cap, st);
	if(err == cudaSuccess)
		err = cub::DeviceRadixSort::SortPairs(nullptr, X.sortBytes, (const int *)nullptr, (int *)nullptr, (const unsigned *)nullptr,
											  (unsigned *)nullptr, n, 0, 31, st);
	if(err != cudaSuccess)
		return err;
	int *base = nullptr;
	unsigned char *temp = nullptr;
	err = cudaMalloc((void **)&base, L.total * sizeof(int));
	if(err == cudaSuccess)
		err = cudaMalloc((void **)&temp, X.scanBytes + X.sortBytes + 512);
	cudaEvent_t e0 = nullptr, e1 = nullptr;
	if(err == cudaSuccess)
		err = cudaEventCreate(&e0);
	if(err == cudaSuccess)
		err = cudaEventCreate(&e1);
	int result = -1;
	if(err == cudaSuccess) {
		X.scanTemp = temp;
		X.sortTemp = temp + ((X.scanBytes + 255) & ~(size_t)255);
		SahState S = L.state(base, inTris, n);
		unsigned *order[2] = {reinterpret_cast<unsigned *>(base + L.order0), reinterpret_cast<unsigned *>(base + L.order1)};
		int orderIdx = 0;
		cudaEventRecord(e0, st);
		result = build_binned_sah(X, S, base + L.levelTotal, base + L.keyTmp, order, inTris, inSh, outTris, outSh, outNodes, heightOut, &orderIdx);
		cudaEventRecord(e1, st);
		err = X.err;
		if(err == cudaSuccess && result > 0) {
			int ord[6];
			err = cudaMemcpyAsync(ord, S.boxLo, 12, cudaMemcpyDeviceToHost, st);
			if(err == cudaSuccess)
				err = cudaMemcpyAsync(ord + 3, S.boxHi, 12, cudaMemcpyDeviceToHost, st);
			if(err == cudaSuccess)
				err = cudaStreamSynchronize(st);
			for(int c = 0; c < 6 && err == cudaSuccess; c++)
				bboxOut[c] = o2f(ord[c]);
			if(err == cudaSuccess && kernelMs)
				err = cudaEventElapsedTime(kernelMs, e0, e1);
		}
	}
	if(e0)
		cudaEventDestroy(e0);
	if(e1)
		cudaEventDestroy(e1);
	cudaFree(base);
	cudaFree(temp);
	if(err != cudaSuccess)
		return err;
	if(result <= 0)  // -2: node capacity, -3: deeper than the traversal stack
		return cudaErrorInvalidValue;
	*nNodesOut = result;
	if(launches)
		*launches = X.launches;
	return cudaSuccess;
}

// The same build as plain loops on the host (no device needed): pins the algorithm in the CPU test-suite.
int host_build_binned_sah(const GenTri *inTris, int n, std::vector<GpuNode> &nodes, std::vector<unsigned> &orderOut, int *heightOut) {
	if(n <= 0)
		return -1;
	const SahLayout L(n);
	std::vector<int> mem(L.total, 0);
	std::vector<GenTri> outTris(n);
	nodes.assign(std::max(n - 1, 1), GpuNode());
	HostExec X;
	SahState S = L.state(mem.data(), inTris, n);
	unsigned *order[2] = {reinterpret_cast<unsigned *>(mem.data() + L.order0), reinterpret_cast<unsigned *>(mem.data() + L.order1)};
	int orderIdx = 0;
	const int result = build_binned_sah(X, S, mem.data() + L.levelTotal, mem.data() + L.keyTmp, order, inTris, (const RefShTri *)nullptr,
										outTris.data(), (RefShTri *)nullptr, nodes.data(), heightOut, &orderIdx);
	if(result <= 0)
		return result;
	nodes.resize(result);
	orderOut.assign(order[orderIdx], order[orderIdx] + n);
	return result;
}

}  // namespace snl
