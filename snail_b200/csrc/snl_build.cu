// snail_b200 — device-side BVH build: replaces BVH::Construct + FindSplitSweep
// (src/bvh/tree.cpp:51-159,289-328) for scenes where the host's O(N log^2 N) full-sweep SAH build dominates load time.
//
// Not the reference's tree (that one stays the default, bit-identical, built by snl_host_scene.cpp): a linear BVH over
// 63-bit Morton codes of the triangle-box centres (Karras 2012), built entirely in HBM:
//   k_lbvh_bounds     triangle boxes (Triangle::GetBBox: a, a+ba, a+ca) + scene box         read 64 B, write 32 B / tri
//   k_lbvh_morton     63-bit Morton key of the box centre                                   read 32 B, write 12 B / tri
//   cub radix sort    (key, index) pairs                                                    8 passes over 12 B / tri
//   k_lbvh_permute    triangles / shading records / boxes into Morton order                 read+write 160 B / tri
//   k_lbvh_hierarchy  one thread per inner node: range, split, children, parents
//   k_lbvh_refit      bottom-up: boxes, triangle counts, SAH cost, leaf collapse, height
//   k_lbvh_emit       GpuNode records (both child boxes + links), the layout the traversal kernels read
// Subtrees of <= 4 triangles become leaves like the reference's `count <= 4` rule (src/bvh/tree.cpp:54); larger ones
// (<= 32) collapse when the SAH says a leaf is cheaper, with the reference's costs (traverse 0, intersect 1, :70-71).
// Any valid BVH yields the same nearest hit (t, u, v come from the triangle test alone), so parity with the reference
// is checked on hits / frames, not on the tree (tests/test_gpu_build.py).
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <cstring>
#include <vector>

#include "snl_device.cuh"
#include "snl_kernels.h"

namespace snl {

namespace {

// float <-> int with the same ordering, for atomicMin / atomicMax on floats
__device__ __forceinline__ int f2ord(float f) {
	const int i = __float_as_int(f);
	return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

struct Box {
	float4 lo, hi;  // xyz used
};

__device__ __forceinline__ Box box_union(const Box &a, const Box &b) {
	Box r;
	r.lo = make_float4(fminf(a.lo.x, b.lo.x), fminf(a.lo.y, b.lo.y), fminf(a.lo.z, b.lo.z), 0.0f);
	r.hi = make_float4(fmaxf(a.hi.x, b.hi.x), fmaxf(a.hi.y, b.hi.y), fmaxf(a.hi.z, b.hi.z), 0.0f);
	return r;
}

__device__ __forceinline__ Box load_box_cg(const Box *p) {
	Box b;
	b.lo = __ldcg(&p->lo);
	b.hi = __ldcg(&p->hi);
	return b;
}

// BoxSA (src/bvh/tree.cpp:40-42): (w * (d + h) + d * h) * 2
__device__ __forceinline__ float box_sa(const Box &b) {
	const float w = b.hi.x - b.lo.x, h = b.hi.y - b.lo.y, d = b.hi.z - b.lo.z;
	return (w * (d + h) + d * h) * 2.0f;
}

__global__ void __launch_bounds__(256) k_lbvh_bounds(const GenTri *__restrict__ tris, int n, Box *__restrict__ boxes,
													 int *__restrict__ sceneOrd /* [6] min xyz, max xyz */) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	float lo[3] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F}, hi[3] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
	if(i < n) {
		float4 a, ba, ca, nn;
		ldg8(&tris[i].a, a, ba);
		ldg8(&tris[i].ca, ca, nn);
		const float p[3][3] = {{a.x, a.y, a.z}, {a.x + ba.x, a.y + ba.y, a.z + ba.z}, {a.x + ca.x, a.y + ca.y, a.z + ca.z}};
#pragma unroll
		for(int c = 0; c < 3; c++) {
			lo[c] = fminf(p[0][c], fminf(p[1][c], p[2][c]));
			hi[c] = fmaxf(p[0][c], fmaxf(p[1][c], p[2][c]));
		}
		Box b;
		b.lo = make_float4(lo[0], lo[1], lo[2], 0.0f);
		b.hi = make_float4(hi[0], hi[1], hi[2], 0.0f);
		boxes[i] = b;
	}
#pragma unroll
	for(int c = 0; c < 3; c++) {
#pragma unroll
		for(int o = 16; o; o >>= 1) {
			lo[c] = fminf(lo[c], __shfl_xor_sync(SNL_FULL, lo[c], o));
			hi[c] = fmaxf(hi[c], __shfl_xor_sync(SNL_FULL, hi[c], o));
		}
	}
	if((threadIdx.x & 31) == 0) {
#pragma unroll
		for(int c = 0; c < 3; c++) {
			atomicMin(&sceneOrd[c], f2ord(lo[c]));
			atomicMax(&sceneOrd[3 + c], f2ord(hi[c]));
		}
	}
}

__device__ __forceinline__ unsigned long long spread21(unsigned v) {
	unsigned long long x = v & 0x1fffffu;
	x = (x | x << 32) & 0x1f00000000ffffull;
	x = (x | x << 16) & 0x1f0000ff0000ffull;
	x = (x | x << 8) & 0x100f00f00f00f00full;
	x = (x | x << 4) & 0x10c30c30c30c30c3ull;
	x = (x | x << 2) & 0x1249249249249249ull;
	return x;
}

__global__ void __launch_bounds__(256) k_lbvh_morton(const Box *__restrict__ boxes, int n, const int *__restrict__ sceneOrd,
													 unsigned long long *__restrict__ keys, unsigned *__restrict__ vals) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n)
		return;
	const Box b = boxes[i];
	const float c[3] = {(b.lo.x + b.hi.x) * 0.5f, (b.lo.y + b.hi.y) * 0.5f, (b.lo.z + b.hi.z) * 0.5f};
	unsigned q[3];
#pragma unroll
	for(int a = 0; a < 3; a++) {
		const float lo = ord2f(sceneOrd[a]), hi = ord2f(sceneOrd[3 + a]);
		const float ext = hi - lo;
		float t = ext > 0.0f ? (c[a] - lo) / ext : 0.0f;
		t = fminf(fmaxf(t, 0.0f), 1.0f);
		q[a] = min((unsigned)(t * 2097152.0f), 2097151u);
	}
	keys[i] = spread21(q[0]) << 2 | spread21(q[1]) << 1 | spread21(q[2]);
	vals[i] = (unsigned)i;
}

__global__ void __launch_bounds__(256) k_lbvh_permute(const unsigned *__restrict__ order, int n, const GenTri *__restrict__ inTris,
													  const RefShTri *__restrict__ inSh, const Box *__restrict__ inBox,
													  GenTri *__restrict__ outTris, RefShTri *__restrict__ outSh,
													  Box *__restrict__ outBox) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n)
		return;
	const unsigned src = order[i];
	const float4 *t = reinterpret_cast<const float4 *>(inTris + src);
	float4 *o = reinterpret_cast<float4 *>(outTris + i);
	o[0] = __ldg(t + 0);
	o[1] = __ldg(t + 1);
	o[2] = __ldg(t + 2);
	o[3] = __ldg(t + 3);
	if(inSh) {
		const float4 *s = reinterpret_cast<const float4 *>(inSh + src);
		float4 *so = reinterpret_cast<float4 *>(outSh + i);
		so[0] = __ldg(s + 0);
		so[1] = __ldg(s + 1);
		so[2] = __ldg(s + 2);
		so[3] = __ldg(s + 3);
	}
	outBox[i] = inBox[src];
}

// ---- hierarchy (Karras 2012, "Maximizing Parallelism in the Construction of BVHs, Octrees, and k-d Trees") -------------
// Leaves are the n sorted triangles, inner node i (0 .. n-2) covers the key range [first[i], last[i]]; node 0 is the root.
// Child references: >= 0 inner node, < 0 leaf ~index.

struct Hier {
	int *left, *right;    // [n-1]
	int *first, *last;    // [n-1]
	int *parent;          // [n-1] parent of inner node (root: -1)
	int *leafParent;      // [n]
};

__device__ __forceinline__ int delta(const unsigned long long *__restrict__ keys, int n, int i, int j) {
	if(j < 0 || j >= n)
		return -1;
	const unsigned long long a = keys[i], b = keys[j];
	if(a == b)  // duplicate keys: fall back to the index, as the paper suggests
		return 64 + __clz(i ^ j);
	return __clzll((long long)(a ^ b));
}

__global__ void __launch_bounds__(256) k_lbvh_hierarchy(const unsigned long long *__restrict__ keys, int n, Hier H) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n - 1)
		return;
	const int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
	const int dmin = delta(keys, n, i, i - d);
	int lmax = 2;
	while(delta(keys, n, i, i + lmax * d) > dmin)
		lmax <<= 1;
	int l = 0;
	for(int t = lmax >> 1; t >= 1; t >>= 1)
		if(delta(keys, n, i, i + (l + t) * d) > dmin)
			l += t;
	const int j = i + l * d;
	const int dnode = delta(keys, n, i, j);
	int s = 0;
	for(int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
		if(delta(keys, n, i, i + (s + t) * d) > dnode)
			s += t;
		if(t == 1)
			break;
	}
	const int gamma = i + s * d + min(d, 0);
	const int lo = min(i, j), hi = max(i, j);
	const int L = lo == gamma ? ~gamma : gamma;
	const int R = hi == gamma + 1 ? ~(gamma + 1) : gamma + 1;
	H.left[i] = L;
	H.right[i] = R;
	H.first[i] = lo;
	H.last[i] = hi;
	if(L >= 0)
		H.parent[L] = i;
	else
		H.leafParent[~L] = i;
	if(R >= 0)
		H.parent[R] = i;
	else
		H.leafParent[~R] = i;
	if(i == 0)
		H.parent[0] = -1;
}

struct NodeInfo {
	Box *box;        // [n-1]
	float *cost;     // [n-1] SAH cost of the subtree after collapsing
	int *height;     // [n-1] inner-node levels below and including this node (0 when collapsed into a leaf)
	int *collapsed;  // [n-1] 1: the subtree is one leaf
	int *visits;     // [n-1] zeroed; second arrival processes the node
};

// One thread per leaf climbs towards the root; the second thread to reach a node owns it (both children are final).
__global__ void __launch_bounds__(256) k_lbvh_refit(int n, Hier H, const Box *__restrict__ leafBox, NodeInfo N) {
	const int leaf = blockIdx.x * blockDim.x + threadIdx.x;
	if(leaf >= n)
		return;
	int p = H.leafParent[leaf];
	while(p >= 0) {
		__threadfence();
		if(atomicAdd(&N.visits[p], 1) == 0)
			return;
		__threadfence();
		const int L = H.left[p], R = H.right[p];
		// children were written by other threads of this launch: read them from L2 (ld.global.cg)
		const Box bl = L < 0 ? leafBox[~L] : load_box_cg(&N.box[L]), br = R < 0 ? leafBox[~R] : load_box_cg(&N.box[R]);
		const Box b = box_union(bl, br);
		const float sa = box_sa(b);
		const int cnt = H.last[p] - H.first[p] + 1;
		const float cl = L < 0 ? box_sa(bl) : __ldcg(&N.cost[L]), cr = R < 0 ? box_sa(br) : __ldcg(&N.cost[R]);
		const int hl = L < 0 ? 0 : __ldcg(&N.height[L]), hr = R < 0 ? 0 : __ldcg(&N.height[R]);
		const float leafCost = (float)cnt * sa, splitCost = cl + cr;  // traverseCost 0, intersectCost 1
		const bool collapse = cnt <= 4 || (cnt <= kLeafMaxCount && leafCost <= splitCost);
		N.box[p] = b;
		N.cost[p] = collapse ? leafCost : splitCost;
		N.height[p] = collapse ? 0 : 1 + max(hl, hr);
		N.collapsed[p] = collapse;
		p = H.parent[p];
	}
}

__device__ __forceinline__ int leaf_link(int first, int count) { return ~(first | ((count - 1) << kLeafFirstBits)); }

__global__ void __launch_bounds__(256) k_lbvh_emit(int n, Hier H, const Box *__restrict__ leafBox, NodeInfo N,
												   GpuNode *__restrict__ out) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= max(n - 1, 1))
		return;
	Box b[2];
	int link[2];
	if(n == 1 || (i == 0 && N.collapsed[0])) {
		// the whole scene is one leaf: node 0 = two links to it (the second visit cannot change the result)
		b[0] = b[1] = n == 1 ? leafBox[0] : N.box[0];
		link[0] = link[1] = leaf_link(0, n);
	}
	else {
		const int c[2] = {H.left[i], H.right[i]};
#pragma unroll
		for(int k = 0; k < 2; k++) {
			if(c[k] < 0) {
				b[k] = leafBox[~c[k]];
				link[k] = leaf_link(~c[k], 1);
			}
			else {
				b[k] = N.box[c[k]];
				link[k] = N.collapsed[c[k]] ? leaf_link(H.first[c[k]], H.last[c[k]] - H.first[c[k]] + 1) : c[k];
			}
		}
	}
	float4 *o = reinterpret_cast<float4 *>(out + i);
	o[0] = make_float4(b[0].lo.x, b[0].hi.x, b[0].lo.y, b[0].hi.y);
	o[1] = make_float4(b[1].lo.x, b[1].hi.x, b[1].lo.y, b[1].hi.y);
	o[2] = make_float4(b[0].lo.z, b[0].hi.z, b[1].lo.z, b[1].hi.z);
	reinterpret_cast<int4 *>(out + i)[3] = make_int4(link[0], link[1], 0, 0);
}

__global__ void k_lbvh_init(int *sceneOrd) {
	if(threadIdx.x < 3) {
		sceneOrd[threadIdx.x] = f2ord(CUDART_INF_F);
		sceneOrd[3 + threadIdx.x] = f2ord(-CUDART_INF_F);
	}
}

}  // namespace


// One allocation for all the scratch arrays of a build (cudaMalloc costs more than the whole build of a small scene).
struct BuildArena {
	unsigned char *base = nullptr;
	size_t size = 0, used = 0;
	~BuildArena() {
		if(base)
			cudaFree(base);
	}
	static size_t pad(size_t bytes) { return (bytes + 255) & ~(size_t)255; }
	template <class T> void need(size_t count) { size += pad(std::max<size_t>(count, 1) * sizeof(T)); }
	template <class T> T *take(size_t count) {
		T *p = reinterpret_cast<T *>(base + used);
		used += pad(std::max<size_t>(count, 1) * sizeof(T));
		return p;
	}
};

#define SNL_BUILD_TRY(call)       \
	do {                          \
		cudaError_t e_ = (call);  \
		if(e_ != cudaSuccess)     \
			return e_;            \
	} while(0)

cudaError_t device_build_lbvh(const GenTri *inTris, const RefShTri *inSh, int n, GenTri *outTris, RefShTri *outSh,
							  GpuNode *outNodes, int *heightOut, float bboxOut[6], cudaStream_t st, int *launches,
							  float *kernelMs) {
	if(n <= 0)
		return cudaErrorInvalidValue;
	const int ni = std::max(n - 1, 1);
	const unsigned gridN = (unsigned)((n + 255) / 256), gridI = (unsigned)((ni + 255) / 256);
	size_t tempBytes = 0;
	{
		cub::DoubleBuffer<unsigned long long> dk(nullptr, nullptr);
		cub::DoubleBuffer<unsigned> dv(nullptr, nullptr);
		SNL_BUILD_TRY(cub::DeviceRadixSort::SortPairs(nullptr, tempBytes, dk, dv, n, 0, 63, st));
	}
	BuildArena A;
	A.need<int>(6);
	A.need<Box>(n);
	A.need<Box>(n);
	A.need<unsigned long long>(n);
	A.need<unsigned long long>(n);
	A.need<unsigned>(n);
	A.need<unsigned>(n);
	for(int k = 0; k < 5; k++)
		A.need<int>(ni);
	A.need<int>(n);
	A.need<Box>(ni);
	A.need<float>(ni);
	for(int k = 0; k < 3; k++)
		A.need<int>(ni);
	A.need<unsigned char>(tempBytes);
	SNL_BUILD_TRY(cudaMalloc((void **)&A.base, A.size));

	int *sceneOrd = A.take<int>(6);
	Box *triBox = A.take<Box>(n), *leafBox = A.take<Box>(n);
	unsigned long long *keys[2] = {A.take<unsigned long long>(n), A.take<unsigned long long>(n)};
	unsigned *vals[2] = {A.take<unsigned>(n), A.take<unsigned>(n)};
	Hier H;
	NodeInfo N;
	H.left = A.take<int>(ni);
	H.right = A.take<int>(ni);
	H.first = A.take<int>(ni);
	H.last = A.take<int>(ni);
	H.parent = A.take<int>(ni);
	H.leafParent = A.take<int>(n);
	N.box = A.take<Box>(ni);
	N.cost = A.take<float>(ni);
	// height, collapsed, visits are adjacent: one memset
	N.height = A.take<int>(ni);
	N.collapsed = A.take<int>(ni);
	N.visits = A.take<int>(ni);
	unsigned char *temp = A.take<unsigned char>(tempBytes);

	cudaEvent_t e0 = nullptr, e1 = nullptr;
	SNL_BUILD_TRY(cudaEventCreate(&e0));
	SNL_BUILD_TRY(cudaEventCreate(&e1));
	struct EvGuard {
		cudaEvent_t a, b;
		~EvGuard() {
			cudaEventDestroy(a);
			cudaEventDestroy(b);
		}
	} guard{e0, e1};
	SNL_BUILD_TRY(cudaEventRecord(e0, st));
	SNL_BUILD_TRY(cudaMemsetAsync(N.height, 0, (size_t)((unsigned char *)(N.visits + ni) - (unsigned char *)N.height), st));

	k_lbvh_init<<<1, 32, 0, st>>>(sceneOrd);
	k_lbvh_bounds<<<gridN, 256, 0, st>>>(inTris, n, triBox, sceneOrd);
	k_lbvh_morton<<<gridN, 256, 0, st>>>(triBox, n, sceneOrd, keys[0], vals[0]);
	SNL_BUILD_TRY(cudaGetLastError());

	cub::DoubleBuffer<unsigned long long> dk(keys[0], keys[1]);
	cub::DoubleBuffer<unsigned> dv(vals[0], vals[1]);
	SNL_BUILD_TRY(cub::DeviceRadixSort::SortPairs(temp, tempBytes, dk, dv, n, 0, 63, st));

	k_lbvh_permute<<<gridN, 256, 0, st>>>(dv.Current(), n, inTris, inSh, triBox, outTris, outSh, leafBox);
	if(n >= 2) {
		k_lbvh_hierarchy<<<gridI, 256, 0, st>>>(dk.Current(), n, H);
		k_lbvh_refit<<<gridN, 256, 0, st>>>(n, H, leafBox, N);
	}
	k_lbvh_emit<<<gridI, 256, 0, st>>>(n, H, leafBox, N, outNodes);
	SNL_BUILD_TRY(cudaGetLastError());
	if(launches)
		*launches = 5 + (n >= 2 ? 2 : 0) + 8;  // + the radix sort's passes (upsweep / scan / downsweep per digit, approx.)

	SNL_BUILD_TRY(cudaEventRecord(e1, st));
	int ord[6], height = 0;
	SNL_BUILD_TRY(cudaMemcpyAsync(ord, sceneOrd, sizeof(ord), cudaMemcpyDeviceToHost, st));
	if(n >= 2)
		SNL_BUILD_TRY(cudaMemcpyAsync(&height, N.height, sizeof(int), cudaMemcpyDeviceToHost, st));
	SNL_BUILD_TRY(cudaStreamSynchronize(st));
	for(int c = 0; c < 6; c++) {
		const int i = ord[c];
		const int bits = i >= 0 ? i : i ^ 0x7fffffff;
		memcpy(&bboxOut[c], &bits, 4);
	}
	*heightOut = std::max(height, 1);
	if(kernelMs)
		SNL_BUILD_TRY(cudaEventElapsedTime(kernelMs, e0, e1));
	return cudaSuccess;
}

}  // namespace snl
