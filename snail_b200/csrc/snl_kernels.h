// snail_b200 — host-callable launch wrappers of the sm_100a kernels (defined in snl_kernels.cu).
#pragma once

#include <cuda_runtime.h>

#include "snl_types.h"

namespace snl {

struct SceneDev;
struct PreSet;
struct WfBuffers;
struct CommCtl;

cudaError_t launch_pretransform(const GenTri *gtris, int nTris, const float origin[3], PreTri *out, cudaStream_t st);
cudaError_t launch_primary(const SceneDev &S, const PreSet &pre, const RayGenParams &g, int W, int H, int *triId, float *t,
						   float *bary, int *instId, cudaStream_t st);
cudaError_t launch_trace_rays(const SceneDev &S, const PreSet &pre, const float *o3, const float *d3,
							  const unsigned char *mask, int n, bool shared, int *triId, float *t, float *bary,
							  cudaStream_t st);
cudaError_t launch_trace_shadow(const SceneDev &S, const PreSet &pre, const float origin[3], const float *d3, float *tmax,
								int n, cudaStream_t st);
cudaError_t launch_raytrace(const SceneDev &S, const PreSet &pre, const float *o3, const float *d3,
							const unsigned char *mask, int n, bool shared, float *rgb3, cudaStream_t st);
// packets == null: the full image, packet p = (p % packetsX, p / packetsX); first / count then pick a sub-range of the
// packets (count < 0: all of them).
cudaError_t launch_render(const SceneDev &S, const PreSet &pre, const RayGenParams &g, int W, int H, const int4 *packets,
						  int nPackets, unsigned char *out, size_t pitch, int variant, cudaStream_t st, int first = 0,
						  int count = -1, unsigned *costOut = nullptr);
// snl_render_multi: every band [bandFirst[b], bandFirst[b + 1]) of a rank's packet list re-ordered by descending cost[]
cudaError_t launch_sort_band_packets(const int4 *in, const unsigned *cost, int4 *out, const int *bandFirst, int nBands,
									 cudaStream_t st);

// Multi-GPU frame protocol (snl_comm.cu; CommCtl lives in rank 0's memory):
// rank 0, before its k_render: waits until every rank has finished frame seq - 2, clears slot seq & 1, opens it
cudaError_t launch_frame_open(CommCtl *ctl, unsigned seq, int world, cudaStream_t st);
// every rank, after its k_render: adds its ray counters to the frame's, marks the rank done with frame seq
cudaError_t launch_frame_publish(CommCtl *ctl, const Counters *local, unsigned seq, int rank, cudaStream_t st);
// every rank, in-stream behind its k_render of a row band: one more rank's share of that band is stored
cudaError_t launch_band_signal(unsigned *counter, cudaStream_t st);
// rank 0 (and a peer waiting for a frame slot to open): returns once *counter >= expected (all packets of a band stored / all ranks' counters in)
cudaError_t launch_wait_counter(const unsigned *counter, unsigned expected, unsigned *error, cudaStream_t st);

// Farm tile format (src/render.cpp:140-166): frame tiles -> planar x, y - x, z - x byte planes at out + offsets[i].
cudaError_t launch_farm_pack(const unsigned char *frame, size_t pitch, const int4 *tiles, const int *offsets, int nTiles,
							 unsigned char *out, cudaStream_t st);

// Wavefront form of launch_render for frames with the reflection bounce (snl_wavefront.cu). *launches = kernels issued.
cudaError_t launch_render_wavefront(const SceneDev &S, const PreSet &pre, const RayGenParams &g, int W, int H,
									const int4 *packets, int nPackets, const WfBuffers &B, int persistentCtas,
									unsigned char *out, size_t pitch, cudaStream_t st, int *launches);

// Device-side BVH build (snl_build.cu): Morton-order LBVH over `n` triangles. Writes the triangles / shading records
// in leaf order, max(n-1, 1) GpuNode records (node 0 = root), the inner-node height of the tree and the scene box
// {min xyz, max xyz}. Synchronises the stream. *kernelMs = device time of the kernels (events on `st`).
cudaError_t device_build_lbvh(const GenTri *inTris, const RefShTri *inSh, int n, GenTri *outTris, RefShTri *outSh,
							  GpuNode *outNodes, int *heightOut, float bboxOut[6], cudaStream_t st, int *launches,
							  float *kernelMs);

// Device-side binned-SAH build (snl_build_sah.cu): same outputs as device_build_lbvh plus the number of GpuNode records
// written (<= n - 1). Synchronises the stream once per tree level.
cudaError_t device_build_sah(const GenTri *inTris, const RefShTri *inSh, int n, GenTri *outTris, RefShTri *outSh,
							 GpuNode *outNodes, int *nNodesOut, int *heightOut, float bboxOut[6], cudaStream_t st, int *launches,
							 float *kernelMs);

}  // namespace snl
