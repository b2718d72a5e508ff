// snail_b200 — builder-side measurement of the ceilings k_render is held against (snl_measure_peaks): the scene is
// L2-resident and the walk is served by L1, so the HBM copy peak of MEASURED_PEAKS.json is not its roofline. These are
// plain streaming microbenchmarks timed with CUDA events:
//   L1   : every CTA re-reads its own 16 KB window (ld.global.ca.v4, coalesced) - bytes / s delivered by L1 to the SMs
//   L2   : the whole grid streams a 48 MB buffer (> all L1s together, < L2) with ld.global.cg.v4, many passes
//   HBM  : one pass over 2 GiB (> L2) with ld.global.cg.v4 - read-only counterpart of the driver's copy figure
//   issue: independent FFMA / integer-ALU chains, 1:1 - warp instructions issued per second (4 schedulers per SM)
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstring>

#include "../../include/snail_b200.h"

namespace snl {
void set_error(const char *fmt, ...);
}

namespace {

__device__ __forceinline__ uint4 ld_ca(const uint4 *p) {
	uint4 v;
	asm volatile("ld.global.ca.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
	return v;
}
__device__ __forceinline__ uint4 ld_cg(const uint4 *p) {
	uint4 v;
	asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
	return v;
}

// window = 1024 uint4 (16 KB) per CTA, 256 threads, 4 loads in flight per thread and iteration
__global__ void __launch_bounds__(256) k_peak_l1(const uint4 *__restrict__ buf, int iters, unsigned *sink) {
	const uint4 *w = buf + (size_t)blockIdx.x * 1024;
	unsigned acc = 0;
	for(int it = 0; it < iters; it++) {
#pragma unroll
		for(int k = 0; k < 4; k++) {
			const uint4 v = ld_ca(w + ((threadIdx.x + k * 256 + it * 32) & 1023));
			acc += v.x ^ v.y ^ v.z ^ v.w;
		}
	}
	if(acc == 0x12345678u)
		*sink = acc;
}

__global__ void __launch_bounds__(256) k_peak_stream(const uint4 *__restrict__ buf, size_t n, int passes, unsigned *sink) {
	unsigned acc = 0;
	const size_t stride = (size_t)gridDim.x * blockDim.x;
	for(int p = 0; p < passes; p++)
		for(size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i + 3 * stride < n; i += 4 * stride) {
			const uint4 a = ld_cg(buf + i), b = ld_cg(buf + i + stride), c = ld_cg(buf + i + 2 * stride), d = ld_cg(buf + i + 3 * stride);
			acc += a.x ^ b.y ^ c.z ^ d.w;
		}
	if(acc == 0x12345678u)
		*sink = acc;
}

// 8 independent FFMA chains interleaved with 8 independent integer min/max chains: 16 counted instructions per step
__global__ void __launch_bounds__(256) k_peak_issue(int iters, float seed, float *sink) {
	float f[8];
	int a[8];
#pragma unroll
	for(int k = 0; k < 8; k++) {
		f[k] = seed + k + threadIdx.x;
		a[k] = (int)threadIdx.x + k;
	}
	for(int it = 0; it < iters; it++) {
#pragma unroll
		for(int u = 0; u < 8; u++)
#pragma unroll
			for(int k = 0; k < 8; k++) {
				asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(f[k]) : "f"(seed));
				asm volatile("max.s32 %0, %0, %1;" : "+r"(a[k]) : "r"(it + u));
			}
	}
	float s = 0.0f;
#pragma unroll
	for(int k = 0; k < 8; k++)
		s += f[k] + (float)a[k];
	if(s == 0.12345f)
		*sink = s;
}

template <class F> cudaError_t time_ms(F launch, int reps, float *best) {
	cudaEvent_t e0, e1;
	cudaError_t err = cudaEventCreate(&e0);
	if(err != cudaSuccess)
		return err;
	err = cudaEventCreate(&e1);
	if(err != cudaSuccess)
		return err;
	*best = 1e30f;
	for(int r = 0; r < reps + 1 && err == cudaSuccess; r++) {  // first repetition = warm-up
		cudaEventRecord(e0);
		launch();
		cudaEventRecord(e1);
		err = cudaEventSynchronize(e1);
		float ms = 0.0f;
		if(err == cudaSuccess)
			err = cudaEventElapsedTime(&ms, e0, e1);
		if(r > 0)
			*best = std::min(*best, ms);
	}
	if(err == cudaSuccess)
		err = cudaGetLastError();
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	return err;
}

}  // namespace

extern "C" int snl_measure_peaks(int device, snl_peaks *out) {
	if(!out) {
		snl::set_error("null argument");
		return SNL_ERR_INVALID;
	}
	memset(out, 0, sizeof(*out));
	cudaError_t err = cudaSetDevice(device);
	cudaDeviceProp prop;
	if(err == cudaSuccess)
		err = cudaGetDeviceProperties(&prop, device);
	if(err != cudaSuccess) {
		snl::set_error("no CUDA device %d: %s", device, cudaGetErrorString(err));
		return SNL_ERR_NO_DEVICE;
	}
	const int sms = prop.multiProcessorCount;
	int clockKHz = 0;
	cudaDeviceGetAttribute(&clockKHz, cudaDevAttrClockRate, device);
	out->sms = sms;
	out->sm_clock_mhz_max = clockKHz * 1e-3f;
	unsigned *sink = nullptr;
	uint4 *buf = nullptr;
	const size_t hbmBytes = (size_t)2 << 30, l2Bytes = (size_t)48 << 20;
	err = cudaMalloc(&sink, 64);
	if(err == cudaSuccess)
		err = cudaMalloc(&buf, hbmBytes);
	if(err == cudaSuccess)
		err = cudaMemset(buf, 1, hbmBytes);
	float ms = 0.0f;
	if(err == cudaSuccess) {  // L1
		const int grid = sms * 8, iters = 2048;
		err = time_ms([&] { k_peak_l1<<<grid, 256>>>(buf, iters, sink); }, 5, &ms);
		out->l1_gbs = (float)((double)grid * 256 * 4 * 16 * iters / (ms * 1e-3) / 1e9);
	}
	if(err == cudaSuccess) {  // L2
		const int passes = 40;
		const size_t n = l2Bytes / 16;
		const int grid = sms * 8;
		const size_t per = (size_t)grid * 256 * 4;
		const size_t covered = n / per * per;  // elements the strided loop really touches
		err = time_ms([&] { k_peak_stream<<<grid, 256>>>(buf, n, passes, sink); }, 5, &ms);
		out->l2_gbs = (float)((double)covered * 16 * passes / (ms * 1e-3) / 1e9);
	}
	if(err == cudaSuccess) {  // HBM
		const size_t n = hbmBytes / 16;
		const int grid = sms * 16;
		const size_t per = (size_t)grid * 256 * 4;
		const size_t covered = n / per * per;
		err = time_ms([&] { k_peak_stream<<<grid, 256>>>(buf, n, 1, sink); }, 5, &ms);
		out->hbm_read_gbs = (float)((double)covered * 16 / (ms * 1e-3) / 1e9);
	}
	if(err == cudaSuccess) {  // issue
		const int grid = sms * 8, iters = 512;
		err = time_ms([&] { k_peak_issue<<<grid, 256>>>(iters, 1.0001f, (float *)sink); }, 5, &ms);
		const double warpInst = (double)grid * 8 /*warps*/ * iters * 8 * 16;
		out->issue_gwarpinst_s = (float)(warpInst / (ms * 1e-3) / 1e9);
	}
	cudaFree(buf);
	cudaFree(sink);
	if(err != cudaSuccess) {
		snl::set_error("snl_measure_peaks: %s", cudaGetErrorString(err));
		return SNL_ERR_CUDA;
	}
	// 4 schedulers per SM, one warp instruction per cycle each, at the maximum SM clock
	out->issue_peak_gwarpinst_s = (float)((double)sms * 4 * clockKHz * 1e3 / 1e9);
	return SNL_OK;
}
