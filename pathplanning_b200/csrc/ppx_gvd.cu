// libppx.so: obstacle squared-distance map on the device.
//
// Replaces the cold build of GVD::ObstacleDistanceMap (state_validator/gvd.cpp:30-89,105-118): the
// reference propagates nearest-obstacle cells through 8-neighbours from a priority queue
// (brushfire) and stores the SQUARED CELL DISTANCE to the nearest occupied cell as int
// (m_distance, INT_MAX when there is no obstacle).  The device computes the exact squared
// Euclidean distance transform in two separable integer passes; the brushfire's result equals it
// on every map of the parity tests (SURVEY trap H14 records that this equality is empirical for the
// reference), and tests/test_gvd_gpu.py compares bit for bit with the compiled reference.
//
//   pass 1  k_edt_columns: g[r][c] = distance in cells from (r, c) to the nearest occupied cell of
//           column c (thread per column, coalesced up/down scans)
//   pass 2  k_edt_rows:    d2[r][c] = min over c' of (c - c')^2 + g[r][c']^2  (CTA per row, g-row in
//           shared memory, outward search that stops when (c - c')^2 >= best)
#include <algorithm>
#include "ppx_ctx.h"

namespace {

constexpr int kNone = 1 << 29; // "no occupied cell in this column"

__global__ void __launch_bounds__(128) k_edt_columns(const int* __restrict__ occupancy, int rows, int cols, int* __restrict__ g)
{
	int c = blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= cols)
		return;
	int d = kNone;
	for (int r = 0; r < rows; r++) { // nearest occupied cell above (smaller row)
		bool occ = occupancy[(size_t)r * cols + c] >= 0; // IsOccupied, obstacle_list_occupancy_map.cpp:63-69
		d = occ ? 0 : (d >= kNone ? kNone : d + 1);
		g[(size_t)r * cols + c] = d;
	}
	d = kNone;
	for (int r = rows - 1; r >= 0; r--) { // ... and below
		size_t i = (size_t)r * cols + c;
		int up = g[i];
		d = up == 0 ? 0 : (d >= kNone ? kNone : d + 1);
		g[i] = min(up, d);
	}
}

__global__ void __launch_bounds__(256) k_edt_rows(const int* __restrict__ g, int rows, int cols, int* __restrict__ sqdist)
{
	extern __shared__ int s_g[];
	for (int r = blockIdx.x; r < rows; r += gridDim.x) {
		__syncthreads();
		for (int c = threadIdx.x; c < cols; c += blockDim.x)
			s_g[c] = g[(size_t)r * cols + c];
		__syncthreads();
		for (int c = threadIdx.x; c < cols; c += blockDim.x) {
			long long best = 0x7fffffffLL; // INT_MAX: no obstacle anywhere (gvd.cpp:22)
			int gc = s_g[c];
			if (gc < kNone)
				best = (long long)gc * gc;
			for (int d = 1; d < cols; d++) {
				long long dd = (long long)d * d;
				if (dd >= best)
					break;
				int lo = c - d, hi = c + d;
				if (lo < 0 && hi >= cols)
					break;
				if (lo >= 0) {
					int v = s_g[lo];
					if (v < kNone)
						best = min(best, dd + (long long)v * v);
				}
				if (hi < cols) {
					int v = s_g[hi];
					if (v < kNone)
						best = min(best, dd + (long long)v * v);
				}
			}
			sqdist[(size_t)r * cols + c] = (int)min(best, 0x7fffffffLL);
		}
	}
}

}

extern "C" int ppx_map_build(ppx_ctx* ctx, const ppx_map_desc* desc, const int32_t* occupancy, const float* pathcost,
	int32_t* sqdist_out)
{
	if (!ctx || !desc)
		return PPX_ERR_INVALID;
	PPX_REQUIRE(ctx, occupancy, "ppx_map_build needs the occupancy grid");
	PPX_REQUIRE(ctx, desc->rows > 0 && desc->cols > 0 && desc->rows < 32768 && desc->cols < 32768, "map size out of range");
	PPX_REQUIRE(ctx, desc->resolution > 0.0f, "resolution must be positive");
	PPX_REQUIRE(ctx, (size_t)desc->cols * sizeof(int) <= 160 * 1024, "map too wide for the row pass (cols <= 40960)");
	PPX_CHECK(ppx_map_alloc(ctx, desc));
	const size_t cells = (size_t)desc->rows * desc->cols;
	int* d_g;
	PPX_CHECK(ppx_scratch(ctx, 6, cells * sizeof(int), (void**)&d_g));
	PPX_CUDA(ctx, cudaMemcpyAsync(ctx->d_occupancy, occupancy, cells * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
	k_edt_columns<<<(desc->cols + 127) / 128, 128, 0, ctx->stream>>>(ctx->d_occupancy, desc->rows, desc->cols, d_g);
	PPX_LAUNCHED(ctx);
	const size_t smem = (size_t)desc->cols * sizeof(int);
	PPX_CUDA(ctx, cudaFuncSetAttribute(k_edt_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
	int blocks = std::min(desc->rows, ctx->num_sms * 4);
	k_edt_rows<<<blocks, 256, smem, ctx->stream>>>(d_g, desc->rows, desc->cols, ctx->d_sqdist);
	PPX_LAUNCHED(ctx);
	if (pathcost)
		PPX_CUDA(ctx, cudaMemcpyAsync(ctx->d_pathcost, pathcost, cells * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
	if (sqdist_out)
		PPX_CUDA(ctx, cudaMemcpyAsync(sqdist_out, ctx->d_sqdist, cells * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
	return ppx_map_derive(ctx, desc, true, true);
}
