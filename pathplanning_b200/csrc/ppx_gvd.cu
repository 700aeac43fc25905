// libppx.so: obstacle squared-distance map on the device and obstacle outline rasterisation (N4).
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

// N4: Shape::RasterizeLine (state_validator/obstacle.cpp:7-61) for every edge of every polygon outline
// (PolygonShape::GetGridCellsPosition, :77-85) and the occupancy write of ObstacleListOccupancyMap::AddObstacle
// (obstacle_list_occupancy_map.cpp:29-45).  One thread per outline vertex v rasterises the edge v -> next(v).
// The reference adds obstacles one after the other, so a cell shared by several outlines keeps the id of the
// LAST one; ids grow with the insertion order (FindSmallestIDAvailable on a map without gaps), hence atomicMax.
__global__ void __launch_bounds__(128) k_rasterize(int rows, int cols, double ox, double oy, double resd, int nPolys,
	const int* __restrict__ offsets, const double* __restrict__ vertices, int firstId, int* __restrict__ occupancy)
{
	const int nVerts = offsets[nPolys];
	const int v = blockIdx.x * blockDim.x + threadIdx.x;
	if (v >= nVerts)
		return;
	int lo = 0, hi = nPolys; // polygon of vertex v: the last p with offsets[p] <= v
	while (hi - lo > 1) {
		int mid = (lo + hi) >> 1;
		if (offsets[mid] <= v)
			lo = mid;
		else
			hi = mid;
	}
	const int p = lo, first = offsets[p], count = offsets[p + 1] - first;
	if (count <= 0)
		return;
	const int w = first + (v - first + 1) % count;
	// WorldPositionToGridCell(p, false), occupancy_map.h:106-117,180-183
	int x0 = ppx::trunc_i32((vertices[2 * v] - ox) / resd), y0 = ppx::trunc_i32((vertices[2 * v + 1] - oy) / resd);
	int x1 = ppx::trunc_i32((vertices[2 * w] - ox) / resd), y1 = ppx::trunc_i32((vertices[2 * w + 1] - oy) / resd);
	const bool steep = abs(y1 - y0) > abs(x1 - x0);
	if (steep) {
		int t = x0; x0 = y0; y0 = t;
		t = x1; x1 = y1; y1 = t;
	}
	if (x0 > x1) {
		int t = x0; x0 = x1; x1 = t;
		t = y0; y0 = y1; y1 = t;
	}
	const int dx = x1 - x0, dy = abs(y1 - y0), ystep = y0 < y1 ? 1 : -1;
	int err = dx / 2, y = y0;
	const int id = firstId + p;
	for (int x = x0; x <= x1; x++) {
		const int r = steep ? y : x, c = steep ? x : y;
		if (r >= 0 && r < rows && c >= 0 && c < cols) {
			if (firstId >= 0)
				atomicMax(&occupancy[(size_t)r * cols + c], id);
			else
				occupancy[(size_t)r * cols + c] = -1; // RemoveObstacle (obstacle_list_occupancy_map.cpp:47-61): cleared whoever owns it
		}
		err -= dy;
		if (err < 0) {
			y += ystep;
			err += dx;
		}
	}
}

}

extern "C" int ppx_map_rasterize(ppx_ctx* ctx, const ppx_map_desc* desc, int32_t n_polygons, const int32_t* offsets,
	const double* vertices, int32_t first_id, int32_t* occupancy)
{
	if (!ctx || !desc)
		return PPX_ERR_INVALID;
	PPX_REQUIRE(ctx, n_polygons >= 0 && offsets && occupancy && first_id >= -1, "bad arguments");
	PPX_REQUIRE(ctx, desc->rows > 0 && desc->cols > 0 && desc->resolution > 0.0f, "bad map description");
	if (n_polygons == 0)
		return PPX_OK;
	const int nVerts = offsets[n_polygons];
	PPX_REQUIRE(ctx, nVerts > 0 && vertices, "polygons without vertices");
	const size_t cells = (size_t)desc->rows * desc->cols;
	size_t off = 0;
	auto carve = [&](size_t bytes) {
		size_t o = off;
		off += (bytes + 255) & ~(size_t)255;
		return o;
	};
	const size_t oOcc = carve(cells * sizeof(int32_t)), oOff = carve((size_t)(n_polygons + 1) * sizeof(int32_t)),
				 oVer = carve((size_t)nVerts * 2 * sizeof(double));
	char* base;
	PPX_CHECK(ppx_scratch(ctx, 6, off, (void**)&base));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oOcc, occupancy, cells * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oOff, offsets, (size_t)(n_polygons + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oVer, vertices, (size_t)nVerts * 2 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
	k_rasterize<<<(nVerts + 127) / 128, 128, 0, ctx->stream>>>(desc->rows, desc->cols, desc->world_origin[0], desc->world_origin[1],
		(double)desc->resolution, n_polygons, (const int*)(base + oOff), (const double*)(base + oVer), first_id, (int*)(base + oOcc));
	PPX_LAUNCHED(ctx);
	PPX_CUDA(ctx, cudaMemcpyAsync(occupancy, base + oOcc, cells * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
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
