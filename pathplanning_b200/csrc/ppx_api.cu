// libppx.so: context, memory helpers and world-model upload.
#include <algorithm>
#include <cstdarg>
#include <cstring>
#include "ppx_ctx.h"

int ppx_fail(ppx_ctx* ctx, int code, const char* fmt, ...)
{
	char buf[512];
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(buf, sizeof(buf), fmt, ap);
	va_end(ap);
	if (ctx)
		ctx->error = buf;
	else
		fprintf(stderr, "ppx: %s\n", buf);
	return code;
}

int ppx_scratch(ppx_ctx* ctx, int slot, size_t bytes, void** out)
{
	auto& s = ctx->scratch[slot];
	if (s.bytes < bytes) {
		if (s.ptr) {
			PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
			PPX_CUDA(ctx, cudaFree(s.ptr));
			s.ptr = nullptr;
			s.bytes = 0;
		}
		size_t want = bytes + bytes / 4 + 256;
		PPX_CUDA(ctx, cudaMalloc(&s.ptr, want));
		s.bytes = want;
	}
	*out = s.ptr;
	return PPX_OK;
}

int ppx_pipeline_init(ppx_ctx* ctx)
{
	if (ctx->s_h2d)
		return PPX_OK;
	PPX_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->s_h2d, cudaStreamNonBlocking));
	PPX_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->s_d2h, cudaStreamNonBlocking));
	PPX_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_start, cudaEventDisableTiming));
	for (int i = 0; i < ppx_ctx::kMaxChunks; i++) {
		PPX_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_h2d[i], cudaEventDisableTiming));
		PPX_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_k[i], cudaEventDisableTiming));
	}
	return PPX_OK;
}

namespace {
// GVD::ObstacleDistanceMap::GetDistanceToNearestObstacle, gvd.h:38:
//   std::sqrt(m_distance[row][col]) * resolution   -- sqrt(int) is the double overload, the
//   product is double*float->double, the return type rounds to float.
__global__ void k_clearance(const int* __restrict__ sqdist, float* __restrict__ clearance, size_t n, float res)
{
	size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
	size_t stride = (size_t)gridDim.x * blockDim.x;
	for (; i < n; i += stride)
		clearance[i] = (float)(sqrt((double)sqdist[i]) * (double)res);
}

// ObstacleListOccupancyMap::IsOccupied, obstacle_list_occupancy_map.cpp:63-69
__global__ void k_occupied(const int* __restrict__ occupancy, uint8_t* __restrict__ occ8, size_t n)
{
	size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
	size_t stride = (size_t)gridDim.x * blockDim.x;
	for (; i < n; i += stride)
		occ8[i] = occupancy[i] >= 0 ? 1 : 0;
}
}

// (Re)allocates the world-model grids for desc's size; a new size resets to the defaults of an empty
// map: free cells (-1), zero potential.
int ppx_map_alloc(ppx_ctx* ctx, const ppx_map_desc* desc)
{
	size_t cells = (size_t)desc->rows * desc->cols;
	if (cells == ctx->cells)
		return PPX_OK;
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	cudaFree(ctx->d_sqdist);
	cudaFree(ctx->d_occupancy);
	cudaFree(ctx->d_pathcost);
	cudaFree(ctx->d_clearance);
	cudaFree(ctx->d_occ8);
	ctx->d_occ8 = nullptr;
	ctx->d_sqdist = ctx->d_occupancy = nullptr;
	ctx->d_pathcost = ctx->d_clearance = nullptr;
	ctx->cells = 0;
	ctx->has_map = false;
	PPX_CUDA(ctx, cudaMalloc(&ctx->d_sqdist, cells * sizeof(int32_t)));
	PPX_CUDA(ctx, cudaMalloc(&ctx->d_occupancy, cells * sizeof(int32_t)));
	PPX_CUDA(ctx, cudaMalloc(&ctx->d_pathcost, cells * sizeof(float)));
	PPX_CUDA(ctx, cudaMalloc(&ctx->d_clearance, cells * sizeof(float)));
	PPX_CUDA(ctx, cudaMalloc(&ctx->d_occ8, cells));
	PPX_CUDA(ctx, cudaMemsetAsync(ctx->d_occ8, 0, cells, ctx->stream));
	PPX_CUDA(ctx, cudaMemsetAsync(ctx->d_occupancy, 0xff, cells * sizeof(int32_t), ctx->stream));
	PPX_CUDA(ctx, cudaMemsetAsync(ctx->d_pathcost, 0, cells * sizeof(float), ctx->stream));
	ctx->cells = cells;
	return PPX_OK;
}

// Derived grids (clearance from sqdist, occ8 from occupancy) after the device copies changed.
int ppx_map_derive(ppx_ctx* ctx, const ppx_map_desc* desc, bool newSqdist, bool newOccupancy)
{
	size_t cells = (size_t)desc->rows * desc->cols;
	ctx->desc = *desc;
	int threads = 256;
	int blocks = (int)std::min<size_t>((cells + threads - 1) / threads, (size_t)ctx->num_sms * 8);
	if (newSqdist) {
		k_clearance<<<blocks, threads, 0, ctx->stream>>>(ctx->d_sqdist, ctx->d_clearance, cells, desc->resolution);
		PPX_LAUNCHED(ctx);
	}
	if (newOccupancy) {
		k_occupied<<<blocks, threads, 0, ctx->stream>>>(ctx->d_occupancy, ctx->d_occ8, cells);
		PPX_LAUNCHED(ctx);
	}
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	ctx->has_map = true;
	return PPX_OK;
}

extern "C" {

int ppx_version(void) { return PPX_VERSION; }

int ppx_ctx_create(int device, ppx_ctx** out)
{
	if (!out)
		return PPX_ERR_INVALID;
	*out = nullptr;
	int count = 0;
	cudaError_t e = cudaGetDeviceCount(&count);
	if (e != cudaSuccess || count <= 0)
		return ppx_fail(nullptr, PPX_ERR_CUDA, "no CUDA device (%s); libppx has no CPU fallback", cudaGetErrorString(e));
	if (device < 0 || device >= count)
		return ppx_fail(nullptr, PPX_ERR_INVALID, "device %d out of range (%d devices)", device, count);
	auto* ctx = new ppx_ctx();
	ctx->device = device;
	PPX_GUARD(ctx); // the caller's current device is restored on return
	PPX_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
	ctx->stream = ctx->own_stream;
	PPX_CUDA(ctx, cudaDeviceGetAttribute(&ctx->num_sms, cudaDevAttrMultiProcessorCount, device));
	PPX_CUDA(ctx, cudaMalloc(&ctx->d_queue, ppx_ctx::kMaxChunks * sizeof(unsigned long long)));
	PPX_CUDA(ctx, cudaMalloc(&ctx->d_rsq, ppx_ctx::kRsqSlots * sizeof(unsigned long long)));
	PPX_CUDA(ctx, cudaMemset(ctx->d_rsq, 0, ppx_ctx::kRsqSlots * sizeof(unsigned long long)));
	*out = ctx;
	return PPX_OK;
}

int ppx_ctx_destroy(ppx_ctx* ctx)
{
	if (!ctx)
		return PPX_ERR_INVALID;
	PPX_GUARD(ctx);
	cudaStreamSynchronize(ctx->stream);
	if (ctx->comm)
		ppx_comm_destroy(ctx);
	cudaFree(ctx->d_sqdist);
	cudaFree(ctx->d_occupancy);
	cudaFree(ctx->d_pathcost);
	cudaFree(ctx->d_clearance);
	cudaFree(ctx->d_occ8);
	cudaFree(ctx->d_voro_sq);
	cudaFree(ctx->d_near_obst);
	cudaFree(ctx->d_near_edge);
	cudaFree(ctx->d_queue);
	cudaFree(ctx->d_rsq);
	cudaFree(ctx->d_nonholo);
	cudaFree(ctx->d_wf_cost);
	cudaFree(ctx->d_wf_explored);
	cudaFree(ctx->d_wf_goals);
	for (auto& s : ctx->scratch)
		cudaFree(s.ptr);
	if (ctx->s_h2d) {
		cudaStreamDestroy(ctx->s_h2d);
		cudaStreamDestroy(ctx->s_d2h);
		cudaEventDestroy(ctx->ev_start);
		for (int i = 0; i < ppx_ctx::kMaxChunks; i++) {
			cudaEventDestroy(ctx->ev_h2d[i]);
			cudaEventDestroy(ctx->ev_k[i]);
		}
	}
	cudaStreamDestroy(ctx->own_stream);
	delete ctx;
	return PPX_OK;
}

int ppx_ctx_set_stream(ppx_ctx* ctx, void* cuda_stream)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
	return PPX_OK;
}

int ppx_ctx_synchronize(ppx_ctx* ctx)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

int ppx_ctx_set_option(ppx_ctx* ctx, const char* name, int64_t value)
{
	PPX_GUARD(ctx);
	if (!ctx || !name)
		return PPX_ERR_INVALID;
	if (!strcmp(name, "rs_filter")) {
		ctx->rs_filter = value != 0;
		return PPX_OK;
	}
	if (!strcmp(name, "rs_overlap")) {
		PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
		PPX_CUDA(ctx, cudaMemset(ctx->d_rsq, 0, ppx_ctx::kRsqSlots * sizeof(unsigned long long)));
		ctx->rs_slot = 0;
		ctx->rs_overlap = value != 0;
		return PPX_OK;
	}
	if (!strcmp(name, "rs_engine")) {
		if (value < 0 || value > 2)
			return ppx_fail(ctx, PPX_ERR_INVALID, "rs_engine must be 0 (k_check_rs), 1 (k_check_rs2) or 2 (1 + quad solver for ppx_rs_solve)");
		ctx->rs_engine = (int)value;
		return PPX_OK;
	}
	if (!strcmp(name, "search_engine")) {
		if (value != 0 && value != 1)
			return ppx_fail(ctx, PPX_ERR_INVALID, "search_engine must be 0 (warp per query) or 1 (CTA per query)");
		ctx->search_engine = (int)value;
		return PPX_OK;
	}
	return ppx_fail(ctx, PPX_ERR_INVALID, "unknown option %s", name);
}

const char* ppx_last_error(const ppx_ctx* ctx) { return ctx ? ctx->error.c_str() : "null context"; }

int64_t ppx_ctx_launch_count(const ppx_ctx* ctx) { return ctx ? ctx->launches : 0; }

int ppx_dev_alloc(ppx_ctx* ctx, uint64_t bytes, void** dptr)
{
	PPX_GUARD(ctx);
	if (!ctx || !dptr)
		return PPX_ERR_INVALID;
	PPX_CUDA(ctx, cudaMalloc(dptr, bytes ? bytes : 1));
	return PPX_OK;
}

int ppx_dev_free(ppx_ctx* ctx, void* dptr)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	PPX_CUDA(ctx, cudaFree(dptr));
	return PPX_OK;
}

int ppx_host_alloc(ppx_ctx* ctx, uint64_t bytes, void** hptr)
{
	PPX_GUARD(ctx);
	if (!ctx || !hptr)
		return PPX_ERR_INVALID;
	PPX_CUDA(ctx, cudaHostAlloc(hptr, bytes ? bytes : 1, cudaHostAllocDefault));
	return PPX_OK;
}

int ppx_host_free(ppx_ctx* ctx, void* hptr)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	PPX_CUDA(ctx, cudaFreeHost(hptr));
	return PPX_OK;
}

int ppx_memcpy_h2d(ppx_ctx* ctx, void* dst, const void* src, uint64_t bytes)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	PPX_CUDA(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

int ppx_memcpy_d2h(ppx_ctx* ctx, void* dst, const void* src, uint64_t bytes)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	PPX_CUDA(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

int ppx_map_upload(ppx_ctx* ctx, const ppx_map_desc* desc, const int32_t* sqdist, const int32_t* occupancy,
	const float* pathcost)
{
	PPX_GUARD(ctx);
	if (!ctx || !desc)
		return PPX_ERR_INVALID;
	PPX_REQUIRE(ctx, desc->rows > 0 && desc->cols > 0, "map size must be positive");
	PPX_REQUIRE(ctx, desc->resolution > 0.0f, "resolution must be positive");
	size_t cells = (size_t)desc->rows * desc->cols;
	if (cells != ctx->cells && !sqdist)
		return ppx_fail(ctx, PPX_ERR_INVALID, "first upload of a map of this size needs sqdist");
	PPX_CHECK(ppx_map_alloc(ctx, desc));
	if (sqdist || occupancy)
		ctx->has_gvd = false; // the device-built Voronoi grids describe the previous map
	if (sqdist)
		PPX_CUDA(ctx, cudaMemcpyAsync(ctx->d_sqdist, sqdist, cells * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
	if (occupancy)
		PPX_CUDA(ctx, cudaMemcpyAsync(ctx->d_occupancy, occupancy, cells * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
	if (pathcost)
		PPX_CUDA(ctx, cudaMemcpyAsync(ctx->d_pathcost, pathcost, cells * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
	else if (sqdist || occupancy) // a potential belongs to the map it was computed for: a new map without one searches with zero
		PPX_CUDA(ctx, cudaMemsetAsync(ctx->d_pathcost, 0, cells * sizeof(float), ctx->stream));
	return ppx_map_derive(ctx, desc, sqdist != nullptr, occupancy != nullptr);
}

} // extern "C"
