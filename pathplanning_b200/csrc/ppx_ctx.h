// Host-side context shared by the translation units of libppx.so.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>
#include "ppx.h"
#include "ppx_device.cuh"

struct ppx_ctx {
	int device = 0;
	cudaStream_t own_stream = nullptr;
	cudaStream_t stream = nullptr;
	int num_sms = 0;
	std::string error;
	int64_t launches = 0;
	int rs_engine = 1; // batched Reeds-Shepp edge check: 1 = k_check_rs2 (knot records, bulk-copy staging), 0 = k_check_rs
	int rs_filter = 1; // 0: k_check_rs evaluates every sample with the exact FP64 path (A/B tests)
	int rs_overlap = 0; // 1: consecutive ppx_check_edges_rs_dev launches overlap (programmatic dependent launch)
	int rs_slot = 0; // cursor slot of the next overlapped launch
	int search_engine = 1; // Hybrid A* graph search: 1 = one CTA per query (master + worker warps), 0 = one warp per query
	// cursors of the overlapped launches, a ring: each launch zeroes its successor's.  More slots than kernels can be resident
	// at once (128), so a slot is never reused while an earlier launch still reads it, however small the launches are
	static constexpr int kRsqSlots = 256;
	unsigned long long* d_rsq = nullptr;

	// multi-GPU (ppx_comm.cu): ncclComm_t of this rank, NULL without ppx_comm_init
	void* comm = nullptr;
	int comm_rank = 0, comm_world = 1;

	// world model
	bool has_map = false;
	ppx_map_desc desc {};
	int32_t* d_sqdist = nullptr;
	int32_t* d_occupancy = nullptr;
	float* d_pathcost = nullptr;
	float* d_clearance = nullptr;
	uint8_t* d_occ8 = nullptr; // occupancy >= 0 (IsOccupied), one byte per cell
	size_t cells = 0;
	// the rest of the GVD when it was built on the device (ppx_gvd_build): squared distance to the nearest Voronoi
	// edge cell, nearest obstacle / edge cells packed row << 16 | col (-1 none)
	bool has_gvd = false;
	size_t gvd_cells = 0;
	int* d_voro_sq = nullptr;
	int* d_near_obst = nullptr;
	int* d_near_edge = nullptr;

	// grow-only device scratch for the host-buffer entry points
	struct Scratch {
		void* ptr = nullptr;
		size_t bytes = 0;
	};
	Scratch scratch[8];
	unsigned long long* d_queue = nullptr; // work-queue cursors of the lane-refill kernels (kMaxChunks of them)
	// copy/compute pipeline of the host-buffer entry points: H2D and D2H run on their own streams
	// (the two copy engines) and are chained to the kernels on `stream` by events, chunk by chunk
	static constexpr int kMaxChunks = 16;
	cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
	cudaEvent_t ev_h2d[kMaxChunks] {}, ev_k[kMaxChunks] {}, ev_start = nullptr;

	// non-holonomic table (A9)
	bool has_nonholo = false;
	double* d_nonholo = nullptr;

	// resident wavefronts (A8)
	int32_t wf_slots = 0;
	float* d_wf_cost = nullptr;
	uint8_t* d_wf_explored = nullptr;
	double* d_wf_goals = nullptr;
	int* d_wf_status = nullptr; // device status word of the last wavefront launch (points into scratch)

	ppx::MapDev map_dev() const
	{
		ppx::MapDev m;
		m.rows = desc.rows;
		m.cols = desc.cols;
		m.res = desc.resolution;
		m.resd = (double)desc.resolution;
		m.ox = desc.world_origin[0];
		m.oy = desc.world_origin[1];
		m.lox = desc.local_origin[0];
		m.loy = desc.local_origin[1];
		m.lbx = desc.lb[0]; m.lby = desc.lb[1]; m.lbt = desc.lb[2];
		m.ubx = desc.ub[0]; m.uby = desc.ub[1]; m.ubt = desc.ub[2];
		m.clearance = d_clearance;
		m.occupancy = d_occupancy;
		m.pathcost = d_pathcost;
		return m;
	}
};

// Every extern "C" entry point runs on its context's device whatever device is current in the calling
// thread (a process may hold one context per GPU, and torch changes the current device freely), and
// leaves the caller's current device as it found it.
struct ppx_device_guard {
	int prev = -1;
	bool changed = false;
	explicit ppx_device_guard(const ppx_ctx* ctx)
	{
		if (ctx && cudaGetDevice(&prev) == cudaSuccess && prev != ctx->device)
			changed = cudaSetDevice(ctx->device) == cudaSuccess;
	}
	~ppx_device_guard()
	{
		if (changed)
			cudaSetDevice(prev);
	}
	ppx_device_guard(const ppx_device_guard&) = delete;
	ppx_device_guard& operator=(const ppx_device_guard&) = delete;
};
#define PPX_GUARD(ctx) ppx_device_guard ppx_guard_(ctx)

int ppx_fail(ppx_ctx* ctx, int code, const char* fmt, ...);
int ppx_scratch(ppx_ctx* ctx, int slot, size_t bytes, void** out);
int ppx_pipeline_init(ppx_ctx* ctx);
int ppx_map_alloc(ppx_ctx* ctx, const ppx_map_desc* desc);
int ppx_map_derive(ppx_ctx* ctx, const ppx_map_desc* desc, bool newSqdist, bool newOccupancy);
int ppx_gvd_unpack_cells(ppx_ctx* ctx, int* d_nearObstacle, int* d_nearEdge);

#define PPX_CUDA(ctx, call)                                                                           \
	do {                                                                                              \
		cudaError_t e_ = (call);                                                                      \
		if (e_ != cudaSuccess)                                                                        \
			return ppx_fail(ctx, PPX_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
	} while (0)

#define PPX_CHECK(expr)      \
	do {                     \
		int rc_ = (expr);    \
		if (rc_ != PPX_OK)   \
			return rc_;      \
	} while (0)

#define PPX_REQUIRE(ctx, cond, msg) \
	do {                            \
		if (!(cond))                \
			return ppx_fail(ctx, PPX_ERR_INVALID, "%s", msg); \
	} while (0)

#define PPX_LAUNCHED(ctx)                                                                             \
	do {                                                                                              \
		(ctx)->launches++;                                                                            \
		cudaError_t e_ = cudaGetLastError();                                                          \
		if (e_ != cudaSuccess)                                                                        \
			return ppx_fail(ctx, PPX_ERR_CUDA, "%s:%d launch: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
	} while (0)
