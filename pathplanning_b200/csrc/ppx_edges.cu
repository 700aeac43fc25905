// libppx.so: batched StateValidatorOccupancyMap::IsStateValid / IsPathValid (A7) over
// Reeds-Shepp, constant-steer and straight edges, and the batched Reeds-Shepp solver (A3).
//
// Mapping.  k_check_rs2 (ppx_check_rs2.cuh, the default) / k_check_rs: every lane of a warp runs the sphere-tracing loop of
// one edge at a time in FP32 with certified cells (ppx_fastlane.cuh) and picks up a new edge when it finishes, so a warp
// stays on one instruction stream although edges need 1 to ~200 dependent samples; uncertified samples are evaluated by
// the exact FP64 restatement of the reference.  k_check_rs2 stages the 72-byte records of 32 edges per warp with one
// bulk async copy (cp.async.bulk + mbarrier) and turns them into FP32 knot records with all lanes active.
// k_check_csteer / k_check_line / k_rs_solve / k_rs_interpolate: one item per thread, persistent grid.
#include <algorithm>
#include <cstring>
#include "ppx_ctx.h"
#include "ppx_fastlane.cuh"

using namespace ppx;

namespace {

constexpr int kThreads = 128;

__device__ __forceinline__ void load_rs_edge(const ppx_rs_edge* __restrict__ e, Pose& start, RSSeg& seg)
{
	start.x = e->start[0];
	start.y = e->start[1];
	start.th = e->start[2];
	#pragma unroll
	for (int i = 0; i < 5; i++)
		seg.len[i] = e->len[i];
	seg.code = (uint32_t)e->code;
}

__global__ void __launch_bounds__(kThreads) k_is_state_valid(MapDev map, int64_t n, const double* __restrict__ poses,
	float minSafeRadius, uint8_t* __restrict__ valid)
{
	int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
	int64_t stride = (int64_t)gridDim.x * blockDim.x;
	for (; i < n; i += stride) {
		float d;
		valid[i] = is_state_valid(map, poses[3 * i], poses[3 * i + 1], poses[3 * i + 2], minSafeRadius, &d) ? 1 : 0;
	}
}

// Per-thread column of the CTA's shared-memory copy of the motion lengths: the lane machine
// indexes them with a data-dependent motion number, which in registers would spill to local memory.
struct SmemLen {
	const double* col; // &s_len[0][threadIdx.x]
	__device__ __forceinline__ double get(int i) const { return col[i * kThreads]; }
};

// A7 on Reeds-Shepp edges, re-converging + filtered form.  Every lane runs the RSFastLane micro-step
// machine (ppx_fastlane.cuh: FP32 positions, certified cells) on its own edge; finished lanes are
// refilled by warp vote from the warp's chunk of the edge queue (chunks of kGrab edges are taken from
// one global counter), so the 32 lanes of a warp stay on one instruction stream although edges need
// between 1 and ~200 samples.  Samples the filter cannot certify PARK their lane; when kParkBatch
// lanes of the warp are parked (or nothing else is runnable) the warp evaluates those samples with
// the exact FP64 restatement of the reference (rs_exact_sample) in one divergent pass.
constexpr int kGrab = 128;
constexpr int kOverlapCtasPerSm = 2; // per launch in rs_overlap mode, see launch_check_rs
constexpr int kCheckCtasPerSm = 6; // 80 registers (k_check_rs<0,6>: 32 B of spill stores, ppx_edges.ptxas.log); 24 warps/SM measured best (r1g sweep)
constexpr int kParkBatch = 8;
constexpr int kRefillBatch = 8;

struct AuditOut {
	unsigned long long certified; // samples decided by the filter
	unsigned long long parked; // samples decided by the exact path
	unsigned long long cell_mismatch; // certified samples whose exact cell differs (must be 0)
	unsigned int max_dev_bits; // max |fast - exact| position, cells, as float bits
};

// State of IsPathValid's loop that the exact path reads and writes, passed by value so that the
// lane machine stays in registers across the out-of-line call.
struct ExactIO {
	double pathLength, length, lastValidLength;
	unsigned ns;
	float last;
	bool verdict, done;
};

__device__ __noinline__ ExactIO exact_sample_lane(ExactIO io, uint32_t code, const double* __restrict__ lencol, MapDev map,
	const ppx_rs_edge* __restrict__ e, double r, float minSafeRadius, float minStep)
{
	Pose start = { e->start[0], e->start[1], e->start[2] };
	RSSeg seg;
	#pragma unroll
	for (int i = 0; i < 5; i++)
		seg.len[i] = lencol[i * kThreads];
	seg.code = code;
	double ex, ey;
	io.done = rs_exact_sample(io, map, start, seg, r, minSafeRadius, minStep, &ex, &ey);
	return io;
}

// audit only: exact position of the sample the lane has just certified, compared with (px, py)
__device__ __noinline__ void audit_sample(const RSFastLane<SmemLen>& lane, const SmemLen& L, const MapDev& map,
	const FastMap& fm, const ppx_rs_edge* __restrict__ e, double r, double lengthAtSample, AuditOut* out)
{
	Pose start = { e->start[0], e->start[1], e->start[2] };
	RSSeg seg;
	for (int i = 0; i < 5; i++)
		seg.len[i] = L.get(i);
	seg.code = lane.code;
	RSWalker w;
	w.init(start);
	Pose p = w.at(seg, r, (lengthAtSample / lane.pathLength) * lane.pathLength);
	int row = trunc_i32((p.x - map.ox) / map.resd), col = trunc_i32((p.y - map.oy) / map.resd);
	int frow = lane.rowb + (int)floorf(lane.px), fcol = lane.colb + (int)floorf(lane.py);
	double dx = (p.x - map.ox) / map.resd - (double)lane.rowb - (double)lane.px;
	double dy = (p.y - map.oy) / map.resd - (double)lane.colb - (double)lane.py;
	float dev = (float)fmax(fabs(dx), fabs(dy));
	atomicMax(&out->max_dev_bits, __float_as_uint(dev));
	atomicAdd(&out->certified, 1ull);
	if (row != frow || col != fcol)
		atomicAdd(&out->cell_mismatch, 1ull);
}

template <bool kAudit, int kCtas>
__global__ void __launch_bounds__(kThreads, kCtas) k_check_rs(MapDev map, FastMap fm, int64_t n,
	const ppx_rs_edge* __restrict__ edges, double r, float minSafeRadius, float minStep, uint8_t* __restrict__ valid,
	float* __restrict__ last, uint32_t* __restrict__ samples, unsigned long long* __restrict__ queue, AuditOut* audit,
	unsigned long long* nextQueue)
{
	__shared__ double s_len[5][kThreads];
	if (nextQueue) { // overlapped launches ("rs_overlap"): zero the successor's cursor, then let the successor in
		if (blockIdx.x == 0 && threadIdx.x == 0) {
			atomicExch(nextQueue, 0ull);
			__threadfence();
		}
		__syncthreads();
		asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
	}
	const unsigned lane_id = threadIdx.x & 31;
	const unsigned lt_mask = (1u << lane_id) - 1;
	double* mycol = &s_len[0][threadIdx.x];
	SmemLen L { mycol };
	RSFastLane<SmemLen> lane;
	int state = LANE_IDLE;
	int64_t idx = 0;
	int64_t cursor = 0, limit = 0; // warp-uniform
	bool exhausted = false; // warp-uniform
	const float inv_r = (float)(1.0 / r);
	const float rc = (float)(r * fm.inv_resd);

	while (true) {
		unsigned idle = __ballot_sync(0xffffffffu, state == LANE_IDLE);
		unsigned parked = __ballot_sync(0xffffffffu, state == LANE_PARKED);
		unsigned running = ~(idle | parked);
		if (!exhausted && (__popc(idle) >= kRefillBatch || !running)) {
			if (cursor >= limit) {
				unsigned long long base = 0;
				if (lane_id == 0)
					base = atomicAdd(queue, (unsigned long long)kGrab);
				base = __shfl_sync(0xffffffffu, base, 0);
				cursor = (int64_t)base;
				limit = cursor + kGrab < n ? cursor + kGrab : n;
				if (cursor >= n) {
					exhausted = true;
					limit = cursor;
				}
			}
			int64_t avail = limit - cursor;
			int rank = __popc(idle & lt_mask);
			if (state == LANE_IDLE && rank < avail) {
				idx = cursor + rank;
				const ppx_rs_edge* e = edges + idx;
				Pose start = { e->start[0], e->start[1], e->start[2] };
				#pragma unroll
				for (int i = 0; i < 5; i++)
					mycol[i * kThreads] = e->len[i];
				lane.begin(L, fm, map, start, (uint32_t)e->code, r);
				state = lane.slowOnly ? LANE_PARKED : LANE_ACTIVE;
			}
			int need = __popc(idle);
			cursor += need < avail ? need : avail;
			idle = __ballot_sync(0xffffffffu, state == LANE_IDLE);
			parked = __ballot_sync(0xffffffffu, state == LANE_PARKED);
			running = ~(idle | parked);
		}
		if (parked && (__popc(parked) >= kParkBatch || !running)) {
			if (state == LANE_PARKED) {
				ExactIO io { lane.pathLength, lane.length, lane.lastValidLength, lane.ns, 0.0f, false, false };
				io = exact_sample_lane(io, lane.code, mycol, map, edges + idx, r, minSafeRadius, minStep);
				lane.length = io.length;
				lane.lastValidLength = io.lastValidLength;
				lane.ns = io.ns;
				lane.verdict = io.verdict;
				lane.last = io.last;
				if (kAudit)
					atomicAdd(&audit->parked, 1ull);
				state = LANE_ACTIVE;
				if (io.done) {
					valid[idx] = lane.verdict ? 1 : 0;
					last[idx] = lane.last;
					if (samples)
						samples[idx] = lane.ns;
					state = LANE_IDLE;
				} else if (lane.slowOnly)
					state = LANE_PARKED;
			}
			continue;
		}
		if (!running) {
			if (exhausted)
				break;
			continue;
		}
		if (state == LANE_ACTIVE) {
			if (lane.advance(L, r, inv_r, rc, fm.inv_res)) {
				double lengthAtSample = lane.length;
				int rcode = lane.sample(map, fm, minSafeRadius, minStep);
				if (kAudit && rcode != 2)
					audit_sample(lane, L, map, fm, edges + idx, r, lengthAtSample, audit);
				if (rcode == 2)
					state = LANE_PARKED;
				else if (rcode == 1) {
					valid[idx] = lane.verdict ? 1 : 0;
					last[idx] = lane.last;
					if (samples)
						samples[idx] = lane.ns;
					state = LANE_IDLE;
				}
			}
		}
	}
	if (nextQueue) // this launch must not complete before its predecessor has (stream order for whatever follows)
		asm volatile("griddepcontrol.wait;" ::: "memory");
}

#include "ppx_check_rs2.cuh"

__global__ void __launch_bounds__(kThreads) k_check_csteer(MapDev map, int64_t n, const ppx_csteer_edge* __restrict__ edges,
	float minSafeRadius, float minStep, uint8_t* __restrict__ valid, float* __restrict__ last, double* __restrict__ finalPose)
{
	int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
	int64_t stride = (int64_t)gridDim.x * blockDim.x;
	for (; i < n; i += stride) {
		CSteer p;
		p.init(edges[i]);
		float l;
		bool ok = csteer_is_path_valid(map, p, minSafeRadius, minStep, &l, nullptr);
		valid[i] = ok ? 1 : 0;
		last[i] = l;
		if (finalPose) {
			// PathConstantSteer ctor: m_final = Interpolate(1.0); Truncate(ratio): m_final = Interpolate(ratio)
			Pose f = p.at(ok ? 1.0 : (double)l);
			finalPose[3 * i] = f.x;
			finalPose[3 * i + 1] = f.y;
			finalPose[3 * i + 2] = f.th;
		}
	}
}

__global__ void __launch_bounds__(kThreads) k_check_line(MapDev map, int64_t n, const ppx_line_edge* __restrict__ edges,
	float minSafeRadius, float minStep, uint8_t* __restrict__ valid, float* __restrict__ last)
{
	int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
	int64_t stride = (int64_t)gridDim.x * blockDim.x;
	for (; i < n; i += stride) {
		Pose a = { edges[i].from[0], edges[i].from[1], edges[i].from[2] };
		Pose b = { edges[i].to[0], edges[i].to[1], edges[i].to[2] };
		float l;
		bool ok = line_is_path_valid(map, a, b, minSafeRadius, minStep, &l);
		valid[i] = ok ? 1 : 0;
		last[i] = l;
	}
}

__device__ __forceinline__ void store_rs_edge(ppx_rs_edge* e, const Pose& start, const RSSeg& seg)
{
	e->start[0] = start.x;
	e->start[1] = start.y;
	e->start[2] = start.th;
	#pragma unroll
	for (int i = 0; i < 5; i++)
		e->len[i] = seg.len[i];
	e->code = seg.code;
}

// check != 0: fused PathConnectionReedsShepp::Connect + IsPathValid
template <int kMinBlocks, bool kCheck>
__global__ void __launch_bounds__(kThreads, kMinBlocks) k_rs_solve(MapDev map, int64_t n, const double* __restrict__ starts,
	const double* __restrict__ goals, ppx_rs_cost_params cp, int check, float minSafeRadius, float minStep,
	ppx_rs_edge* __restrict__ edges, int32_t* __restrict__ word, double* __restrict__ length, float* __restrict__ cost,
	uint8_t* __restrict__ valid, float* __restrict__ last)
{
	int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
	int64_t stride = (int64_t)gridDim.x * blockDim.x;
	for (; i < n; i += stride) {
		Pose s = { starts[3 * i], starts[3 * i + 1], starts[3 * i + 2] };
		Pose g = { goals[3 * i], goals[3 * i + 1], goals[3 * i + 2] };
		RSSolution sol;
		rs_solve(s, g, cp.min_turning_radius, cp.reverse_cost_multiplier, cp.forward_cost_multiplier,
			cp.direction_switching_cost, cp.shortest, sol);
		if (edges)
			store_rs_edge(edges + i, s, sol.seg);
		if (word)
			word[i] = sol.word;
		if (length)
			length[i] = sol.mLength * cp.min_turning_radius;
		if (cost)
			cost[i] = sol.cost;
		if (kCheck && check) {
			float l;
			bool ok = rs_is_path_valid(map, s, sol.seg, cp.min_turning_radius, minSafeRadius, minStep, &l, nullptr);
			valid[i] = ok ? 1 : 0;
			last[i] = l;
		}
	}
}

__global__ void __launch_bounds__(kThreads) k_rs_interpolate(int64_t n, const ppx_rs_edge* __restrict__ edges, double r,
	int nratios, const double* __restrict__ ratios, double* __restrict__ poses)
{
	int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
	int64_t total = n * nratios;
	int64_t stride = (int64_t)gridDim.x * blockDim.x;
	for (; idx < total; idx += stride) {
		int64_t i = idx / nratios;
		Pose start;
		RSSeg seg;
		load_rs_edge(edges + i, start, seg);
		double totalLength = rs_total_length(seg) * r;
		Pose p = start;
		if (totalLength != 0) { // path_reeds_shepp.cpp:17-18
			RSWalker w;
			w.init(start);
			p = w.at(seg, r, ratios[idx] * totalLength);
		}
		poses[3 * idx] = p.x;
		poses[3 * idx + 1] = p.y;
		poses[3 * idx + 2] = p.th;
	}
}

// Batched Solver::GetOptimalPath / GetShortestPath, eight pairs per warp (rs_solve_quad): quad k of a warp solves pair
// 8 * w + k, its first lane writes the results.
__global__ void __launch_bounds__(kThreads) k_rs_solve_quad(int64_t n, const double* __restrict__ starts, const double* __restrict__ goals,
	ppx_rs_cost_params cp, ppx_rs_edge* __restrict__ edges, int32_t* __restrict__ word, double* __restrict__ length,
	float* __restrict__ cost)
{
	const int64_t quad = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 2;
	const int64_t nquads = ((int64_t)gridDim.x * blockDim.x) >> 2;
	const int64_t rounds = (n + nquads - 1) / nquads; // all lanes of a warp take part in the quad shuffles
	for (int64_t k = 0; k < rounds; k++) {
		const int64_t i = k * nquads + quad;
		const bool have = i < n;
		Pose s = { 0, 0, 0 }, g = { 0, 0, 0 };
		if (have) {
			s = { starts[3 * i], starts[3 * i + 1], starts[3 * i + 2] };
			g = { goals[3 * i], goals[3 * i + 1], goals[3 * i + 2] };
		}
		int w;
		double t, u, v;
		rs_solve_quad(s, g, cp.min_turning_radius, cp.reverse_cost_multiplier, cp.forward_cost_multiplier, cp.direction_switching_cost,
			cp.shortest, w, t, u, v);
		if (!have || (threadIdx.x & 3) != 0)
			continue;
		RSSeg seg;
		double mLength = 0.0;
		float c = __int_as_float(0x7f800000);
		if (w >= 0) {
			mLength = rs_build_segment(w, t, u, v, seg);
			c = rs_compute_cost(seg, mLength, cp.min_turning_radius, cp.reverse_cost_multiplier, cp.forward_cost_multiplier,
				cp.direction_switching_cost);
		} else { // PathSegment(): all motions invalid, m_length 0
			for (int m = 0; m < 5; m++)
				seg.len[m] = PPX_INF;
			seg.code = 0xAAAAA;
		}
		if (edges)
			store_rs_edge(edges + i, s, seg);
		if (word)
			word[i] = w;
		if (length)
			length[i] = mLength * cp.min_turning_radius;
		if (cost)
			cost[i] = c;
	}
}

// rectangular footprint variant (ppx_device.cuh, RectDev): one edge / pose per thread, exact FP64 positions
__global__ void __launch_bounds__(kThreads) k_check_rs_rect(MapDev map, RectDev R, int64_t n, const ppx_rs_edge* __restrict__ edges,
	double r, uint8_t* __restrict__ valid, float* __restrict__ last, uint32_t* __restrict__ samples)
{
	int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
	const int64_t stride = (int64_t)gridDim.x * blockDim.x;
	for (; i < n; i += stride) {
		Pose start;
		RSSeg seg;
		load_rs_edge(edges + i, start, seg);
		float l;
		unsigned ns;
		bool ok = rs_is_path_valid_rect(map, R, start, seg, r, &l, &ns);
		valid[i] = ok ? 1 : 0;
		last[i] = l;
		if (samples)
			samples[i] = ns;
	}
}

__global__ void __launch_bounds__(kThreads) k_is_state_valid_rect(MapDev map, RectDev R, int64_t n, const double* __restrict__ poses,
	uint8_t* __restrict__ valid)
{
	int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
	const int64_t stride = (int64_t)gridDim.x * blockDim.x;
	for (; i < n; i += stride) {
		float slack, border;
		valid[i] = rect_state_valid(map, R, poses[3 * i], poses[3 * i + 1], poses[3 * i + 2], &slack, &border) ? 1 : 0;
	}
}

// ppx_footprint_params -> device constants; false when the parameters make no sense
bool rect_derive(const ppx_footprint_params* fp, double rmin, RectDev& R)
{
	if (!fp || !(fp->length > 0.0) || !(fp->width > 0.0) || fp->n_discs < 1 || fp->n_discs > 16 || !(fp->min_step > 0.0f) || !(rmin > 0.0))
		return false;
	R.n = fp->n_discs;
	const double seg = fp->length / fp->n_discs; // every disc covers a length / n by width slice of the rectangle
	R.pitch = seg;
	R.first = fp->rear_to_center - fp->length / 2.0 + seg / 2.0;
	R.rho = (float)sqrt(seg * seg / 4.0 + fp->width * fp->width / 4.0);
	double omax = 0.0;
	for (int i = 0; i < R.n; i++)
		omax = fmax(omax, fabs(R.first + i * R.pitch));
	R.gamma = (float)sqrt(1.0 + (omax / rmin) * (omax / rmin)); // a disc centre at offset o moves ds * sqrt(1 + (o * kappa)^2), kappa <= 1 / rmin
	R.minStep = fp->min_step;
	return true;
}

int grid_for(const ppx_ctx* ctx, int64_t n, int threads, int blocksPerSm)
{
	int64_t need = (n + threads - 1) / threads;
	int64_t cap = (int64_t)ctx->num_sms * blocksPerSm;
	return (int)std::max<int64_t>(1, std::min(need, cap));
}

int require_map(ppx_ctx* ctx)
{
	if (!ctx)
		return PPX_ERR_INVALID;
	if (!ctx->has_map)
		return ppx_fail(ctx, PPX_ERR_STATE, "no map uploaded (ppx_map_upload)");
	return PPX_OK;
}

}

extern "C" {

int ppx_is_state_valid(ppx_ctx* ctx, int64_t n, const double* poses, const ppx_validator_params* vp, uint8_t* valid)
{
	PPX_GUARD(ctx);
	PPX_CHECK(require_map(ctx));
	PPX_REQUIRE(ctx, n >= 0 && vp && (n == 0 || (poses && valid)), "bad arguments");
	if (n == 0)
		return PPX_OK;
	double* d_poses;
	uint8_t* d_valid;
	PPX_CHECK(ppx_scratch(ctx, 0, n * 3 * sizeof(double), (void**)&d_poses));
	PPX_CHECK(ppx_scratch(ctx, 1, n, (void**)&d_valid));
	PPX_CUDA(ctx, cudaMemcpyAsync(d_poses, poses, n * 3 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
	k_is_state_valid<<<grid_for(ctx, n, kThreads, 16), kThreads, 0, ctx->stream>>>(ctx->map_dev(), n, d_poses,
		vp->min_safe_radius, d_valid);
	PPX_LAUNCHED(ctx);
	PPX_CUDA(ctx, cudaMemcpyAsync(valid, d_valid, n, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

static int launch_check_rs(ppx_ctx* ctx, int64_t n, const ppx_rs_edge* d_edges, double min_turning_radius,
	const ppx_validator_params* vp, uint8_t* d_valid, float* d_last, uint32_t* d_samples, int queueSlot, bool overlap = false)
{
	FastMap fm = fast_map_derive(ctx->desc);
	if (!ctx->rs_filter)
		fm.enabled = 0;
	if (overlap) {
		// programmatic dependent launch: this kernel may start while the previous one on the stream still drains
		cudaLaunchConfig_t cfg = {};
		// Two CTAs per SM and launch (the SM holds six): up to three launches are resident at once, and the warps of a launch
		// are a third as many -- each works through three times the edges before it reaches its tail of long sample chains,
		// which is where lanes idle (1M-edge launches, CTAs per SM and launch 6 / 4 / 2 / 1: 7.98 / 8.65 / 9.36-9.59 / 9.92 G
		// edges/s; 2 rather than 1 so that the last launch of a stream drains at a third of the GPU, not a sixth)
		static const int overlapCtas = getenv("PPX_OVERLAP_CTAS") ? std::max(1, std::min(6, atoi(getenv("PPX_OVERLAP_CTAS")))) : kOverlapCtasPerSm;
		cfg.gridDim = dim3((unsigned)grid_for(ctx, n, kThreads, overlapCtas));
		cfg.blockDim = dim3(kThreads);
		cfg.dynamicSmemBytes = 0;
		cfg.stream = ctx->stream;
		cudaLaunchAttribute attr[1];
		attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
		attr[0].val.programmaticStreamSerializationAllowed = 1;
		cfg.attrs = attr;
		cfg.numAttrs = 1;
		unsigned long long* q = ctx->d_rsq + queueSlot;
		unsigned long long* next = ctx->d_rsq + (queueSlot + 1) % ppx_ctx::kRsqSlots;
		if (ctx->rs_engine >= 1)
			PPX_CUDA(ctx, cudaLaunchKernelEx(&cfg, k_check_rs2<false, false>, ctx->map_dev(), fm, n, d_edges, min_turning_radius,
				vp->min_safe_radius, vp->min_path_interpolation_distance, d_valid, d_last, d_samples, q, (AuditOut*)nullptr, next, SolveArgs {}));
		else
			PPX_CUDA(ctx, cudaLaunchKernelEx(&cfg, k_check_rs<false, kCheckCtasPerSm>, ctx->map_dev(), fm, n, d_edges, min_turning_radius,
				vp->min_safe_radius, vp->min_path_interpolation_distance, d_valid, d_last, d_samples, q, (AuditOut*)nullptr, next));
		ctx->launches++;
		return PPX_OK;
	}
	if (ctx->rs_engine >= 1)
		k_check_rs2<false, false><<<grid_for(ctx, n, kThreads, kCheck2Ctas), kThreads, 0, ctx->stream>>>(ctx->map_dev(), fm, n, d_edges,
			min_turning_radius, vp->min_safe_radius, vp->min_path_interpolation_distance, d_valid, d_last, d_samples,
			ctx->d_queue + queueSlot, nullptr, nullptr, SolveArgs {});
	else
		k_check_rs<false, kCheckCtasPerSm><<<grid_for(ctx, n, kThreads, kCheckCtasPerSm), kThreads, 0, ctx->stream>>>(ctx->map_dev(), fm, n, d_edges,
			min_turning_radius, vp->min_safe_radius, vp->min_path_interpolation_distance, d_valid, d_last, d_samples,
			ctx->d_queue + queueSlot, nullptr, nullptr);
	PPX_LAUNCHED(ctx);
	return PPX_OK;
}

int ppx_check_edges_rs_dev(ppx_ctx* ctx, int64_t n, const ppx_rs_edge* d_edges, double min_turning_radius,
	const ppx_validator_params* vp, uint8_t* d_valid, float* d_last, uint32_t* d_samples)
{
	PPX_GUARD(ctx);
	PPX_CHECK(require_map(ctx));
	PPX_REQUIRE(ctx, n >= 0 && vp && (n == 0 || (d_edges && d_valid && d_last)), "bad arguments");
	if (n == 0)
		return PPX_OK;
	if (ctx->rs_overlap) { // no memset between the kernels: each launch zeroes the cursor of the next one
		const int slot = ctx->rs_slot;
		ctx->rs_slot = (slot + 1) % ppx_ctx::kRsqSlots;
		return launch_check_rs(ctx, n, d_edges, min_turning_radius, vp, d_valid, d_last, d_samples, slot, true);
	}
	PPX_CUDA(ctx, cudaMemsetAsync(ctx->d_queue, 0, sizeof(unsigned long long), ctx->stream));
	return launch_check_rs(ctx, n, d_edges, min_turning_radius, vp, d_valid, d_last, d_samples, 0);
}

int ppx_check_edges_rs_audit(ppx_ctx* ctx, int64_t n, const ppx_rs_edge* edges, double min_turning_radius,
	const ppx_validator_params* vp, uint8_t* valid, float* last, uint32_t* samples, uint64_t* counters, float* max_dev_cells)
{
	PPX_GUARD(ctx);
	PPX_CHECK(require_map(ctx));
	PPX_REQUIRE(ctx, n > 0 && vp && edges && valid && last && samples && counters && max_dev_cells, "bad arguments");
	ppx_rs_edge* d_edges;
	uint8_t* d_valid;
	float* d_last;
	uint32_t* d_samples;
	AuditOut* d_audit;
	PPX_CHECK(ppx_scratch(ctx, 0, n * sizeof(ppx_rs_edge), (void**)&d_edges));
	PPX_CHECK(ppx_scratch(ctx, 1, n, (void**)&d_valid));
	PPX_CHECK(ppx_scratch(ctx, 2, n * sizeof(float), (void**)&d_last));
	PPX_CHECK(ppx_scratch(ctx, 3, n * sizeof(uint32_t), (void**)&d_samples));
	PPX_CHECK(ppx_scratch(ctx, 4, sizeof(AuditOut), (void**)&d_audit));
	PPX_CUDA(ctx, cudaMemcpyAsync(d_edges, edges, n * sizeof(ppx_rs_edge), cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemsetAsync(d_audit, 0, sizeof(AuditOut), ctx->stream));
	PPX_CUDA(ctx, cudaMemsetAsync(ctx->d_queue, 0, sizeof(unsigned long long), ctx->stream));
	FastMap fm = fast_map_derive(ctx->desc);
	if (ctx->rs_engine >= 1)
		k_check_rs2<true, false><<<grid_for(ctx, n, kThreads, kCheck2Ctas), kThreads, 0, ctx->stream>>>(ctx->map_dev(), fm, n, d_edges,
			min_turning_radius, vp->min_safe_radius, vp->min_path_interpolation_distance, d_valid, d_last, d_samples, ctx->d_queue,
			d_audit, nullptr, SolveArgs {});
	else
		k_check_rs<true, kCheckCtasPerSm><<<grid_for(ctx, n, kThreads, kCheckCtasPerSm), kThreads, 0, ctx->stream>>>(ctx->map_dev(), fm, n, d_edges,
			min_turning_radius, vp->min_safe_radius, vp->min_path_interpolation_distance, d_valid, d_last, d_samples, ctx->d_queue,
			d_audit, nullptr);
	PPX_LAUNCHED(ctx);
	AuditOut h {};
	PPX_CUDA(ctx, cudaMemcpyAsync(valid, d_valid, n, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(last, d_last, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(samples, d_samples, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(&h, d_audit, sizeof(AuditOut), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	counters[0] = h.certified;
	counters[1] = h.parked;
	counters[2] = h.cell_mismatch;
	memcpy(max_dev_cells, &h.max_dev_bits, 4);
	return PPX_OK;
}

// Host-buffer entry: the batch is cut into chunks; chunk c's H2D copy (stream s_h2d), kernel (ctx->stream)
// and D2H copies (stream s_d2h) are chained by events, so that the copy engines and the SMs work on
// different chunks at the same time.  With pinned host buffers the call runs at PCIe H2D speed.
int ppx_check_edges_rs(ppx_ctx* ctx, int64_t n, const ppx_rs_edge* edges, double min_turning_radius,
	const ppx_validator_params* vp, uint8_t* valid, float* last, uint32_t* samples)
{
	PPX_GUARD(ctx);
	PPX_CHECK(require_map(ctx));
	PPX_REQUIRE(ctx, n >= 0 && vp && (n == 0 || (edges && valid && last)), "bad arguments");
	if (n == 0)
		return PPX_OK;
	ppx_rs_edge* d_edges;
	uint8_t* d_valid;
	float* d_last;
	uint32_t* d_samples = nullptr;
	PPX_CHECK(ppx_scratch(ctx, 0, n * sizeof(ppx_rs_edge), (void**)&d_edges));
	PPX_CHECK(ppx_scratch(ctx, 1, n, (void**)&d_valid));
	PPX_CHECK(ppx_scratch(ctx, 2, n * sizeof(float), (void**)&d_last));
	if (samples)
		PPX_CHECK(ppx_scratch(ctx, 3, n * sizeof(uint32_t), (void**)&d_samples));
	PPX_CHECK(ppx_pipeline_init(ctx));
	const int64_t minChunk = 250000; // a launch has a fixed tail of ~0.1 ms (its longest sample chains): few, large chunks
	int nchunks = (int)std::min<int64_t>(ppx_ctx::kMaxChunks, std::max<int64_t>(1, n / minChunk));
	int64_t chunk = ((n + nchunks - 1) / nchunks + 1) & ~(int64_t)1; // even: every chunk starts on a 16-byte boundary (bulk copy)
	PPX_CUDA(ctx, cudaMemsetAsync(ctx->d_queue, 0, ppx_ctx::kMaxChunks * sizeof(unsigned long long), ctx->stream));
	PPX_CUDA(ctx, cudaEventRecord(ctx->ev_start, ctx->stream));
	PPX_CUDA(ctx, cudaStreamWaitEvent(ctx->s_h2d, ctx->ev_start, 0));
	PPX_CUDA(ctx, cudaStreamWaitEvent(ctx->s_d2h, ctx->ev_start, 0));
	for (int c = 0; c < nchunks; c++) {
		int64_t off = c * chunk, m = std::min(chunk, n - off);
		if (m <= 0)
			break;
		PPX_CUDA(ctx, cudaMemcpyAsync(d_edges + off, edges + off, m * sizeof(ppx_rs_edge), cudaMemcpyHostToDevice, ctx->s_h2d));
		PPX_CUDA(ctx, cudaEventRecord(ctx->ev_h2d[c], ctx->s_h2d));
		PPX_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_h2d[c], 0));
		PPX_CHECK(launch_check_rs(ctx, m, d_edges + off, min_turning_radius, vp, d_valid + off, d_last + off,
			d_samples ? d_samples + off : nullptr, c));
		PPX_CUDA(ctx, cudaEventRecord(ctx->ev_k[c], ctx->stream));
		PPX_CUDA(ctx, cudaStreamWaitEvent(ctx->s_d2h, ctx->ev_k[c], 0));
		PPX_CUDA(ctx, cudaMemcpyAsync(valid + off, d_valid + off, m, cudaMemcpyDeviceToHost, ctx->s_d2h));
		PPX_CUDA(ctx, cudaMemcpyAsync(last + off, d_last + off, m * sizeof(float), cudaMemcpyDeviceToHost, ctx->s_d2h));
		if (samples)
			PPX_CUDA(ctx, cudaMemcpyAsync(samples + off, d_samples + off, m * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->s_d2h));
	}
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->s_d2h));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

int ppx_check_edges_rs_rect_dev(ppx_ctx* ctx, int64_t n, const ppx_rs_edge* d_edges, double min_turning_radius,
	const ppx_footprint_params* fp, uint8_t* d_valid, float* d_last, uint32_t* d_samples)
{
	PPX_GUARD(ctx);
	PPX_CHECK(require_map(ctx));
	PPX_REQUIRE(ctx, n >= 0 && (n == 0 || (d_edges && d_valid && d_last)), "bad arguments");
	RectDev R;
	PPX_REQUIRE(ctx, rect_derive(fp, min_turning_radius, R), "bad footprint parameters");
	if (n == 0)
		return PPX_OK;
	k_check_rs_rect<<<grid_for(ctx, n, kThreads, 16), kThreads, 0, ctx->stream>>>(ctx->map_dev(), R, n, d_edges, min_turning_radius, d_valid,
		d_last, d_samples);
	PPX_LAUNCHED(ctx);
	return PPX_OK;
}

int ppx_check_edges_rs_rect(ppx_ctx* ctx, int64_t n, const ppx_rs_edge* edges, double min_turning_radius,
	const ppx_footprint_params* fp, uint8_t* valid, float* last, uint32_t* samples)
{
	PPX_GUARD(ctx);
	PPX_CHECK(require_map(ctx));
	PPX_REQUIRE(ctx, n >= 0 && (n == 0 || (edges && valid && last)), "bad arguments");
	if (n == 0)
		return PPX_OK;
	ppx_rs_edge* d_edges;
	uint8_t* d_valid;
	float* d_last;
	uint32_t* d_samples = nullptr;
	PPX_CHECK(ppx_scratch(ctx, 0, n * sizeof(ppx_rs_edge), (void**)&d_edges));
	PPX_CHECK(ppx_scratch(ctx, 1, n, (void**)&d_valid));
	PPX_CHECK(ppx_scratch(ctx, 2, n * sizeof(float), (void**)&d_last));
	if (samples)
		PPX_CHECK(ppx_scratch(ctx, 3, n * sizeof(uint32_t), (void**)&d_samples));
	PPX_CUDA(ctx, cudaMemcpyAsync(d_edges, edges, n * sizeof(ppx_rs_edge), cudaMemcpyHostToDevice, ctx->stream));
	PPX_CHECK(ppx_check_edges_rs_rect_dev(ctx, n, d_edges, min_turning_radius, fp, d_valid, d_last, d_samples));
	PPX_CUDA(ctx, cudaMemcpyAsync(valid, d_valid, n, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(last, d_last, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
	if (samples)
		PPX_CUDA(ctx, cudaMemcpyAsync(samples, d_samples, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

int ppx_is_state_valid_rect(ppx_ctx* ctx, int64_t n, const double* poses, const ppx_footprint_params* fp, uint8_t* valid)
{
	PPX_GUARD(ctx);
	PPX_CHECK(require_map(ctx));
	PPX_REQUIRE(ctx, n >= 0 && (n == 0 || (poses && valid)), "bad arguments");
	RectDev R;
	PPX_REQUIRE(ctx, rect_derive(fp, 1.0, R), "bad footprint parameters");
	if (n == 0)
		return PPX_OK;
	double* d_poses;
	uint8_t* d_valid;
	PPX_CHECK(ppx_scratch(ctx, 0, n * 3 * sizeof(double), (void**)&d_poses));
	PPX_CHECK(ppx_scratch(ctx, 1, n, (void**)&d_valid));
	PPX_CUDA(ctx, cudaMemcpyAsync(d_poses, poses, n * 3 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
	k_is_state_valid_rect<<<grid_for(ctx, n, kThreads, 16), kThreads, 0, ctx->stream>>>(ctx->map_dev(), R, n, d_poses, d_valid);
	PPX_LAUNCHED(ctx);
	PPX_CUDA(ctx, cudaMemcpyAsync(valid, d_valid, n, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

int ppx_check_edges_csteer(ppx_ctx* ctx, int64_t n, const ppx_csteer_edge* edges, const ppx_validator_params* vp,
	uint8_t* valid, float* last, double* final_pose)
{
	PPX_GUARD(ctx);
	PPX_CHECK(require_map(ctx));
	PPX_REQUIRE(ctx, n >= 0 && vp && (n == 0 || (edges && valid && last)), "bad arguments");
	if (n == 0)
		return PPX_OK;
	ppx_csteer_edge* d_edges;
	uint8_t* d_valid;
	float* d_last;
	double* d_final = nullptr;
	PPX_CHECK(ppx_scratch(ctx, 0, n * sizeof(ppx_csteer_edge), (void**)&d_edges));
	PPX_CHECK(ppx_scratch(ctx, 1, n, (void**)&d_valid));
	PPX_CHECK(ppx_scratch(ctx, 2, n * sizeof(float), (void**)&d_last));
	if (final_pose)
		PPX_CHECK(ppx_scratch(ctx, 3, n * 3 * sizeof(double), (void**)&d_final));
	PPX_CUDA(ctx, cudaMemcpyAsync(d_edges, edges, n * sizeof(ppx_csteer_edge), cudaMemcpyHostToDevice, ctx->stream));
	k_check_csteer<<<grid_for(ctx, n, kThreads, 16), kThreads, 0, ctx->stream>>>(ctx->map_dev(), n, d_edges,
		vp->min_safe_radius, vp->min_path_interpolation_distance, d_valid, d_last, d_final);
	PPX_LAUNCHED(ctx);
	PPX_CUDA(ctx, cudaMemcpyAsync(valid, d_valid, n, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(last, d_last, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
	if (final_pose)
		PPX_CUDA(ctx, cudaMemcpyAsync(final_pose, d_final, n * 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

int ppx_check_edges_line(ppx_ctx* ctx, int64_t n, const ppx_line_edge* edges, const ppx_validator_params* vp,
	uint8_t* valid, float* last)
{
	PPX_GUARD(ctx);
	PPX_CHECK(require_map(ctx));
	PPX_REQUIRE(ctx, n >= 0 && vp && (n == 0 || (edges && valid && last)), "bad arguments");
	if (n == 0)
		return PPX_OK;
	ppx_line_edge* d_edges;
	uint8_t* d_valid;
	float* d_last;
	PPX_CHECK(ppx_scratch(ctx, 0, n * sizeof(ppx_line_edge), (void**)&d_edges));
	PPX_CHECK(ppx_scratch(ctx, 1, n, (void**)&d_valid));
	PPX_CHECK(ppx_scratch(ctx, 2, n * sizeof(float), (void**)&d_last));
	PPX_CUDA(ctx, cudaMemcpyAsync(d_edges, edges, n * sizeof(ppx_line_edge), cudaMemcpyHostToDevice, ctx->stream));
	k_check_line<<<grid_for(ctx, n, kThreads, 16), kThreads, 0, ctx->stream>>>(ctx->map_dev(), n, d_edges,
		vp->min_safe_radius, vp->min_path_interpolation_distance, d_valid, d_last);
	PPX_LAUNCHED(ctx);
	PPX_CUDA(ctx, cudaMemcpyAsync(valid, d_valid, n, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(last, d_last, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

int ppx_rs_solve_dev(ppx_ctx* ctx, int64_t n, const double* d_starts, const double* d_goals,
	const ppx_rs_cost_params* cp, ppx_rs_edge* d_edges, int32_t* d_word, double* d_length, float* d_cost)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	PPX_REQUIRE(ctx, n >= 0 && cp && (n == 0 || (d_starts && d_goals)), "bad arguments");
	if (n == 0)
		return PPX_OK;
	if (ctx->rs_engine == 2) {
		// quad form (8 pairs per warp, no formula divergence): 1.96 ms per 1M pairs on B200 against 1.62 ms for one pair per
		// thread -- the per-thread loop over the 48 words does not diverge either and needs no cross-lane reduction; kept for
		// A/B (tools/time_check.py), the fused Connect + IsPathValid kernel uses the quad solver because it runs at 168 registers
		k_rs_solve_quad<<<grid_for(ctx, 4 * n, kThreads, 8), kThreads, 0, ctx->stream>>>(n, d_starts, d_goals, *cp, d_edges, d_word, d_length,
			d_cost);
	} else {
		MapDev none {};
		// 4 CTAs per SM (128 registers): 1.25 ms per 1M pairs against 1.61 ms at 2 (236 registers); ncu at 2: FP64 pipe 13 %,
		// issue slots 35 % busy on 8 warps per SM (profiles/r2t_k_rs_solve_*) -- latency-bound, not FP64-bound
		k_rs_solve<4, false><<<grid_for(ctx, n, kThreads, 16), kThreads, 0, ctx->stream>>>(none, n, d_starts, d_goals, *cp, 0, 0.0f, 0.0f,
			d_edges, d_word, d_length, d_cost, nullptr, nullptr);
	}
	PPX_LAUNCHED(ctx);
	return PPX_OK;
}

int ppx_connect_check_rs_dev(ppx_ctx* ctx, int64_t n, const double* d_starts, const double* d_goals,
	const ppx_rs_cost_params* cp, const ppx_validator_params* vp, uint8_t* d_valid, float* d_last, int32_t* d_word,
	double* d_length, float* d_cost)
{
	PPX_GUARD(ctx);
	PPX_CHECK(require_map(ctx));
	PPX_REQUIRE(ctx, n >= 0 && cp && vp && (n == 0 || (d_starts && d_goals && d_valid && d_last)), "bad arguments");
	if (n == 0)
		return PPX_OK;
	if (ctx->rs_engine >= 1) {
		// Connect + IsPathValid in one kernel: quad solver feeding the lane machine (ppx_check_rs2.cuh); the solved segments
		// pass through a scratch array the exact path re-reads
		ppx_rs_edge* d_solved;
		PPX_CHECK(ppx_scratch(ctx, 7, (size_t)n * sizeof(ppx_rs_edge), (void**)&d_solved));
		FastMap fm = fast_map_derive(ctx->desc);
		if (!ctx->rs_filter)
			fm.enabled = 0;
		SolveArgs sa { d_starts, d_goals, d_solved, d_word, d_length, d_cost, cp->reverse_cost_multiplier, cp->forward_cost_multiplier,
			cp->direction_switching_cost, cp->shortest };
		PPX_CUDA(ctx, cudaMemsetAsync(ctx->d_queue, 0, sizeof(unsigned long long), ctx->stream));
		// 5 CTAs per SM (96 registers, some spills): 2.18 ms per 1M pairs against 2.68 ms at 3 (168 registers) -- the solver
		// is a chain of dependent FP64 latencies, more warps hide it (sweep 3 / 4 / 5 / 6: 2.68 / 2.32 / 2.18 / 2.32 ms)
		k_check_rs2<false, true, 5><<<grid_for(ctx, n, kThreads, 5), kThreads, 0, ctx->stream>>>(ctx->map_dev(), fm, n, nullptr,
			cp->min_turning_radius, vp->min_safe_radius, vp->min_path_interpolation_distance, d_valid, d_last, nullptr, ctx->d_queue,
			nullptr, nullptr, sa);
	} else {
		k_rs_solve<2, true><<<grid_for(ctx, n, kThreads, 16), kThreads, 0, ctx->stream>>>(ctx->map_dev(), n, d_starts, d_goals, *cp, 1,
			vp->min_safe_radius, vp->min_path_interpolation_distance, nullptr, d_word, d_length, d_cost, d_valid, d_last);
	}
	PPX_LAUNCHED(ctx);
	return PPX_OK;
}

static int solve_host(ppx_ctx* ctx, int64_t n, const double* starts, const double* goals, const ppx_rs_cost_params* cp,
	const ppx_validator_params* vp, ppx_rs_edge* edges, int32_t* word, double* length, float* cost, uint8_t* valid,
	float* last)
{
	double *d_starts, *d_goals, *d_length = nullptr;
	ppx_rs_edge* d_edges = nullptr;
	int32_t* d_word = nullptr;
	float *d_cost = nullptr, *d_last = nullptr;
	uint8_t* d_valid = nullptr;
	// one scratch slot, carved
	size_t off = 0;
	auto carve = [&](size_t bytes) {
		size_t o = off;
		off += (bytes + 255) & ~(size_t)255;
		return o;
	};
	size_t o_starts = carve(n * 3 * sizeof(double)), o_goals = carve(n * 3 * sizeof(double));
	size_t o_edges = carve(edges ? n * sizeof(ppx_rs_edge) : 0), o_word = carve(word ? n * 4 : 0);
	size_t o_length = carve(length ? n * 8 : 0), o_cost = carve(cost ? n * 4 : 0);
	size_t o_valid = carve(vp ? n : 0), o_last = carve(vp ? n * 4 : 0);
	char* base;
	PPX_CHECK(ppx_scratch(ctx, 0, off, (void**)&base));
	d_starts = (double*)(base + o_starts);
	d_goals = (double*)(base + o_goals);
	if (edges) d_edges = (ppx_rs_edge*)(base + o_edges);
	if (word) d_word = (int32_t*)(base + o_word);
	if (length) d_length = (double*)(base + o_length);
	if (cost) d_cost = (float*)(base + o_cost);
	if (vp) {
		d_valid = (uint8_t*)(base + o_valid);
		d_last = (float*)(base + o_last);
	}
	PPX_CUDA(ctx, cudaMemcpyAsync(d_starts, starts, n * 3 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(d_goals, goals, n * 3 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
	if (vp)
		PPX_CHECK(ppx_connect_check_rs_dev(ctx, n, d_starts, d_goals, cp, vp, d_valid, d_last, d_word, d_length, d_cost));
	else
		PPX_CHECK(ppx_rs_solve_dev(ctx, n, d_starts, d_goals, cp, d_edges, d_word, d_length, d_cost));
	if (edges) PPX_CUDA(ctx, cudaMemcpyAsync(edges, d_edges, n * sizeof(ppx_rs_edge), cudaMemcpyDeviceToHost, ctx->stream));
	if (word) PPX_CUDA(ctx, cudaMemcpyAsync(word, d_word, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (length) PPX_CUDA(ctx, cudaMemcpyAsync(length, d_length, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
	if (cost) PPX_CUDA(ctx, cudaMemcpyAsync(cost, d_cost, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (vp) {
		PPX_CUDA(ctx, cudaMemcpyAsync(valid, d_valid, n, cudaMemcpyDeviceToHost, ctx->stream));
		PPX_CUDA(ctx, cudaMemcpyAsync(last, d_last, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
	}
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

int ppx_rs_solve(ppx_ctx* ctx, int64_t n, const double* starts, const double* goals, const ppx_rs_cost_params* cp,
	ppx_rs_edge* edges, int32_t* word, double* length, float* cost)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	PPX_REQUIRE(ctx, n >= 0 && cp && (n == 0 || (starts && goals)), "bad arguments");
	if (n == 0)
		return PPX_OK;
	return solve_host(ctx, n, starts, goals, cp, nullptr, edges, word, length, cost, nullptr, nullptr);
}

int ppx_connect_check_rs(ppx_ctx* ctx, int64_t n, const double* starts, const double* goals,
	const ppx_rs_cost_params* cp, const ppx_validator_params* vp, uint8_t* valid, float* last, int32_t* word,
	double* length, float* cost)
{
	PPX_GUARD(ctx);
	PPX_CHECK(require_map(ctx));
	PPX_REQUIRE(ctx, n >= 0 && cp && vp && (n == 0 || (starts && goals && valid && last)), "bad arguments");
	if (n == 0)
		return PPX_OK;
	return solve_host(ctx, n, starts, goals, cp, vp, nullptr, word, length, cost, valid, last);
}

int ppx_rs_interpolate(ppx_ctx* ctx, int64_t n, const ppx_rs_edge* edges, double min_turning_radius, int32_t nratios,
	const double* ratios, double* poses)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	PPX_REQUIRE(ctx, n >= 0 && nratios >= 0 && (n * nratios == 0 || (edges && ratios && poses)), "bad arguments");
	if (n * nratios == 0)
		return PPX_OK;
	ppx_rs_edge* d_edges;
	double *d_ratios, *d_poses;
	int64_t total = n * nratios;
	PPX_CHECK(ppx_scratch(ctx, 0, n * sizeof(ppx_rs_edge), (void**)&d_edges));
	PPX_CHECK(ppx_scratch(ctx, 1, total * sizeof(double), (void**)&d_ratios));
	PPX_CHECK(ppx_scratch(ctx, 2, total * 3 * sizeof(double), (void**)&d_poses));
	PPX_CUDA(ctx, cudaMemcpyAsync(d_edges, edges, n * sizeof(ppx_rs_edge), cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(d_ratios, ratios, total * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
	k_rs_interpolate<<<grid_for(ctx, total, kThreads, 16), kThreads, 0, ctx->stream>>>(n, d_edges, min_turning_radius, nratios,
		d_ratios, d_poses);
	PPX_LAUNCHED(ctx);
	PPX_CUDA(ctx, cudaMemcpyAsync(poses, d_poses, total * 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

} // extern "C"
