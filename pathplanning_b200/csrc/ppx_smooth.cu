// libppx.so: batched path smoother, Smoother::Smooth of algo/smoother.cpp:33-158.
//
// One CTA per path.  The path's points live in shared memory (double-buffered positions + the three
// curvature parts of every optimised index); an iteration of the reference's gradient descent is
//   phase A  every optimised index computes its curvature parts           (CalculateCurvatureTerm)
//   phase B  every point assembles its gradient in the reference's summation order, steps, and
//            contributes |gradient| to the block-wide maximum that ends the descent
// with two block barriers per iteration.  See ppx_smooth.cuh for the order argument.
#include <algorithm>
#include "ppx_ctx.h"
#include "ppx_smooth.cuh"

using namespace ppx;

namespace {

constexpr int kSmoothThreads = 256;
constexpr int kBytesPerPoint = 5 * (int)sizeof(V2) + 2; // pos, pos', 3 curvature parts, cusp flag, state flags

struct SmoothConsts {
	SmoothDev P;
	MapDev map;
};

__global__ void __launch_bounds__(kSmoothThreads) k_smooth(const SmoothConsts* __restrict__ consts, int64_t nq,
	const double* __restrict__ poses, const uint8_t* __restrict__ cusp, const int32_t* __restrict__ count, int maxPoints, int capPoints,
	double* __restrict__ out, int32_t* __restrict__ status, int32_t* __restrict__ iterations)
{
	extern __shared__ double s_raw[];
	__shared__ float s_red[2][kSmoothThreads / 32];
	__shared__ int s_unsafe;
	V2* pos = (V2*)s_raw;
	V2* nxt = pos + capPoints; // capPoints: the longest path of the batch
	V2* cm1 = nxt + capPoints;
	V2* cc = cm1 + capPoints;
	V2* cp1 = cc + capPoints;
	uint8_t* s_cusp = (uint8_t*)(cp1 + capPoints);
	uint8_t* flags = s_cusp + capPoints; // bit 0: optimised index, bit 1: curvature term active this iteration
	const SmoothDev& P = consts->P;
	const MapDev& map = consts->map;
	const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	const float unsafeRadius = P.minSafeRadius * (1 + P.collisionRatio);

	for (int64_t q = blockIdx.x; q < nq; q += gridDim.x) {
		const int n = count[q];
		const double* in = poses + (size_t)q * maxPoints * 3;
		double* o = out + (size_t)q * maxPoints * 3;
		__syncthreads();
		pos = (V2*)s_raw;
		nxt = pos + capPoints;
		if (n <= 0) { // no path to smooth
			if (tid == 0) {
				status[q] = -1;
				if (iterations)
					iterations[q] = 0;
			}
			continue;
		}
		if (tid == 0)
			s_unsafe = 0;
		for (int i = tid; i < n; i += kSmoothThreads) {
			pos[i] = v2(in[3 * i], in[3 * i + 1]);
			s_cusp[i] = cusp[(size_t)q * maxPoints + i];
		}
		__syncthreads();
		for (int i = tid; i < n; i += kSmoothThreads)
			flags[i] = smooth_is_index(s_cusp, n, i) ? 1 : 0;
		__syncthreads();

		int st;
		int it = -1;
		if (n < 5) {
			st = 2; // Status::PathSize
		} else {
			float step = P.stepTolerance;
			while (true) {
				it++;
				if (it >= P.maxIterations) {
					st = 0; // Status::MaxIteration
					break;
				}
				if (step < P.stepTolerance) {
					st = 1; // Status::StepTolerance
					break;
				}
				for (int i = tid; i < n; i += kSmoothThreads) {
					if (!(flags[i] & 1))
						continue;
					V2 a, b, c;
					bool on = smooth_curvature(P, pos[i - 1], pos[i], pos[i + 1], a, b, c);
					flags[i] = on ? 3 : 1;
					if (on) {
						cm1[i] = a;
						cc[i] = b;
						cp1[i] = c;
					}
				}
				__syncthreads();
				float mx = 0.0f;
				for (int i = tid; i < n; i += kSmoothThreads) {
					V2 g = v2(0.0, 0.0);
					if (i > 0 && flags[i - 1] == 3)
						g = g + cp1[i - 1];
					if (flags[i] & 1) {
						g = smooth_own_terms(P, map, g, v2(in[3 * i], in[3 * i + 1]), pos[i - 2], pos[i - 1], pos[i], pos[i + 1], pos[i + 2],
							unsafeRadius);
						if (flags[i] == 3)
							g = g + cc[i];
					}
					if (i + 1 < n && flags[i + 1] == 3)
						g = g + cm1[i + 1];
					nxt[i] = pos[i] - (double)P.learningRate * g;
					mx = fmaxf(mx, (float)v2_norm(g));
				}
#pragma unroll
				for (int d = 16; d > 0; d >>= 1)
					mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, d));
				if (lane == 0)
					s_red[it & 1][warp] = mx;
				__syncthreads();
				step = s_red[it & 1][0];
#pragma unroll
				for (int w = 1; w < kSmoothThreads / 32; w++)
					step = fmaxf(step, s_red[it & 1][w]);
				V2* t = pos;
				pos = nxt;
				nxt = t;
			}
		}
		// IsPathSafe (smoother.cpp:215-224) on the smoothed positions with the input headings
		int bad = 0;
		for (int i = tid; i < n; i += kSmoothThreads) {
			double th = in[3 * i + 2];
			float d;
			if (!is_state_valid(map, pos[i].x, pos[i].y, th, P.minSafeRadius, &d))
				bad = 1;
			o[3 * i] = pos[i].x;
			o[3 * i + 1] = pos[i].y;
			o[3 * i + 2] = th;
		}
		if (bad)
			s_unsafe = 1;
		__syncthreads();
		if (tid == 0) {
			status[q] = s_unsafe ? -1 : st; // Status::Failure
			if (iterations)
				iterations[q] = it;
		}
	}
}

}

struct ppx_smoother {
	ppx_ctx* ctx = nullptr;
	ppx_smoother_params params {};
	int* d_near_obstacle = nullptr;
	int* d_near_edge = nullptr;
	SmoothConsts* d_consts = nullptr;
	bool has_maps = false;
};

extern "C" int ppx_smoother_create(ppx_ctx* ctx, const ppx_smoother_params* params, ppx_smoother** out)
{
	PPX_GUARD(ctx);
	if (!ctx || !out)
		return PPX_ERR_INVALID;
	PPX_REQUIRE(ctx, params, "ppx_smoother_create needs parameters");
	PPX_REQUIRE(ctx, ctx->has_map, "upload or build the map before creating a smoother");
	PPX_REQUIRE(ctx, params->max_iterations >= 0 && params->learning_rate == params->learning_rate, "bad smoother parameters");
	auto* s = new ppx_smoother();
	s->ctx = ctx;
	s->params = *params;
	*out = s;
	return PPX_OK;
}

extern "C" int ppx_smoother_destroy(ppx_smoother* s)
{
	PPX_GUARD(s ? s->ctx : nullptr);
	if (!s)
		return PPX_ERR_INVALID;
	cudaFree(s->d_near_obstacle);
	cudaFree(s->d_near_edge);
	cudaFree(s->d_consts);
	delete s;
	return PPX_OK;
}

extern "C" int ppx_smoother_set_gvd(ppx_smoother* s, const int32_t* nearest_obstacle, const int32_t* nearest_edge)
{
	PPX_GUARD(s ? s->ctx : nullptr);
	if (!s)
		return PPX_ERR_INVALID;
	ppx_ctx* ctx = s->ctx;
	PPX_REQUIRE(ctx, nearest_obstacle && nearest_edge, "ppx_smoother_set_gvd needs both cell grids");
	const size_t bytes = (size_t)ctx->desc.rows * ctx->desc.cols * 2 * sizeof(int32_t);
	if (!s->d_near_obstacle) {
		PPX_CUDA(ctx, cudaMalloc(&s->d_near_obstacle, bytes));
		PPX_CUDA(ctx, cudaMalloc(&s->d_near_edge, bytes));
	}
	PPX_CUDA(ctx, cudaMemcpyAsync(s->d_near_obstacle, nearest_obstacle, bytes, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(s->d_near_edge, nearest_edge, bytes, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	s->has_maps = true;
	return PPX_OK;
}

extern "C" int ppx_smoother_use_gvd(ppx_smoother* s)
{
	PPX_GUARD(s ? s->ctx : nullptr);
	if (!s)
		return PPX_ERR_INVALID;
	ppx_ctx* ctx = s->ctx;
	const size_t bytes = (size_t)ctx->desc.rows * ctx->desc.cols * 2 * sizeof(int32_t);
	if (!s->d_near_obstacle) {
		PPX_CUDA(ctx, cudaMalloc(&s->d_near_obstacle, bytes));
		PPX_CUDA(ctx, cudaMalloc(&s->d_near_edge, bytes));
	}
	PPX_CHECK(ppx_gvd_unpack_cells(ctx, s->d_near_obstacle, s->d_near_edge));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	s->has_maps = true;
	return PPX_OK;
}

extern "C" int ppx_smoother_smooth(ppx_smoother* s, int64_t nq, const double* poses, const uint8_t* cusp, const int32_t* count,
	int32_t max_points, double* out_poses, int32_t* status, int32_t* iterations)
{
	PPX_GUARD(s ? s->ctx : nullptr);
	if (!s)
		return PPX_ERR_INVALID;
	ppx_ctx* ctx = s->ctx;
	PPX_REQUIRE(ctx, s->has_maps, "ppx_smoother_set_gvd has not been called");
	PPX_REQUIRE(ctx, nq >= 0 && max_points > 0, "bad arguments");
	PPX_REQUIRE(ctx, nq == 0 || (poses && cusp && count && out_poses && status), "bad arguments");
	if (nq == 0)
		return PPX_OK;
	int cap = 8;
	for (int64_t q = 0; q < nq; q++) {
		PPX_REQUIRE(ctx, count[q] <= max_points, "count[q] exceeds max_points");
		cap = std::max(cap, (int)count[q]);
	}
	cap = (cap + 7) & ~7;
	const size_t smem = (size_t)cap * kBytesPerPoint + 16;
	PPX_REQUIRE(ctx, smem <= 200 * 1024, "path too long for the smoother (<= 2400 points)");
	size_t off = 0;
	auto carve = [&](size_t bytes) {
		size_t o = off;
		off += (bytes + 255) & ~(size_t)255;
		return o;
	};
	const size_t bPoses = (size_t)nq * max_points * 24, bCusp = (size_t)nq * max_points, bCount = (size_t)nq * 4;
	size_t oIn = carve(bPoses), oCusp = carve(bCusp), oCount = carve(bCount), oOut = carve(bPoses), oStatus = carve(bCount),
		   oIter = carve(bCount);
	char* base;
	PPX_CHECK(ppx_scratch(ctx, 5, off, (void**)&base));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oIn, poses, bPoses, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oCusp, cusp, bCusp, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oCount, count, bCount, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemsetAsync(base + oOut, 0, bPoses, ctx->stream));
	SmoothConsts c;
	const ppx_smoother_params& p = s->params;
	c.P.stepTolerance = p.step_tolerance;
	c.P.maxIterations = p.max_iterations;
	c.P.learningRate = p.learning_rate;
	c.P.pathWeight = p.path_weight;
	c.P.smoothWeight = p.smooth_weight;
	c.P.voronoiWeight = p.voronoi_weight;
	c.P.collisionWeight = p.collision_weight;
	c.P.curvatureWeight = p.curvature_weight;
	c.P.collisionRatio = p.collision_ratio;
	c.P.maxCurvature = p.max_curvature;
	c.P.alpha = p.alpha;
	c.P.dMax = p.d_max;
	c.P.minSafeRadius = p.min_safe_radius;
	c.P.nearObstacle = s->d_near_obstacle;
	c.P.nearEdge = s->d_near_edge;
	c.map = ctx->map_dev();
	if (!s->d_consts)
		PPX_CUDA(ctx, cudaMalloc(&s->d_consts, sizeof(SmoothConsts)));
	PPX_CUDA(ctx, cudaMemcpyAsync(s->d_consts, &c, sizeof(SmoothConsts), cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaFuncSetAttribute(k_smooth, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
	int perSm = (int)std::max<size_t>(1, std::min<size_t>(8, (200 * 1024) / smem));
	int ctas = (int)std::min<int64_t>(nq, (int64_t)ctx->num_sms * perSm);
	k_smooth<<<ctas, kSmoothThreads, smem, ctx->stream>>>(s->d_consts, nq, (const double*)(base + oIn), (const uint8_t*)(base + oCusp),
		(const int32_t*)(base + oCount), max_points, cap, (double*)(base + oOut), (int32_t*)(base + oStatus), (int32_t*)(base + oIter));
	PPX_LAUNCHED(ctx);
	PPX_CUDA(ctx, cudaMemcpyAsync(out_poses, base + oOut, bPoses, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(status, base + oStatus, bCount, cudaMemcpyDeviceToHost, ctx->stream));
	if (iterations)
		PPX_CUDA(ctx, cudaMemcpyAsync(iterations, base + oIter, bCount, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}
