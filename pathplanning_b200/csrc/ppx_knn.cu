// libppx.so: exact brute-force k nearest neighbours (A10), the replacement of the flann kd-tree
// behind Tree::GetNearestNode / GetNearestNodes (utils/tree.h:73-116).
//
// Semantics kept from the reference's use of flann: metric L2_Simple = sum of squared coordinate
// differences over the raw state coordinates, accumulated in coordinate order (utils/tree.h:174),
// exact search, neighbours returned by increasing distance.  flann leaves the order of EQUAL
// distances unspecified; here ties go to the lower point index (the choice the oracle's flann
// stand-in makes too, oracle/shim/flann/flann.hpp).
//
// Two mappings.  Many queries: one thread per query, the point set streamed through shared memory
// in tiles that every thread of the CTA reuses (the dense distance tile of the batch); D is 2 or 3,
// so a tile costs 3 FP64 FMAs per pair -- there is no K dimension to feed tensor cores with, and the
// neighbour order must be exact in FP64 (DESIGN.md "kNN").  Few queries (the inner loop of a single
// RRT tree): one CTA per query, threads stride over the points and a k-round block argmin merges
// their private lists.
//
// The per-thread candidate list has a COMPILE-TIME length K >= k (1, 4, 8, 16 or 32): it lives in registers with static
// indexing (insertion = one pass of compare-and-swap with a carried element, pop = shift).  With a run-time length it sat
// in local memory (384-400 byte stack frames) and every pair paid a local load for the threshold compare: the round-1
// kernels issued on 31 % of the cycles with 3.3 long-scoreboard stall cycles per issue (profiles/r2o_k_knn_*).
#include <algorithm>
#include "ppx_ctx.h"

namespace {

constexpr int kMaxK = 32;
constexpr int kTile = 1024; // points per shared-memory tile (24 KB at D = 3)

__device__ __forceinline__ bool closer(double da, int ia, double db, int ib) { return da < db || (da == db && ia < ib); }

constexpr int kNoIdx = 0x7fffffff;

// ascending candidate list of K entries in registers
template <int K>
struct RegList {
	double d[K];
	int idx[K];
	__device__ __forceinline__ void clear()
	{
		#pragma unroll
		for (int i = 0; i < K; i++) {
			d[i] = __longlong_as_double(0x7ff0000000000000LL);
			idx[i] = kNoIdx;
		}
	}
	// keeps the k smallest (k <= K): entries beyond k are never read, so inserting against the whole list is harmless,
	// but the early-out threshold is the k-th entry
	__device__ __forceinline__ void offer(double nd, int nidx, int k)
	{
		if (!closer(nd, nidx, d[K - 1], idx[K - 1]))
			return;
		#pragma unroll
		for (int i = 0; i < K; i++) {
			if (closer(nd, nidx, d[i], idx[i])) {
				const double td = d[i];
				const int ti = idx[i];
				d[i] = nd;
				idx[i] = nidx;
				nd = td;
				nidx = ti;
			}
		}
		(void)k;
	}
	__device__ __forceinline__ void pop()
	{
		#pragma unroll
		for (int i = 0; i + 1 < K; i++) {
			d[i] = d[i + 1];
			idx[i] = idx[i + 1];
		}
		d[K - 1] = __longlong_as_double(0x7ff0000000000000LL);
		idx[K - 1] = kNoIdx;
	}
};

template <int D>
__device__ __forceinline__ double dist2(const double* __restrict__ a, const double* __restrict__ b)
{
	double r = 0.0;
	#pragma unroll
	for (int i = 0; i < D; i++) {
		double diff = a[i] - b[i];
		r += diff * diff; // flann::L2_Simple: result += diff * diff
	}
	return r;
}

template <int D, int K>
__global__ void __launch_bounds__(256) k_knn_thread_per_query(int64_t np, const double* __restrict__ points, int64_t nq,
	const double* __restrict__ queries, int k, int* __restrict__ indices, double* __restrict__ dists)
{
	__shared__ double s_pts[kTile * D];
	const int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
	const bool active = q < nq;
	double qv[D];
	#pragma unroll
	for (int i = 0; i < D; i++)
		qv[i] = active ? queries[q * D + i] : 0.0;
	RegList<K> best;
	best.clear();
	for (int64_t base = 0; base < np; base += kTile) {
		const int m = (int)min((int64_t)kTile, np - base);
		__syncthreads();
		for (int i = threadIdx.x; i < m * D; i += blockDim.x)
			s_pts[i] = points[base * D + i];
		__syncthreads();
		if (!active)
			continue;
		#pragma unroll 4
		for (int j = 0; j < m; j++)
			best.offer(dist2<D>(qv, s_pts + j * D), (int)(base + j), k);
	}
	if (active) {
		#pragma unroll
		for (int i = 0; i < K; i++) {
			if (i < k) {
				const bool have = best.idx[i] != kNoIdx;
				indices[q * k + i] = have ? best.idx[i] : -1;
				if (dists)
					dists[q * k + i] = best.d[i];
			}
		}
	}
}

// One CTA per query.
template <int D, int K>
__global__ void __launch_bounds__(256) k_knn_cta_per_query(int64_t np, const double* __restrict__ points, int64_t nq,
	const double* __restrict__ queries, int k, int* __restrict__ indices, double* __restrict__ dists)
{
	__shared__ double s_d[8];
	__shared__ int s_i[8];
	for (int64_t q = blockIdx.x; q < nq; q += gridDim.x) {
		double qv[D];
		#pragma unroll
		for (int i = 0; i < D; i++)
			qv[i] = queries[q * D + i];
		RegList<K> best;
		best.clear();
		for (int64_t j = threadIdx.x; j < np; j += blockDim.x)
			best.offer(dist2<D>(qv, points + j * D), (int)j, k);
		// k-way merge of the private ascending lists: k rounds of block argmin over the list heads; point indices are
		// unique, so exactly the thread whose head is the minimum pops
		for (int round = 0; round < k; round++) {
			double d = best.d[0];
			int idx = best.idx[0];
			#pragma unroll
			for (int o = 16; o > 0; o >>= 1) {
				const double od = __shfl_xor_sync(0xffffffffu, d, o);
				const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
				if (closer(od, oi, d, idx)) {
					d = od;
					idx = oi;
				}
			}
			if ((threadIdx.x & 31) == 0) {
				s_d[threadIdx.x >> 5] = d;
				s_i[threadIdx.x >> 5] = idx;
			}
			__syncthreads();
			double bd = s_d[0];
			int bi = s_i[0];
			#pragma unroll
			for (int w = 1; w < 8; w++)
				if (closer(s_d[w], s_i[w], bd, bi)) {
					bd = s_d[w];
					bi = s_i[w];
				}
			if (threadIdx.x == 0) {
				indices[q * k + round] = bi != kNoIdx ? bi : -1;
				if (dists)
					dists[q * k + round] = bd;
			}
			if (bi != kNoIdx && best.idx[0] == bi)
				best.pop();
			__syncthreads();
		}
	}
}

template <int D>
void knn_launch(ppx_ctx* ctx, bool perThread, int blocks, int64_t np, const double* points, int64_t nq, const double* queries, int k,
	int* indices, double* dists)
{
	#define PPX_KNN_CASE(K) \
		if (perThread) \
			k_knn_thread_per_query<D, K><<<blocks, 256, 0, ctx->stream>>>(np, points, nq, queries, k, indices, dists); \
		else \
			k_knn_cta_per_query<D, K><<<blocks, 256, 0, ctx->stream>>>(np, points, nq, queries, k, indices, dists);
	if (k == 1) {
		PPX_KNN_CASE(1)
	} else if (k <= 4) {
		PPX_KNN_CASE(4)
	} else if (k <= 8) {
		PPX_KNN_CASE(8)
	} else if (k <= 16) {
		PPX_KNN_CASE(16)
	} else {
		PPX_KNN_CASE(32)
	}
	#undef PPX_KNN_CASE
}

}

extern "C" {

int ppx_knn_dev(ppx_ctx* ctx, int32_t dim, int64_t npoints, const double* d_points, int64_t nqueries, const double* d_queries,
	int32_t k, int32_t* d_indices, double* d_dist2)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	PPX_REQUIRE(ctx, (dim == 2 || dim == 3) && npoints >= 0 && nqueries >= 0 && k >= 1 && k <= kMaxK, "knn: dim must be 2 or 3, 1 <= k <= 32");
	PPX_REQUIRE(ctx, nqueries == 0 || (d_queries && d_indices && (npoints == 0 || d_points)), "bad arguments");
	PPX_REQUIRE(ctx, npoints < 0x7fffffff, "knn: too many points");
	if (nqueries == 0)
		return PPX_OK;
	const bool perThread = nqueries >= 4 * (int64_t)ctx->num_sms;
	const int blocks = perThread ? (int)((nqueries + 255) / 256) : (int)nqueries;
	if (dim == 2)
		knn_launch<2>(ctx, perThread, blocks, npoints, d_points, nqueries, d_queries, k, d_indices, d_dist2);
	else
		knn_launch<3>(ctx, perThread, blocks, npoints, d_points, nqueries, d_queries, k, d_indices, d_dist2);
	PPX_LAUNCHED(ctx);
	return PPX_OK;
}

int ppx_knn(ppx_ctx* ctx, int32_t dim, int64_t npoints, const double* points, int64_t nqueries, const double* queries, int32_t k,
	int32_t* indices, double* dist2)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	PPX_REQUIRE(ctx, (dim == 2 || dim == 3) && npoints >= 0 && nqueries >= 0 && k >= 1 && k <= kMaxK, "knn: dim must be 2 or 3, 1 <= k <= 32");
	PPX_REQUIRE(ctx, nqueries == 0 || (queries && indices && (npoints == 0 || points)), "bad arguments");
	if (nqueries == 0)
		return PPX_OK;
	double *d_points, *d_queries, *d_dist;
	int32_t* d_idx;
	PPX_CHECK(ppx_scratch(ctx, 0, std::max<int64_t>(1, npoints) * dim * sizeof(double), (void**)&d_points));
	PPX_CHECK(ppx_scratch(ctx, 1, nqueries * dim * sizeof(double), (void**)&d_queries));
	PPX_CHECK(ppx_scratch(ctx, 2, nqueries * k * sizeof(int32_t), (void**)&d_idx));
	PPX_CHECK(ppx_scratch(ctx, 3, nqueries * k * sizeof(double), (void**)&d_dist));
	if (npoints)
		PPX_CUDA(ctx, cudaMemcpyAsync(d_points, points, npoints * dim * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(d_queries, queries, nqueries * dim * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
	PPX_CHECK(ppx_knn_dev(ctx, dim, npoints, d_points, nqueries, d_queries, k, d_idx, d_dist));
	PPX_CUDA(ctx, cudaMemcpyAsync(indices, d_idx, nqueries * k * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
	if (dist2)
		PPX_CUDA(ctx, cudaMemcpyAsync(dist2, d_dist, nqueries * k * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

} // extern "C"
