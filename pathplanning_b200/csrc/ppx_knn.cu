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
#include <algorithm>
#include "ppx_ctx.h"

namespace {

constexpr int kMaxK = 32;
constexpr int kTile = 1024; // points per shared-memory tile (24 KB at D = 3)

struct Best {
	double d[kMaxK];
	int idx[kMaxK];
};

__device__ __forceinline__ bool closer(double da, int ia, double db, int ib) { return da < db || (da == db && ia < ib); }

// insert (d, idx) into the ascending list of length k (the caller checked it beats the last entry)
__device__ __forceinline__ void insert_sorted(double* bd, int* bi, int k, double d, int idx)
{
	int p = k - 1;
	while (p > 0 && closer(d, idx, bd[p - 1], bi[p - 1])) {
		bd[p] = bd[p - 1];
		bi[p] = bi[p - 1];
		p--;
	}
	bd[p] = d;
	bi[p] = idx;
}

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

template <int D>
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
	double bd[kMaxK];
	int bi[kMaxK];
	for (int i = 0; i < k; i++) {
		bd[i] = __longlong_as_double(0x7ff0000000000000LL);
		bi[i] = 0x7fffffff;
	}
	for (int64_t base = 0; base < np; base += kTile) {
		const int m = (int)min((int64_t)kTile, np - base);
		__syncthreads();
		for (int i = threadIdx.x; i < m * D; i += blockDim.x)
			s_pts[i] = points[base * D + i];
		__syncthreads();
		if (!active)
			continue;
		for (int j = 0; j < m; j++) {
			double d = dist2<D>(qv, s_pts + j * D);
			int idx = (int)(base + j);
			if (closer(d, idx, bd[k - 1], bi[k - 1]))
				insert_sorted(bd, bi, k, d, idx);
		}
	}
	if (active)
		for (int i = 0; i < k; i++) {
			bool have = bi[i] != 0x7fffffff;
			indices[q * k + i] = have ? bi[i] : -1;
			if (dists)
				dists[q * k + i] = bd[i];
		}
}

// One CTA per query.
template <int D>
__global__ void __launch_bounds__(256) k_knn_cta_per_query(int64_t np, const double* __restrict__ points, int64_t nq,
	const double* __restrict__ queries, int k, int* __restrict__ indices, double* __restrict__ dists)
{
	__shared__ double s_d[8];
	__shared__ int s_i[8];
	__shared__ int s_owner[8];
	__shared__ int s_winner;
	for (int64_t q = blockIdx.x; q < nq; q += gridDim.x) {
		double qv[D];
		#pragma unroll
		for (int i = 0; i < D; i++)
			qv[i] = queries[q * D + i];
		double bd[kMaxK];
		int bi[kMaxK];
		for (int i = 0; i < k; i++) {
			bd[i] = __longlong_as_double(0x7ff0000000000000LL);
			bi[i] = 0x7fffffff;
		}
		for (int64_t j = threadIdx.x; j < np; j += blockDim.x) {
			double d = dist2<D>(qv, points + j * D);
			if (closer(d, (int)j, bd[k - 1], bi[k - 1]))
				insert_sorted(bd, bi, k, d, (int)j);
		}
		// k-way merge of the private ascending lists: k rounds of block argmin over the list heads
		int head = 0;
		for (int round = 0; round < k; round++) {
			double d = head < k ? bd[head] : __longlong_as_double(0x7ff0000000000000LL);
			int idx = head < k ? bi[head] : 0x7fffffff;
			int owner = threadIdx.x;
			#pragma unroll
			for (int o = 16; o > 0; o >>= 1) {
				double od = __shfl_down_sync(0xffffffffu, d, o);
				int oi = __shfl_down_sync(0xffffffffu, idx, o);
				int oo = __shfl_down_sync(0xffffffffu, owner, o);
				if (closer(od, oi, d, idx)) {
					d = od;
					idx = oi;
					owner = oo;
				}
			}
			if ((threadIdx.x & 31) == 0) {
				s_d[threadIdx.x >> 5] = d;
				s_i[threadIdx.x >> 5] = idx;
				s_owner[threadIdx.x >> 5] = owner;
			}
			__syncthreads();
			if (threadIdx.x == 0) {
				int best = 0;
				for (int w = 1; w < (int)(blockDim.x >> 5); w++)
					if (closer(s_d[w], s_i[w], s_d[best], s_i[best]))
						best = w;
				bool have = s_i[best] != 0x7fffffff;
				indices[q * k + round] = have ? s_i[best] : -1;
				if (dists)
					dists[q * k + round] = s_d[best];
				s_winner = have ? s_owner[best] : -1;
			}
			__syncthreads();
			if ((int)threadIdx.x == s_winner)
				head++;
			__syncthreads();
		}
	}
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
	if (nqueries >= 4 * (int64_t)ctx->num_sms) {
		int blocks = (int)((nqueries + 255) / 256);
		if (dim == 2)
			k_knn_thread_per_query<2><<<blocks, 256, 0, ctx->stream>>>(npoints, d_points, nqueries, d_queries, k, d_indices, d_dist2);
		else
			k_knn_thread_per_query<3><<<blocks, 256, 0, ctx->stream>>>(npoints, d_points, nqueries, d_queries, k, d_indices, d_dist2);
	} else {
		int blocks = (int)nqueries;
		if (dim == 2)
			k_knn_cta_per_query<2><<<blocks, 256, 0, ctx->stream>>>(npoints, d_points, nqueries, d_queries, k, d_indices, d_dist2);
		else
			k_knn_cta_per_query<3><<<blocks, 256, 0, ctx->stream>>>(npoints, d_points, nqueries, d_queries, k, d_indices, d_dist2);
	}
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
