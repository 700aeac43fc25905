// libppx.so: Hybrid A* on the device -- NonHolonomicHeuristic table build (A9), batched successor
// generation (A1/A2/A5/A6), heuristic lookups (A8/A9) and the batched graph search, one warp per
// query (algo/hybrid_a_star.cpp:59-173,206-257, algo/a_star.h:326-434).
//
// Search kernel mapping: a query is a chain of dependent expansions (pop -> <=11 children ->
// frontier updates), so the parallelism inside a query is the children of one node: lane c of the
// query's warp generates child c (constant-steer rollout + adaptive collision check + Voronoi term
// + heuristics, or the Reeds-Shepp shot), then lane 0 commits the children IN THE REFERENCE'S ORDER
// to the frontier heap / key table, which is what makes expansion order -- and hence the expanded
// set -- reproducible.  Throughput comes from thousands of independent queries in flight (one warp
// each, persistent grid fed from one atomic cursor).
#include <algorithm>
#include <unordered_map>
#include <vector>
#include <cuda_runtime.h>
// One out-of-line sincos for this translation unit: the search kernel's warps each sit in a different
// phase of a different query, so its instruction footprint -- not arithmetic -- is what stalls it
// (ncu r1c: 6.6 "no instruction" stall cycles per issued instruction with everything inlined).
__device__ __noinline__ void ppx_sincos_ool(double x, double* s, double* c);
__device__ __noinline__ double ppx_fmod_ool(double a, double b);
__device__ __noinline__ double ppx_atan2_ool(double y, double x);
__device__ __noinline__ double ppx_acos_ool(double x);
__device__ __noinline__ double ppx_asin_ool(double x);
__device__ __noinline__ double ppx_sin_ool(double x);
#define PPX_SINCOS ppx_sincos_ool
#define PPX_FMOD ppx_fmod_ool
#define PPX_ATAN2 ppx_atan2_ool
#define PPX_ACOS ppx_acos_ool
#define PPX_ASIN ppx_asin_ool
#define PPX_SIN ppx_sin_ool
#include <cstdlib>
#include "ppx_ctx.h"
#include "ppx_hybrid_host.h"
#include "ppx_resample.cuh"

using namespace ppx;

__device__ __noinline__ void ppx_sincos_ool(double x, double* s, double* c) { sincos(x, s, c); }
__device__ __noinline__ double ppx_fmod_ool(double a, double b) { return fmod(a, b); }
__device__ __noinline__ double ppx_atan2_ool(double y, double x) { return atan2(y, x); }
__device__ __noinline__ double ppx_acos_ool(double x) { return acos(x); }
__device__ __noinline__ double ppx_asin_ool(double x) { return asin(x); }
__device__ __noinline__ double ppx_sin_ool(double x) { return sin(x); }

int ppx_wavefront_launch(ppx_ctx* ctx, int64_t n, const double* d_goals, float* d_cost, uint8_t* d_explored);

struct ppx_hybrid {
	ppx_ctx* ctx = nullptr;
	ppx_hybrid_params params {};
	ppx_validator_params vp {};
	HybridDev dev {};
	double* d_table = nullptr;
	int x0 = 0, y0 = 0, tx = 0, ty = 0;
	// grow-only device buffers
	void* d_slabs = nullptr;
	size_t slabBytes = 0;
	float* d_wf = nullptr;
	size_t wfBytes = 0;
	void* d_io = nullptr;
	size_t ioBytes = 0;
	void* d_consts = nullptr;
	cudaEvent_t ev[3] {}; // start / wavefronts done / search done of the last ppx_hybrid_search[_dev] call
	bool timed = false;
};

namespace {

constexpr int kGenSlots = 3; // children sets per warp in shared memory (see k_hybrid_search)
constexpr int kSearchWarps = 4; // warps (queries) per CTA of the one-warp-per-query engine

// NonHolonomicHeuristic::Build, heuristics.cpp:62-73: one Reeds-Shepp solve per table entry
__global__ void __launch_bounds__(128) k_nonholo_build(HybridDev P, double* __restrict__ values)
{
	const size_t total = (size_t)P.nhX * P.nhY * P.nhA;
	size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
	const size_t stride = (size_t)gridDim.x * blockDim.x;
	for (; idx < total; idx += stride) {
		int k = (int)(idx % P.nhA);
		int j = (int)((idx / P.nhA) % P.nhY);
		int i = (int)(idx / ((size_t)P.nhA * P.nhY));
		// Pose2d pose(i * spatialResolution - offsetX, j * spatialResolution - offsetY, k * angularResolution): ctor wraps
		Pose pose = { i * P.spatialRes - P.nhOffX, j * P.spatialRes - P.nhOffY, wrap_theta(k * P.angRes) };
		Pose goal = { 0.0, 0.0, 0.0 };
		RSSolution sol;
		rs_solve(pose, goal, P.rmin, P.revF, P.fwdF, P.swF, 0, sol);
		values[idx] = (double)sol.cost;
	}
}

__global__ void __launch_bounds__(128) k_hybrid_heuristic(HybridDev P, MapDev map, const float* __restrict__ wfCost, Pose goal,
	int64_t n, const double* __restrict__ poses, double* __restrict__ combined, double* __restrict__ nonholo,
	double* __restrict__ obstacles)
{
	int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
	const int64_t stride = (int64_t)gridDim.x * blockDim.x;
	for (; i < n; i += stride) {
		Pose s = { poses[3 * i], poses[3 * i + 1], poses[3 * i + 2] };
		double h0 = nonholo_heuristic(P, goal, s);
		double h1 = obstacle_heuristic(P, map, wfCost, goal, s);
		if (nonholo)
			nonholo[i] = h0;
		if (obstacles)
			obstacles[i] = h1;
		if (combined)
			combined[i] = combined_heuristic(P, map, wfCost, goal, s);
	}
}

// GetNeighborStates for arbitrary parents: one warp per parent, lane c = child slot c.
__global__ void __launch_bounds__(128) k_hybrid_expand(HybridDev P, MapDev map, const float* __restrict__ wfCost, Pose goal,
	int64_t n, const double* __restrict__ parents, const unsigned long long* __restrict__ seeds,
	int32_t* __restrict__ nChildren, ppx_hybrid_child* __restrict__ children, double* __restrict__ parentHeuristic,
	unsigned long long* __restrict__ mtScratch)
{
	const unsigned lane = threadIdx.x & 31;
	const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
	const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
	for (int64_t i = warp; i < n; i += nwarps) {
		Pose parent = { parents[3 * i], parents[3 * i + 1], parents[3 * i + 2] };
		Key3 pk = discretize_pose(P, parent);
		double hCost = combined_heuristic(P, map, wfCost, goal, parent);
		int want = 0;
		if (lane == 0) {
			int idx = 0;
			Mt64 rng { mtScratch + (size_t)warp * 312, &idx };
			rng.seed(seeds[i]);
			want = rs_shot_wanted(rng, hCost) ? 1 : 0;
			if (parentHeuristic)
				parentHeuristic[i] = hCost;
		}
		want = __shfl_sync(0xffffffffu, want, 0);
		const int nslots = 2 * P.numDeltas + want;
		Child c;
		c.ok = false;
		double h = 0.0;
		if ((int)lane < nslots) {
			make_child(P, map, parent, pk, goal, (int)lane, c);
			if (c.ok)
				h = combined_heuristic(P, map, wfCost, goal, c.pose);
		}
		unsigned okMask = __ballot_sync(0xffffffffu, c.ok);
		if (c.ok) {
			int pos = __popc(okMask & ((1u << lane) - 1));
			ppx_hybrid_child& o = children[i * PPX_MAX_CHILDREN + pos];
			o.pose[0] = c.pose.x;
			o.pose[1] = c.pose.y;
			o.pose[2] = c.pose.th;
			o.cost = c.cost;
			o.heuristic = h;
			o.length = c.length;
			o.key[0] = c.key.x;
			o.key[1] = c.key.y;
			o.key[2] = c.key.k;
			o.direction = c.dir;
			o.kind = c.kind;
			o.last = c.last;
		}
		if (lane == 0)
			nChildren[i] = __popc(okMask);
		__syncwarp();
	}
}

}

namespace {

struct WsLayout {
	size_t bytes; // per warp
	int maxNodes, tableSize;
	int x0, y0, tx, ty;
};

__host__ __device__ inline size_t align16(size_t v) { return (v + 15) & ~(size_t)15; }

__host__ __device__ inline size_t ws_bytes(int maxNodes, int tableSize)
{
	size_t n = (size_t)maxNodes;
	return align16(n * 24) + 5 * align16(n * 8) + 6 * align16(n * 4) + align16((size_t)tableSize * 4) + align16(312 * 8);
}

__device__ inline SearchWs ws_carve(char* base, const WsLayout& L)
{
	SearchWs W;
	size_t n = (size_t)L.maxNodes;
	char* p = base;
	W.pose = (double*)p; p += align16(n * 24);
	W.g = (double*)p; p += align16(n * 8);
	W.f = (double*)p; p += align16(n * 8);
	W.h = (double*)p; p += align16(n * 8);
	W.length = (double*)p; p += align16(n * 8);
	W.heapF = (double*)p; p += align16(n * 8);
	W.mt = (unsigned long long*)p; p += align16(312 * 8);
	W.parent = (int*)p; p += align16(n * 4);
	W.keyIdx = (int*)p; p += align16(n * 4);
	W.meta = (uint32_t*)p; p += align16(n * 4);
	W.last = (float*)p; p += align16(n * 4);
	W.heapSeq = (uint32_t*)p; p += align16(n * 4);
	W.heapNode = (uint32_t*)p; p += align16(n * 4);
	W.table = (uint32_t*)p;
	W.maxNodes = L.maxNodes;
	W.x0 = L.x0; W.y0 = L.y0; W.tx = L.tx; W.ty = L.ty;
	return W;
}

// Out-of-line copies for the search kernel (see the note on PPX_SINCOS above).
__device__ __noinline__ double hy_heuristic(const HybridDev& P, const MapDev& map, const float* wf, const Pose& goal, const Pose& s)
{
	return combined_heuristic(P, map, wf, goal, s);
}
__device__ __noinline__ void hy_rs_child(const HybridDev& P, const MapDev& map, const Pose& parent, const RSSolution& sol, Child& out,
	const FastMap* fm)
{
	rs_child_from(P, map, parent, sol, out, fm);
}
__device__ __noinline__ void hy_rs_solve_warp(const HybridDev& P, const Pose& parent, const Pose& goal, RSSolution& sol)
{
	rs_solve_warp(parent, goal, P.rmin, P.revF, P.fwdF, P.swF, sol);
}

// Parameters of the search kernel live in device memory: the out-of-line helpers take them by
// reference, and a reference to a by-value kernel parameter would force a local-memory copy.
struct SearchConsts {
	HybridDev P;
	MapDev map;
	FastMap fm; // FP32 cell filter of the primitives' collision checks (ppx_fastlane.cuh)
};

// N3: hybrid_a_star.cpp:260-291 for a batch of solved queries, one warp per query: lane 0 plans (composite
// length, cusp ratios, sample ratios -- a sequential double accumulation), the 32 lanes interpolate.
constexpr int kResampleWarps = 4;
__global__ void __launch_bounds__(kResampleWarps * 32) k_hybrid_resample(const SearchConsts* __restrict__ consts, int64_t nq,
	const double* __restrict__ goals, const ppx_hybrid_result* __restrict__ results, const ppx_hybrid_path_state* __restrict__ paths,
	int maxPath, float pathInterpolation, int maxPoints, double* __restrict__ finalLength, double* __restrict__ poses,
	uint8_t* __restrict__ cusp, int32_t* __restrict__ count, double* __restrict__ lengths)
{
	__shared__ ResamplePath s_rp[kResampleWarps];
	const HybridDev& P = consts->P;
	const unsigned lane = threadIdx.x & 31;
	const int wib = threadIdx.x >> 5;
	for (int64_t q = (int64_t)blockIdx.x * kResampleWarps + wib; q < nq; q += (int64_t)gridDim.x * kResampleWarps) {
		double* out = poses + (size_t)q * maxPoints * 3;
		uint8_t* cq = cusp + (size_t)q * maxPoints;
		int n = 0;
		if (lane == 0) {
			const ppx_hybrid_result& r = results[q];
			s_rp[wib].length = 0.0;
			if (r.status == 0 && r.path_len > maxPath)
				n = -1; // the chain was cut by max_path
			else if (r.status == 0 && r.path_len >= 2) {
				const Pose goal = { goals[3 * q], goals[3 * q + 1], goals[3 * q + 2] };
				n = resample_plan(P, paths + (size_t)q * maxPath, r.path_len, goal, pathInterpolation,
					finalLength + (size_t)q * maxPath, s_rp[wib], out, cq, maxPoints);
			}
			count[q] = n;
			if (lengths)
				lengths[q] = s_rp[wib].length;
		}
		__syncwarp();
		n = __shfl_sync(0xffffffffu, n, 0);
		for (int i = (int)lane; i < n; i += 32) {
			Pose p = resample_pose(P, s_rp[wib], out[3 * i]);
			out[3 * i] = p.x;
			out[3 * i + 1] = p.y;
			out[3 * i + 2] = p.th;
		}
		__syncwarp();
	}
}

// Schedule of a batch: queries are independent, and the batch ends when its longest query does.  A
// query whose start cell the goal's wavefront never reached exhausts its whole reachable state set
// (about 7000 expansions on C2, 10x the mean), and otherwise the expansion count grows with the
// wavefront cost-to-go of the start cell -- both are known before the search starts.  One CTA
// bucket-sorts the queries by that key, largest first (unreached = +inf first).
constexpr int kOrderBins = 1024;
__global__ void __launch_bounds__(1024) k_query_order(MapDev map, int64_t nq, const double* __restrict__ starts,
	const int* __restrict__ wfSlot, const float* __restrict__ wfCost, float* __restrict__ keys, int* __restrict__ order)
{
	__shared__ unsigned s_bins[kOrderBins + 1];
	__shared__ float s_max;
	const size_t cells = (size_t)map.rows * map.cols;
	const float INF = __int_as_float(0x7f800000);
	if (threadIdx.x == 0)
		s_max = 0.0f;
	for (int i = threadIdx.x; i <= kOrderBins; i += blockDim.x)
		s_bins[i] = 0u;
	__syncthreads();
	for (int64_t q = threadIdx.x; q < nq; q += blockDim.x) {
		int row = trunc_i32((starts[3 * q] - map.ox) / map.resd);
		int col = trunc_i32((starts[3 * q + 1] - map.oy) / map.resd);
		float key = INF;
		if (row >= 0 && row < map.rows && col >= 0 && col < map.cols)
			key = wfCost[(size_t)(wfSlot ? wfSlot[q] : q) * cells + (size_t)row * map.cols + col];
		if (!(key >= 0.0f))
			key = INF;
		keys[q] = key;
		if (key < INF)
			atomicMax((int*)&s_max, __float_as_int(key)); // non-negative floats order like their bits
	}
	__syncthreads();
	const float scale = s_max > 0.0f ? (float)(kOrderBins - 1) / s_max : 0.0f;
	for (int64_t q = threadIdx.x; q < nq; q += blockDim.x) {
		float key = keys[q];
		int bin = key < INF ? min(kOrderBins - 1, (int)(key * scale)) : kOrderBins;
		atomicAdd(&s_bins[bin], 1u);
	}
	__syncthreads();
	if (threadIdx.x == 0) { // descending exclusive scan (1025 bins)
		unsigned acc = 0;
		for (int b = kOrderBins; b >= 0; b--) {
			unsigned c = s_bins[b];
			s_bins[b] = acc;
			acc += c;
		}
	}
	__syncthreads();
	for (int64_t q = threadIdx.x; q < nq; q += blockDim.x) {
		float key = keys[q];
		int bin = key < INF ? min(kOrderBins - 1, (int)(key * scale)) : kOrderBins;
		order[atomicAdd(&s_bins[bin], 1u)] = (int)q;
	}
}

#include "ppx_hybrid_cta.cuh"

// The graph search of one query per warp (ppx_ctx_set_option "search_engine" = 0; the default engine is the
// CTA-per-query kernel above, same results).
__global__ void __launch_bounds__(kSearchWarps * 32) k_hybrid_search(const SearchConsts* __restrict__ consts, int64_t nq,
	const double* __restrict__ starts, const double* __restrict__ goals, const unsigned long long* __restrict__ seeds,
	const int* __restrict__ wfSlot, const float* __restrict__ wfCost, WsLayout L, char* __restrict__ slabs,
	ppx_hybrid_limits limits, ppx_hybrid_result* __restrict__ results, ppx_hybrid_path_state* __restrict__ paths,
	ppx_hybrid_log_entry* __restrict__ logs, unsigned long long* __restrict__ queue, const int* __restrict__ order)
{
	// Children of up to kGenSlots nodes per warp: slot 2 holds the children of the node being expanded when
	// they had to be generated now, slots 0/1 cache the children of frontier nodes generated AHEAD of their
	// expansion (children are a pure function of the node; only the Reeds-Shepp shot depends on the RNG order).
	// A generation pass therefore fills up to 30 lanes (3 nodes x 10 primitives) instead of 10, and an
	// expansion whose node is cached is pop + commit only.
	__shared__ Child s_child[kSearchWarps][kGenSlots][kMaxChildren - 1];
	__shared__ double s_h[kSearchWarps][kGenSlots][kMaxChildren - 1];
	__shared__ Child s_rs[kSearchWarps];
	__shared__ double s_rsh[kSearchWarps];
	const HybridDev& P = consts->P;
	const MapDev& map = consts->map;
	const unsigned lane = threadIdx.x & 31;
	const int wib = threadIdx.x >> 5;
	const size_t gwarp = (size_t)blockIdx.x * kSearchWarps + wib;
	const SearchWs W = ws_carve(slabs + gwarp * L.bytes, L);
	const size_t cells = (size_t)map.rows * map.cols;
	const int nCs = 2 * P.numDeltas; // constant-steer children per node
	const int perPass = min(kGenSlots, 32 / nCs); // nodes one generation pass can take

	while (true) {
		long long q = 0;
		if (lane == 0)
			q = (long long)atomicAdd(queue, 1ull);
		q = __shfl_sync(0xffffffffu, q, 0);
		if (q >= nq)
			break;
		if (order)
			q = order[q]; // longest-expected-first schedule (k_query_order); results do not depend on it
		for (int i = (int)lane; i < L.tableSize; i += 32)
			W.table[i] = 0u;
		__syncwarp();
		const Pose start = { starts[3 * q], starts[3 * q + 1], starts[3 * q + 2] };
		const Pose goal = { goals[3 * q], goals[3 * q + 1], goals[3 * q + 2] };
		const float* wf = wfCost + (size_t)(wfSlot ? wfSlot[q] : q) * cells;
		SearchState S;
		Mt64 rng { W.mt, &S.mtIdx };
		if (lane == 0) {
			rng.seed(seeds ? seeds[q] : 0ull);
			search_init(P, W, S, start, hy_heuristic(P, map, wf, goal, start));
		}
		__syncwarp();
		int status = __shfl_sync(0xffffffffu, S.status, 0);
		ppx_hybrid_log_entry* log = (logs && limits.max_log > 0) ? logs + (size_t)q * limits.max_log : nullptr;
		int cache0 = -1, cache1 = -1; // nodes whose children sit in slots 0 / 1 (warp-uniform)

		while (status == PPX_SEARCH_RUNNING) {
			int n = -1;
#ifdef PPX_SEARCH_STATS
			long long tk0 = clock64();
#endif
			if (lane == 0) {
				if (limits.max_expansions > 0 && S.nExpanded >= limits.max_expansions)
					S.status = PPX_ERR_LIMIT;
				else
					n = search_pop(W, S, goal);
			}
			n = __shfl_sync(0xffffffffu, n, 0);
			if (n < 0)
				break;
			// node n was written by lane 0 in an earlier iteration (ordered by the __syncwarp below)
			const Pose parent = { W.pose[3 * n], W.pose[3 * n + 1], W.pose[3 * n + 2] };
			const Key3 pk = discretize_pose(P, parent);
			const double hCost = W.h[n];
			int want = 0;
			if (lane == 0) {
				if (log && S.nExpanded <= limits.max_log) {
					ppx_hybrid_log_entry& e = log[S.nExpanded - 1];
					e.key[0] = pk.x;
					e.key[1] = pk.y;
					e.key[2] = pk.k;
					e.direction = (int)(W.meta[n] & 3u);
					e.pose[0] = parent.x;
					e.pose[1] = parent.y;
					e.pose[2] = parent.th;
				}
				want = rs_shot_wanted(rng, hCost) ? 1 : 0;
				S.nRsAttempts += want;
			}
			want = __shfl_sync(0xffffffffu, want, 0);
#ifdef PPX_SEARCH_STATS
			long long tk1 = clock64();
#endif
			RSSolution sol;
			if (want) // warp-uniform: the 48 Reeds-Shepp words are evaluated by the 32 lanes
				hy_rs_solve_warp(P, parent, goal, sol);
#ifdef PPX_SEARCH_STATS
			long long tk2 = clock64();
#endif
			const int slotP = cache0 == n ? 0 : (cache1 == n ? 1 : 2);
#ifdef PPX_SEARCH_STATS
			if (lane == 0) atomicAdd(queue + 8 + (slotP == 2 ? 0 : 1), 1ull);
#endif
			if (slotP == 2) {
				// generation pass: node n (slot 2) and, if there are lanes to spare, the two best frontier nodes
				int candA = -1, candB = -1;
				if (perPass >= 2 && lane == 0) {
					if (S.heapSize > 0)
						candA = (int)W.heapNode[0];
					if (perPass >= 3 && S.heapSize > 1) {
						int i = 1;
						if (S.heapSize > 2 && heap_before(W.heapF[2], W.heapSeq[2], W.heapF[1], W.heapSeq[1]))
							i = 2;
						candB = (int)W.heapNode[i];
					}
					if (candA >= 0 && (W.meta[candA] & 0x100u))
						candA = -1;
					if (candB >= 0 && (W.meta[candB] & 0x100u))
						candB = -1;
				}
				candA = __shfl_sync(0xffffffffu, candA, 0);
				candB = __shfl_sync(0xffffffffu, candB, 0);
				// keep cached candidates where they are, give the others the remaining slots
				const bool keep0 = cache0 >= 0 && (cache0 == candA || cache0 == candB);
				const bool keep1 = cache1 >= 0 && (cache1 == candA || cache1 == candB);
				int gen0 = -1, gen1 = -1; // nodes to generate into slot 0 / slot 1
				if (candA >= 0 && candA != cache0 && candA != cache1) {
					if (!keep0) gen0 = candA;
					else if (!keep1) gen1 = candA;
				}
				if (candB >= 0 && candB != cache0 && candB != cache1 && candB != candA) {
					if (!keep0 && gen0 < 0) gen0 = candB;
					else if (!keep1 && gen1 < 0) gen1 = candB;
				}
				if (!keep0) cache0 = gen0;
				if (!keep1) cache1 = gen1;
#ifdef PPX_SEARCH_STATS
				if (lane == 0) atomicAdd(queue + 10, (unsigned long long)((gen0 >= 0) + (gen1 >= 0)));
#endif
				const int t = (int)lane / nCs, cslot = (int)lane % nCs;
				const int node = t == 0 ? n : (t == 1 ? (gen0 >= 0 ? gen0 : gen1) : (gen0 >= 0 ? gen1 : -1));
				const int slot = t == 0 ? 2 : (node == gen0 ? 0 : 1);
				if (t < perPass && node >= 0) {
					Pose par = parent;
					Key3 key = pk;
					if (t > 0) {
						par = { W.pose[3 * node], W.pose[3 * node + 1], W.pose[3 * node + 2] };
						key = discretize_pose(P, par);
					}
					Child c;
					csteer_child(P, map, par, key, cslot >> 1, cslot & 1, c, &consts->fm);
					double h = 0.0;
					if (c.ok)
						h = hy_heuristic(P, map, wf, goal, c.pose);
					s_child[wib][slot][cslot] = c;
					s_h[wib][slot][cslot] = h;
				}
			}
#ifdef PPX_SEARCH_STATS
			__syncwarp();
			long long tk3 = clock64();
#endif
			if (want && lane == 0) {
				Child c;
				hy_rs_child(P, map, parent, sol, c, &consts->fm);
				s_rsh[wib] = c.ok ? hy_heuristic(P, map, wf, goal, c.pose) : 0.0;
				s_rs[wib] = c;
			}
			__syncwarp();
#ifdef PPX_SEARCH_STATS
			long long tk4 = clock64();
#endif
			if (lane == 0) {
				const double parentG = W.g[n];
				const Child* kids = s_child[wib][slotP];
				const double* kidsH = s_h[wib][slotP];
				for (int sl = 0; sl < nCs && S.status == PPX_SEARCH_RUNNING; sl++)
					if (kids[sl].ok)
						consume_child(W, S, n, parentG, kids[sl], kidsH[sl], sl);
				if (want && S.status == PPX_SEARCH_RUNNING && s_rs[wib].ok)
					consume_child(W, S, n, parentG, s_rs[wib], s_rsh[wib], nCs);
			}
			__syncwarp();
			if (slotP == 0) cache0 = -1;
			if (slotP == 1) cache1 = -1;
#ifdef PPX_SEARCH_STATS
			if (lane == 0) {
				long long tk5 = clock64();
				atomicAdd(queue + 11, (unsigned long long)(tk1 - tk0));
				atomicAdd(queue + 12, (unsigned long long)(tk2 - tk1));
				atomicAdd(queue + 13, (unsigned long long)(tk3 - tk2));
				atomicAdd(queue + 14, (unsigned long long)(tk4 - tk3));
				atomicAdd(queue + 15, (unsigned long long)(tk5 - tk4));
				atomicAdd(queue + 7, (unsigned long long)want);
			}
#endif
			status = __shfl_sync(0xffffffffu, S.status, 0);
		}

		if (lane == 0) {
			ppx_hybrid_result r;
			r.status = S.status == PPX_SEARCH_RUNNING ? -1 : S.status;
			r.n_expanded = S.nExpanded;
			r.n_nodes = S.numNodes;
			r.n_children = S.nChildren;
			r.n_rs_attempts = S.nRsAttempts;
			r.path_len = 0;
			r.cost = PPX_INF;
			r.expanded_hash = S.hash;
			if (S.status == PPX_OK && S.solution >= 0) {
				r.cost = W.g[S.solution]; // GetOptimalCost, a_star.h:300-305
				int depth = 0;
				for (int m = S.solution; m >= 0; m = W.parent[m])
					depth++;
				r.path_len = depth;
				if (paths && limits.max_path > 0) {
					ppx_hybrid_path_state* out = paths + (size_t)q * limits.max_path;
					int pos = depth - 1;
					for (int m = S.solution; m >= 0; m = W.parent[m], pos--) {
						if (pos >= limits.max_path)
							continue;
						ppx_hybrid_path_state& o = out[pos];
						o.pose[0] = W.pose[3 * m];
						o.pose[1] = W.pose[3 * m + 1];
						o.pose[2] = W.pose[3 * m + 2];
						o.length = W.length[m];
						o.last = W.last[m];
						uint32_t meta = W.meta[m];
						o.kind = W.parent[m] < 0 ? (int8_t)-1 : (int8_t)((meta >> 2) & 1u);
						o.direction = (int8_t)(meta & 3u);
						o.slot = (int8_t)((meta >> 3) & 31u);
						o._pad = 0;
					}
				}
			}
			results[q] = r;
		}
		__syncwarp();
	}
}

}

namespace {

int grow(ppx_ctx* ctx, void** ptr, size_t* have, size_t want)
{
	if (*have >= want)
		return PPX_OK;
	if (*ptr) {
		PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
		PPX_CUDA(ctx, cudaFree(*ptr));
		*ptr = nullptr;
		*have = 0;
	}
	size_t bytes = want + want / 8 + 256;
	PPX_CUDA(ctx, cudaMalloc(ptr, bytes));
	*have = bytes;
	return PPX_OK;
}

int check_handle(ppx_hybrid* h)
{
	if (!h || !h->ctx)
		return PPX_ERR_INVALID;
	if (!h->ctx->has_map)
		return ppx_fail(h->ctx, PPX_ERR_STATE, "no map uploaded (ppx_map_upload)");
	return PPX_OK;
}

// One wavefront for `goal` into h->d_wf (slot 0).
int single_wavefront(ppx_hybrid* h, const double* goal)
{
	ppx_ctx* ctx = h->ctx;
	PPX_CHECK(grow(ctx, (void**)&h->d_wf, &h->wfBytes, ctx->cells * sizeof(float)));
	double* d_goal;
	PPX_CHECK(ppx_scratch(ctx, 5, 3 * sizeof(double), (void**)&d_goal));
	PPX_CUDA(ctx, cudaMemcpyAsync(d_goal, goal, 3 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
	return ppx_wavefront_launch(ctx, 1, d_goal, h->d_wf, nullptr);
}

// Sizes the per-warp slabs and launches the search kernel over device-resident queries.
int launch_search(ppx_hybrid* h, int64_t nq, const double* d_starts, const double* d_goals, const uint64_t* d_seeds,
	const int* d_wfSlot, const float* d_wf, const ppx_hybrid_limits& limits, ppx_hybrid_result* d_results,
	ppx_hybrid_path_state* d_paths, ppx_hybrid_log_entry* d_log)
{
	ppx_ctx* ctx = h->ctx;
	WsLayout L;
	L.x0 = h->x0; L.y0 = h->y0; L.tx = h->tx; L.ty = h->ty;
	L.tableSize = h->tx * h->ty * 7;
	L.maxNodes = limits.max_nodes > 0 ? limits.max_nodes : 2 * L.tableSize;
	L.bytes = (ws_bytes(L.maxNodes, L.tableSize) + 255) & ~(size_t)255;
	const size_t budget = (size_t)32 << 30; // of slabs (180 GB of HBM; the wavefront grids of 4096 goals take 2.6 GB)
	SearchConsts consts { h->dev, ctx->map_dev(), fast_map_derive(ctx->desc) };
	if (!ctx->rs_filter)
		consts.fm.enabled = 0;
	if (!h->d_consts)
		PPX_CUDA(ctx, cudaMalloc(&h->d_consts, sizeof(SearchConsts)));
	PPX_CUDA(ctx, cudaMemcpyAsync(h->d_consts, &consts, sizeof(SearchConsts), cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemsetAsync(ctx->d_queue, 0, sizeof(unsigned long long), ctx->stream));
	int64_t resident; // queries in flight
	int ctas;
	size_t slab;
	using CtaKernel = void (*)(const SearchConsts*, int64_t, const double*, const double*, const unsigned long long*, const int*,
		const float*, WsLayout, size_t, char*, ppx_hybrid_limits, ppx_hybrid_result*, ppx_hybrid_path_state*, ppx_hybrid_log_entry*,
		unsigned long long*, const int*, unsigned long long*);
	CtaKernel ctaKernel = nullptr;
	int ctaWarps = 4;
	size_t ctaSmem = 0;
	if (ctx->search_engine == 1) {
		// one CTA per query: master warp + worker warps (ppx_hybrid_cta.cuh).  The key table lives in shared memory
		// (16-bit entries) when it fits; a batch much larger than the GPU holds runs with one worker per master and
		// twice the CTAs, a smaller batch with three workers per master (shortest critical path).
		const int nCs = 2 * h->dev.numDeltas;
		slab = (cta_ws_bytes(L.maxNodes, L.tableSize, nCs) + 255) & ~(size_t)255;
		const bool smemTable = (size_t)L.tableSize * 2 <= 28 * 1024 && L.maxNodes + 2 <= 65535
			&& !(getenv("PPX_SEARCH_SMEMTAB") && atoi(getenv("PPX_SEARCH_SMEMTAB")) == 0);
		// key table (16-bit entries, padded to 16 bytes) + ready bitmap (2 bits per node)
		ctaSmem = smemTable ? (size_t)((L.tableSize + 7) & ~7) * 2 + (size_t)((L.maxNodes + 15) / 16) * 4 : 0;
		auto pick = [&](int warps) -> CtaKernel {
			if (warps == 2) // throughput mode: small shared-memory footprint, 8 CTAs per SM
				return k_hybrid_search_cta<2, false, kCtaHeapSmemSmall>;
			return smemTable ? k_hybrid_search_cta<4, true, kCtaHeapSmemBig> : k_hybrid_search_cta<4, false, kCtaHeapSmemBig>;
		};
		auto perSmOf = [&](int warps, int* out) -> int {
			PPX_CUDA(ctx, cudaFuncSetAttribute(pick(warps), cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
			PPX_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(out, pick(warps), warps * 32, warps == 2 ? 0 : ctaSmem));
			if (getenv("PPX_SEARCH_CTAS"))
				*out = std::min(*out, atoi(getenv("PPX_SEARCH_CTAS")));
			*out = std::max(*out, 1);
			return PPX_OK;
		};
		int perSm4 = 1, perSm2 = 1;
		PPX_CHECK(perSmOf(4, &perSm4));
		PPX_CHECK(perSmOf(2, &perSm2));
		// measured crossover on C4 (search phase, 2 / 4 warps): 1024 queries 69.6 / 59.3 ms, 1536: 82.3 / 75.7, 2048: 86.4 / 86.8,
		// 3072: 109 / 118, 4096: 134 / 155
		ctaWarps = nq > 4 * (int64_t)ctx->num_sms * perSm4 ? 2 : 4;
		if (getenv("PPX_SEARCH_WARPS"))
			ctaWarps = atoi(getenv("PPX_SEARCH_WARPS")) == 2 ? 2 : 4;
		ctaKernel = pick(ctaWarps);
		if (ctaWarps == 2)
			ctaSmem = 0;
		resident = std::min<int64_t>(nq, (int64_t)ctx->num_sms * (ctaWarps == 2 ? perSm2 : perSm4));
		resident = std::max<int64_t>(1, std::min<int64_t>(resident, (int64_t)(budget / slab)));
		ctas = (int)resident;
	} else {
		// as many queries in flight as are resident (168 registers -> 8 warps per SM), at most 12 GB of slabs
		static const int wps = getenv("PPX_SEARCH_WPS") ? atoi(getenv("PPX_SEARCH_WPS")) : 8;
		slab = L.bytes;
		resident = std::min<int64_t>(nq, (int64_t)ctx->num_sms * wps);
		resident = std::max<int64_t>(1, std::min<int64_t>(resident, (int64_t)(budget / slab)));
		ctas = (int)((resident + kSearchWarps - 1) / kSearchWarps);
		resident = (int64_t)ctas * kSearchWarps;
	}
	PPX_CHECK(grow(ctx, &h->d_slabs, &h->slabBytes, (size_t)resident * slab));
	int* d_order = nullptr;
	if (nq > resident && nq < (int64_t)1 << 31) { // more queries than slots: schedule the long ones first
		float* d_keys;
		PPX_CHECK(ppx_scratch(ctx, 6, (size_t)nq * 8, (void**)&d_keys));
		d_order = (int*)(d_keys + nq);
		k_query_order<<<1, 1024, 0, ctx->stream>>>(ctx->map_dev(), nq, d_starts, d_wfSlot, d_wf, d_keys, d_order);
		PPX_LAUNCHED(ctx);
	}
	static const bool statsOn = getenv("PPX_SEARCH_STATS") != nullptr; // phase cycles of the CTA engine to stderr (tools/time_search.py)
	static unsigned long long* d_stats = nullptr;
	if (statsOn && !d_stats)
		PPX_CUDA(ctx, cudaMalloc(&d_stats, 32 * sizeof(unsigned long long)));
	if (statsOn)
		PPX_CUDA(ctx, cudaMemsetAsync(d_stats, 0, 32 * sizeof(unsigned long long), ctx->stream));
	if (ctx->search_engine == 1) {
		ctaKernel<<<ctas, ctaWarps * 32, ctaSmem, ctx->stream>>>((const SearchConsts*)h->d_consts, nq, d_starts, d_goals,
			(const unsigned long long*)d_seeds, d_wfSlot, d_wf, L, slab, (char*)h->d_slabs, limits, d_results, d_paths, d_log,
			ctx->d_queue, d_order, statsOn ? d_stats : nullptr);
		if (statsOn) {
			unsigned long long st[32];
			cudaStreamSynchronize(ctx->stream);
			cudaMemcpy(st, d_stats, sizeof(st), cudaMemcpyDeviceToHost);
			const double ne = (double)std::max<unsigned long long>(1, st[11]);
			fprintf(stderr, "cta search stats (%lld queries, %d ctas x %d warps, smem table %zu B, %llu expansions): cycles/expansion pop %.0f, "
				"rng+log %.0f, own gen %.0f, rs %.0f, wait %.0f, commit-load %.0f, commit-decide %.0f, commit-write %.0f | kids ready %.1f%%, claimed %.1f%%, "
				"own %.1f%% | rs shots ready %llu, claimed %llu, own %llu | workers: %llu passes, %.2f nodes/pass, %.0f cycles/pass, %llu rs "
				"tasks, %llu idle, nodes generated / expanded %.2f\n",
				(long long)nq, ctas, ctaWarps, ctaSmem, st[11], st[0] / ne, st[1] / ne, st[2] / ne, st[3] / ne, st[4] / ne, st[5] / ne,
				st[7] / ne, st[6] / ne, 100.0 * st[8] / ne, 100.0 * st[9] / ne, 100.0 * st[10] / ne, st[12], st[13], st[14], st[16],
				st[16] ? (double)st[17] / st[16] : 0.0, st[16] ? (double)st[18] / st[16] : 0.0, st[20], st[19], (st[17] + st[10]) / ne);
		}
	} else
		k_hybrid_search<<<ctas, kSearchWarps * 32, 0, ctx->stream>>>((const SearchConsts*)h->d_consts, nq, d_starts, d_goals,
			(const unsigned long long*)d_seeds, d_wfSlot, d_wf, L, (char*)h->d_slabs, limits, d_results, d_paths, d_log, ctx->d_queue,
			d_order);
	PPX_LAUNCHED(ctx);
#ifdef PPX_SEARCH_STATS
	{
		unsigned long long st[9];
		cudaStreamSynchronize(ctx->stream);
		cudaMemcpy(st, ctx->d_queue + 7, sizeof(st), cudaMemcpyDeviceToHost);
		double ne = (double)(st[1] + st[2]);
		fprintf(stderr, "search stats: %llu passes (misses), %llu hits, %llu generated ahead, %llu rs shots; cycles/expansion: pop+rng %.0f, rs solve %.0f, gen pass %.0f, rs child %.0f, commit %.0f\n",
			st[1], st[2], st[3], st[0], st[4] / ne, st[5] / ne, st[6] / ne, st[7] / ne, st[8] / ne);
		cudaMemset(ctx->d_queue + 7, 0, sizeof(st));
	}
#endif
	return PPX_OK;
}

}

extern "C" {

int ppx_hybrid_create(ppx_ctx* ctx, const ppx_hybrid_params* params, const ppx_validator_params* vp, ppx_hybrid** out)
{
	PPX_GUARD(ctx);
	if (!ctx || !out)
		return PPX_ERR_INVALID;
	*out = nullptr;
	PPX_REQUIRE(ctx, params && vp, "bad arguments");
	if (!ctx->has_map)
		return ppx_fail(ctx, PPX_ERR_STATE, "no map uploaded (ppx_map_upload)");
	PPX_REQUIRE(ctx, params->spatial_resolution > 0 && params->angular_resolution > 0 && params->min_turning_radius > 0,
		"resolutions and turning radius must be positive");
	auto* h = new ppx_hybrid();
	h->ctx = ctx;
	h->params = *params;
	h->vp = *vp;
	if (hybrid_derive(*params, *vp, ctx->desc, h->dev) != PPX_OK) {
		delete h;
		return ppx_fail(ctx, PPX_ERR_INVALID, "num_generated_motion too large (max %d steering angles)", kMaxDeltas);
	}
	hybrid_key_table(*params, ctx->desc, h->x0, h->y0, h->tx, h->ty);
	const size_t total = (size_t)h->dev.nhX * h->dev.nhY * h->dev.nhA;
	cudaError_t e = cudaMalloc(&h->d_table, total * sizeof(double));
	if (e != cudaSuccess) {
		delete h;
		return ppx_fail(ctx, PPX_ERR_CUDA, "cudaMalloc(heuristic table): %s", cudaGetErrorString(e));
	}
	h->dev.nh = h->d_table;
	for (auto& e : h->ev)
		cudaEventCreate(&e);
	int blocks = (int)std::min<size_t>((total + 127) / 128, (size_t)ctx->num_sms * 16);
	k_nonholo_build<<<blocks, 128, 0, ctx->stream>>>(h->dev, h->d_table);
	PPX_LAUNCHED(ctx);
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	*out = h;
	return PPX_OK;
}

int ppx_hybrid_destroy(ppx_hybrid* h)
{
	PPX_GUARD(h ? h->ctx : nullptr);
	if (!h)
		return PPX_ERR_INVALID;
	cudaStreamSynchronize(h->ctx->stream);
	cudaFree(h->d_table);
	cudaFree(h->d_slabs);
	cudaFree(h->d_wf);
	cudaFree(h->d_io);
	cudaFree(h->d_consts);
	for (auto& e : h->ev)
		if (e)
			cudaEventDestroy(e);
	delete h;
	return PPX_OK;
}

int ppx_hybrid_table(ppx_hybrid* h, int32_t* dims, double* offsets, double* values)
{
	PPX_GUARD(h ? h->ctx : nullptr);
	if (!h)
		return PPX_ERR_INVALID;
	if (dims) {
		dims[0] = h->dev.nhX;
		dims[1] = h->dev.nhY;
		dims[2] = h->dev.nhA;
	}
	if (offsets) {
		offsets[0] = h->dev.nhOffX;
		offsets[1] = h->dev.nhOffY;
	}
	if (values) {
		const size_t total = (size_t)h->dev.nhX * h->dev.nhY * h->dev.nhA;
		PPX_CUDA(h->ctx, cudaMemcpyAsync(values, h->d_table, total * sizeof(double), cudaMemcpyDeviceToHost, h->ctx->stream));
		PPX_CUDA(h->ctx, cudaStreamSynchronize(h->ctx->stream));
	}
	return PPX_OK;
}

int ppx_hybrid_heuristic(ppx_hybrid* h, const double* goal, int64_t n, const double* poses, double* combined,
	double* nonholo, double* obstacles)
{
	PPX_GUARD(h ? h->ctx : nullptr);
	PPX_CHECK(check_handle(h));
	ppx_ctx* ctx = h->ctx;
	PPX_REQUIRE(ctx, goal && n >= 0 && (n == 0 || poses), "bad arguments");
	if (n == 0)
		return PPX_OK;
	PPX_CHECK(single_wavefront(h, goal));
	double* d_poses;
	double* d_out;
	PPX_CHECK(ppx_scratch(ctx, 0, n * 3 * sizeof(double), (void**)&d_poses));
	PPX_CHECK(ppx_scratch(ctx, 1, n * 3 * sizeof(double), (void**)&d_out));
	PPX_CUDA(ctx, cudaMemcpyAsync(d_poses, poses, n * 3 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
	// Pose2d goal as HybridAStar::SetGoalState stores it (ctor wrap is the caller's, like everywhere in this ABI)
	Pose g = { goal[0], goal[1], goal[2] };
	int blocks = (int)std::min<int64_t>((n + 127) / 128, (int64_t)ctx->num_sms * 16);
	k_hybrid_heuristic<<<blocks, 128, 0, ctx->stream>>>(h->dev, ctx->map_dev(), h->d_wf, g, n, d_poses, d_out, d_out + n, d_out + 2 * n);
	PPX_LAUNCHED(ctx);
	if (combined)
		PPX_CUDA(ctx, cudaMemcpyAsync(combined, d_out, n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
	if (nonholo)
		PPX_CUDA(ctx, cudaMemcpyAsync(nonholo, d_out + n, n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
	if (obstacles)
		PPX_CUDA(ctx, cudaMemcpyAsync(obstacles, d_out + 2 * n, n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

int ppx_hybrid_expand(ppx_hybrid* h, const double* goal, int64_t n, const double* parents, const uint64_t* seeds,
	int32_t* n_children, ppx_hybrid_child* children, double* parent_heuristic)
{
	PPX_GUARD(h ? h->ctx : nullptr);
	PPX_CHECK(check_handle(h));
	ppx_ctx* ctx = h->ctx;
	PPX_REQUIRE(ctx, goal && n >= 0 && (n == 0 || (parents && seeds && n_children && children)), "bad arguments");
	if (n == 0)
		return PPX_OK;
	PPX_CHECK(single_wavefront(h, goal));
	int blocks = (int)std::min<int64_t>((n + 3) / 4, (int64_t)ctx->num_sms * 8);
	const size_t nwarps = (size_t)blocks * 4;
	size_t off = 0;
	auto carve = [&](size_t bytes) {
		size_t o = off;
		off += (bytes + 255) & ~(size_t)255;
		return o;
	};
	size_t oParents = carve(n * 3 * sizeof(double)), oSeeds = carve(n * sizeof(uint64_t)), oCount = carve(n * sizeof(int32_t));
	size_t oChildren = carve(n * PPX_MAX_CHILDREN * sizeof(ppx_hybrid_child)), oHeur = carve(n * sizeof(double));
	size_t oMt = carve(nwarps * 312 * sizeof(unsigned long long));
	char* base;
	PPX_CHECK(ppx_scratch(ctx, 0, off, (void**)&base));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oParents, parents, n * 3 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oSeeds, seeds, n * sizeof(uint64_t), cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemsetAsync(base + oChildren, 0, n * PPX_MAX_CHILDREN * sizeof(ppx_hybrid_child), ctx->stream));
	Pose g = { goal[0], goal[1], goal[2] };
	k_hybrid_expand<<<blocks, 128, 0, ctx->stream>>>(h->dev, ctx->map_dev(), h->d_wf, g, n, (const double*)(base + oParents),
		(const unsigned long long*)(base + oSeeds), (int32_t*)(base + oCount), (ppx_hybrid_child*)(base + oChildren),
		(double*)(base + oHeur), (unsigned long long*)(base + oMt));
	PPX_LAUNCHED(ctx);
	PPX_CUDA(ctx, cudaMemcpyAsync(n_children, base + oCount, n * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(children, base + oChildren, n * PPX_MAX_CHILDREN * sizeof(ppx_hybrid_child), cudaMemcpyDeviceToHost, ctx->stream));
	if (parent_heuristic)
		PPX_CUDA(ctx, cudaMemcpyAsync(parent_heuristic, base + oHeur, n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

int ppx_hybrid_resample(ppx_hybrid* h, int64_t nq, const double* goals, const ppx_hybrid_result* results,
	const ppx_hybrid_path_state* paths, int32_t max_path, float path_interpolation, int32_t max_points, double* poses,
	uint8_t* cusp, int32_t* count, double* lengths)
{
	PPX_GUARD(h ? h->ctx : nullptr);
	PPX_CHECK(check_handle(h));
	ppx_ctx* ctx = h->ctx;
	PPX_REQUIRE(ctx, nq >= 0 && max_path >= 2 && max_points > 0 && path_interpolation > 0.0f, "bad arguments");
	PPX_REQUIRE(ctx, nq == 0 || (goals && results && paths && poses && cusp && count), "bad arguments");
	if (nq == 0)
		return PPX_OK;
	size_t off = 0;
	auto carve = [&](size_t bytes) {
		size_t o = off;
		off += (bytes + 255) & ~(size_t)255;
		return o;
	};
	const size_t bGoals = (size_t)nq * 24, bRes = (size_t)nq * sizeof(ppx_hybrid_result);
	const size_t bPaths = (size_t)nq * max_path * sizeof(ppx_hybrid_path_state), bFinal = (size_t)nq * max_path * 8;
	const size_t bPoses = (size_t)nq * max_points * 24, bCusp = (size_t)nq * max_points, bCount = (size_t)nq * 4, bLen = (size_t)nq * 8;
	size_t oGoals = carve(bGoals), oRes = carve(bRes), oPaths = carve(bPaths), oFinal = carve(bFinal), oPoses = carve(bPoses);
	size_t oCusp = carve(bCusp), oCount = carve(bCount), oLen = carve(bLen);
	char* base;
	PPX_CHECK(ppx_scratch(ctx, 5, off, (void**)&base));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oGoals, goals, bGoals, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oRes, results, bRes, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oPaths, paths, bPaths, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemsetAsync(base + oCusp, 0, bCusp, ctx->stream));
	PPX_CUDA(ctx, cudaMemsetAsync(base + oPoses, 0, bPoses, ctx->stream)); // entries past count[q] read as zero
	SearchConsts consts { h->dev, ctx->map_dev(), fast_map_derive(ctx->desc) };
	if (!h->d_consts)
		PPX_CUDA(ctx, cudaMalloc(&h->d_consts, sizeof(SearchConsts)));
	PPX_CUDA(ctx, cudaMemcpyAsync(h->d_consts, &consts, sizeof(SearchConsts), cudaMemcpyHostToDevice, ctx->stream));
	int ctas = (int)std::min<int64_t>((nq + kResampleWarps - 1) / kResampleWarps, (int64_t)ctx->num_sms * 8);
	k_hybrid_resample<<<ctas, kResampleWarps * 32, 0, ctx->stream>>>((const SearchConsts*)h->d_consts, nq, (const double*)(base + oGoals),
		(const ppx_hybrid_result*)(base + oRes), (const ppx_hybrid_path_state*)(base + oPaths), max_path, path_interpolation,
		max_points, (double*)(base + oFinal), (double*)(base + oPoses), (uint8_t*)(base + oCusp), (int32_t*)(base + oCount),
		(double*)(base + oLen));
	PPX_LAUNCHED(ctx);
	PPX_CUDA(ctx, cudaMemcpyAsync(poses, base + oPoses, bPoses, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(cusp, base + oCusp, bCusp, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(count, base + oCount, bCount, cudaMemcpyDeviceToHost, ctx->stream));
	if (lengths)
		PPX_CUDA(ctx, cudaMemcpyAsync(lengths, base + oLen, bLen, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

int ppx_hybrid_search_dev(ppx_hybrid* h, int64_t nq, const double* d_starts, const double* d_goals,
	const uint64_t* d_seeds, const ppx_hybrid_limits* limits, ppx_hybrid_result* d_results,
	ppx_hybrid_path_state* d_paths, ppx_hybrid_log_entry* d_log)
{
	PPX_GUARD(h ? h->ctx : nullptr);
	PPX_CHECK(check_handle(h));
	ppx_ctx* ctx = h->ctx;
	PPX_REQUIRE(ctx, nq >= 0 && limits && (nq == 0 || (d_starts && d_goals && d_results)), "bad arguments");
	if (nq == 0)
		return PPX_OK;
	// chunks of queries whose wavefront grids fit the budget (8 GB)
	const size_t grid = ctx->cells * sizeof(float);
	const int64_t chunk = std::max<int64_t>(1, std::min<int64_t>(nq, (int64_t)(((size_t)8 << 30) / grid)));
	PPX_CHECK(grow(ctx, (void**)&h->d_wf, &h->wfBytes, (size_t)chunk * grid));
	h->timed = chunk >= nq;
	for (int64_t q0 = 0; q0 < nq; q0 += chunk) {
		const int64_t m = std::min(chunk, nq - q0);
		if (h->timed)
			cudaEventRecord(h->ev[0], ctx->stream);
		PPX_CHECK(ppx_wavefront_launch(ctx, m, d_goals + 3 * q0, h->d_wf, nullptr));
		if (h->timed)
			cudaEventRecord(h->ev[1], ctx->stream);
		PPX_CHECK(launch_search(h, m, d_starts + 3 * q0, d_goals + 3 * q0, d_seeds ? d_seeds + q0 : nullptr, nullptr, h->d_wf, *limits,
			d_results + q0, d_paths ? d_paths + (size_t)q0 * limits->max_path : nullptr,
			d_log ? d_log + (size_t)q0 * limits->max_log : nullptr));
		if (h->timed)
			cudaEventRecord(h->ev[2], ctx->stream);
	}
	return PPX_OK;
}

int ppx_hybrid_timing(ppx_hybrid* h, double* wavefront_ms, double* search_ms)
{
	PPX_GUARD(h ? h->ctx : nullptr);
	if (!h || !h->ctx)
		return PPX_ERR_INVALID;
	if (!h->timed)
		return ppx_fail(h->ctx, PPX_ERR_STATE, "no timed search (the last call was split into chunks, or there was none)");
	PPX_CUDA(h->ctx, cudaEventSynchronize(h->ev[2]));
	float a = 0, b = 0;
	PPX_CUDA(h->ctx, cudaEventElapsedTime(&a, h->ev[0], h->ev[1]));
	PPX_CUDA(h->ctx, cudaEventElapsedTime(&b, h->ev[1], h->ev[2]));
	if (wavefront_ms)
		*wavefront_ms = a;
	if (search_ms)
		*search_ms = b;
	return PPX_OK;
}

int ppx_hybrid_search(ppx_hybrid* h, int64_t nq, const double* starts, const double* goals, const uint64_t* seeds,
	const ppx_hybrid_limits* limits, ppx_hybrid_result* results, ppx_hybrid_path_state* paths, ppx_hybrid_log_entry* log)
{
	PPX_GUARD(h ? h->ctx : nullptr);
	PPX_CHECK(check_handle(h));
	ppx_ctx* ctx = h->ctx;
	PPX_REQUIRE(ctx, nq >= 0 && limits && (nq == 0 || (starts && goals && results)), "bad arguments");
	PPX_REQUIRE(ctx, !paths || limits->max_path > 0, "paths requested with max_path == 0");
	PPX_REQUIRE(ctx, !log || limits->max_log > 0, "log requested with max_log == 0");
	if (nq == 0)
		return PPX_OK;
	const ppx_map_desc& d = ctx->desc;
	// ObstaclesHeuristic::Update depends on the goal CELL only: one wavefront per distinct cell
	std::unordered_map<long long, int> slotOf;
	std::vector<int> wfSlot(nq);
	std::vector<double> uniqueGoals;
	for (int64_t q = 0; q < nq; q++) {
		double gx = (goals[3 * q] - d.world_origin[0]) / (double)d.resolution;
		double gy = (goals[3 * q + 1] - d.world_origin[1]) / (double)d.resolution;
		bool in = gx > -2147483649.0 && gx < 2147483648.0 && gy > -2147483649.0 && gy < 2147483648.0;
		int r = in ? (int)gx : -1, c = in ? (int)gy : -1;
		if (r < 0 || r >= d.rows || c < 0 || c >= d.cols)
			r = c = -1; // every goal outside the map shares one (all-infinite) grid
		long long key = (long long)r * 0x100000000LL + (long long)(unsigned int)c;
		auto it = slotOf.find(key);
		if (it == slotOf.end()) {
			it = slotOf.emplace(key, (int)(uniqueGoals.size() / 3)).first;
			uniqueGoals.insert(uniqueGoals.end(), goals + 3 * q, goals + 3 * q + 3);
		}
		wfSlot[q] = it->second;
	}
	const int64_t nGoals = (int64_t)uniqueGoals.size() / 3;
	const size_t grid = ctx->cells * sizeof(float);
	PPX_REQUIRE(ctx, (size_t)nGoals * grid <= ((size_t)64 << 30), "too many distinct goals for one call (wavefront grids > 64 GB)");
	PPX_CHECK(grow(ctx, (void**)&h->d_wf, &h->wfBytes, (size_t)nGoals * grid));
	size_t off = 0;
	auto carve = [&](size_t bytes) {
		size_t o = off;
		off += (bytes + 255) & ~(size_t)255;
		return o;
	};
	size_t oStarts = carve(nq * 24), oGoals = carve(nq * 24), oSeeds = carve(nq * 8), oSlot = carve(nq * 4);
	size_t oUGoals = carve(nGoals * 24), oRes = carve(nq * sizeof(ppx_hybrid_result));
	size_t oPaths = carve(paths ? (size_t)nq * limits->max_path * sizeof(ppx_hybrid_path_state) : 0);
	size_t oLog = carve(log ? (size_t)nq * limits->max_log * sizeof(ppx_hybrid_log_entry) : 0);
	PPX_CHECK(grow(ctx, &h->d_io, &h->ioBytes, off));
	char* base = (char*)h->d_io;
	std::vector<uint64_t> zeroSeeds;
	if (!seeds) {
		zeroSeeds.assign(nq, 0);
		seeds = zeroSeeds.data();
	}
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oStarts, starts, nq * 24, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oGoals, goals, nq * 24, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oSeeds, seeds, nq * 8, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oSlot, wfSlot.data(), nq * 4, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oUGoals, uniqueGoals.data(), nGoals * 24, cudaMemcpyHostToDevice, ctx->stream));
	h->timed = true;
	cudaEventRecord(h->ev[0], ctx->stream);
	PPX_CHECK(ppx_wavefront_launch(ctx, nGoals, (const double*)(base + oUGoals), h->d_wf, nullptr));
	cudaEventRecord(h->ev[1], ctx->stream);
	ppx_hybrid_path_state* d_paths = paths ? (ppx_hybrid_path_state*)(base + oPaths) : nullptr;
	ppx_hybrid_log_entry* d_log = log ? (ppx_hybrid_log_entry*)(base + oLog) : nullptr;
	if (d_paths)
		PPX_CUDA(ctx, cudaMemsetAsync(d_paths, 0, (size_t)nq * limits->max_path * sizeof(ppx_hybrid_path_state), ctx->stream));
	if (d_log)
		PPX_CUDA(ctx, cudaMemsetAsync(d_log, 0, (size_t)nq * limits->max_log * sizeof(ppx_hybrid_log_entry), ctx->stream));
	PPX_CHECK(launch_search(h, nq, (const double*)(base + oStarts), (const double*)(base + oGoals), (const uint64_t*)(base + oSeeds),
		(const int*)(base + oSlot), h->d_wf, *limits, (ppx_hybrid_result*)(base + oRes), d_paths, d_log));
	cudaEventRecord(h->ev[2], ctx->stream);
	PPX_CUDA(ctx, cudaMemcpyAsync(results, base + oRes, nq * sizeof(ppx_hybrid_result), cudaMemcpyDeviceToHost, ctx->stream));
	if (paths)
		PPX_CUDA(ctx, cudaMemcpyAsync(paths, d_paths, (size_t)nq * limits->max_path * sizeof(ppx_hybrid_path_state), cudaMemcpyDeviceToHost, ctx->stream));
	if (log)
		PPX_CUDA(ctx, cudaMemcpyAsync(log, d_log, (size_t)nq * limits->max_log * sizeof(ppx_hybrid_log_entry), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

} // extern "C"
