// libppx.so: RRT and RRT* (A11) with the brute-force nearest-neighbour scan (A10) in the loop,
// one tree per thread-block CLUSTER, batched over trees (algo/rrt.h:55-95, algo/rrt_star.h:53-112, utils/tree.h).
//
// A tree is a chain of dependent iterations, each dominated by one or two nearest-neighbour queries over all current
// nodes.  The scalar control flow -- RNG draws, steering, duplicate-state rule, goal test -- runs on thread 0 of the
// cluster's CTA 0 in the reference's statement order; the RRT* parent candidates (k <= 16 connections + collision
// checks) are evaluated on that CTA's threads.  The scans are the part that grows with the tree, and they are spread
// over the cluster (1, 2, 4, 8 or 16 CTAs = SMs, by how many trees there are to fill the GPU with):
//   * node i belongs to CTA (i / 256) % C; every CTA keeps the coordinates of its nodes in its own shared memory
//     (SoA, up to 200 KB: 136 k nodes for C = 16), so a scan never leaves the SM -- one CTA scanning 100 k nodes out of
//     L2 was bound by the load latency of a single SM (41 us per scan, profiles/r2l_*);
//   * CTA 0 publishes a command (query, k, tree size, the node appended since the last command) in its shared memory;
//     after a cluster barrier every CTA reads it through DSMEM, appends the node if it owns it, scans its slice (exact
//     FP64 squared distances, ties to the lower insertion index like the oracle's flann stand-in), merges the
//     per-thread candidate lists by warp shuffles and publishes its k best; after a second cluster barrier CTA 0
//     gathers the C lists through DSMEM and merges them.  Two cluster barriers (~0.2 us each) per query.
// Trees too large for the slices are scanned from global memory with the same partition.
// Independent trees (different seeds / queries) fill the other SMs.
#include <algorithm>
#include <cooperative_groups.h>
#include "ppx_ctx.h"
#include "ppx_fastlane.cuh"

using namespace ppx;
namespace cg = cooperative_groups;

namespace {

constexpr int kRrtThreads = 256;
constexpr int kRrtWarps = kRrtThreads / 32;
constexpr int kRrtMaxK = 16;
constexpr int kRrtMaxCluster = 16;
constexpr int kNoNode = 0x7fffffff;

struct RrtDev {
	int dim, connection, star, validator;
	unsigned maxIteration, maxNodes;
	double maxConnectionDistance, goalBias, rmin;
	double lb[3], ub[3];
	float minSafeRadius, minStep;
};

struct RrtWs { // per tree
	double* coords; // [cap][3]
	int* parent; // [cap]
	double* cost; // [cap] NodeInfo::costFromGoal (RRT*)
	int* hash; // [hashCap] node index + 1, open addressing on the state's coordinate bits
	unsigned long long* mt; // 312
	int cap, hashCap;
};

// std::mt19937_64 + Random<double>::SampleUniform (utils/random.h:12-27), see ppx_hybrid.cuh Mt64
struct RrtRng {
	unsigned long long* mt;
	int idx;
	__device__ void seed(unsigned long long s)
	{
		mt[0] = s;
		for (int i = 1; i < 312; i++)
			mt[i] = 6364136223846793005ULL * (mt[i - 1] ^ (mt[i - 1] >> 62)) + (unsigned long long)i;
		idx = 312;
	}
	__device__ unsigned long long next()
	{
		if (idx >= 312) {
			const unsigned long long UM = 0xFFFFFFFF80000000ULL, LM = 0x7FFFFFFFULL, A = 0xB5026F5AA96619E9ULL;
			for (int i = 0; i < 312; i++) {
				unsigned long long x = (mt[i] & UM) | (mt[(i + 1) % 312] & LM);
				mt[i] = mt[(i + 156) % 312] ^ (x >> 1) ^ ((x & 1ULL) ? A : 0ULL);
			}
			idx = 0;
		}
		unsigned long long x = mt[idx++];
		x ^= (x >> 29) & 0x5555555555555555ULL;
		x ^= (x << 17) & 0x71D67FFFEDA60000ULL;
		x ^= (x << 37) & 0xFFF7EEE000000000ULL;
		x ^= (x >> 43);
		return x;
	}
	// s_uniformDistribution(*s_engine): generate_canonical * (nextafter(1) - 0) + 0
	__device__ double dist()
	{
		double canon = (double)next() / 18446744073709551616.0;
		if (canon >= 1.0)
			canon = 0.99999999999999988897769753748;
		return canon * 1.0000000000000002220446049250313 + 0.0;
	}
	// SampleUniform(lb, ub) = lb + (ub - lb) * dist
	__device__ double uniform(double lb, double ub)
	{
		double range = ub - lb;
		return lb + range * dist();
	}
};

__device__ __forceinline__ bool nn_closer(double da, int ia, double db, int ib) { return da < db || (da == db && ia < ib); }

// flann::L2_Simple over the first `dim` raw coordinates
__device__ __forceinline__ double state_dist2(const double* a, const double* b, int dim)
{
	double r = 0.0;
	for (int i = 0; i < dim; i++) {
		double diff = a[i] - b[i];
		r += diff * diff;
	}
	return r;
}

struct NnEntry { // a neighbour candidate: squared distance and node index (kNoNode: none)
	double d;
	int idx;
	int pad;
};

enum { OP_SCAN = 0, OP_EXIT = 1, OP_SELECT = 2, OP_NEAREST = 3 };
constexpr int kSpec = 8; // samples of a speculation batch (one per warp of CTA 0)
constexpr int kCandCap = 64; // candidates per CTA of a threshold selection (overflow: the list scan answers the query)

struct alignas(8) RrtCmd { // CTA 0's command to its cluster, read by the other CTAs through DSMEM after the cluster barrier
	int op, k, n; // n: tree size the scan covers
	int nq; // OP_NEAREST: number of queries (nearest node of each); else 1
	int nNew, pad; // nodes appended since the last command; their owners store them in their slices
	int newIdx[kSpec];
	double q[kSpec][3], newState[kSpec][3];
};

struct RrtSlice { // the cluster's partition of the node coordinates: node i -> CTA (i >> 8) % C, local slot ((i >> 8) / C) << 8 | i & 255
	double *x, *y, *th; // shared memory of this CTA (SoA) or null: scan W.coords in global memory
	int rank, C;
};

struct RrtXch { // exchange areas in shared memory
	NnEntry warpList[kRrtWarps][kRrtMaxK];
	NnEntry ctaList[kRrtMaxK]; // read by CTA 0 through DSMEM
	NnEntry gather[kRrtMaxCluster][kRrtMaxK]; // CTA 0 only
	// threshold selection (cta_select_*): per-warp minima and the candidates under the threshold, read through DSMEM
	NnEntry warpMin[kRrtWarps];
	NnEntry allMin[kRrtMaxCluster * kRrtWarps];
	NnEntry cand[kCandCap];
	NnEntry thr;
	int candCount;
	int counts[kRrtMaxCluster];
};

// Cluster barrier: arrive.release + wait.acquire (what cg::cluster_group::sync() is; SASS MEMBAR.ALL.GPU + UCGABAR_ARV +
// UCGABAR_WAIT + CCTL.IVALL).  The release orders this CTA's shared-memory writes (commands, slices, candidate lists) and
// global writes (the tree, when the slices are scanned from global memory) before the other CTAs' reads behind the barrier.
// A relaxed arrive behind a CTA-scope fence passed every test and was 1.5 % faster on the 100k-node trees; it is not what
// the PTX memory model asks for when another CTA reads the data, so it was not kept.
__device__ __forceinline__ void cluster_barrier(bool /*globalData*/)
{
	asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// Warp-wide minimum of (d, idx) under nn_closer; every lane gets it.
__device__ __forceinline__ void warp_nn_min(double& d, int& idx)
{
	#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		const double od = __shfl_xor_sync(0xffffffffu, d, o);
		const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
		if (nn_closer(od, oi, d, idx)) {
			d = od;
			idx = oi;
		}
	}
}

// Merge of up to 32 ascending lists (one per lane, `mine` = this lane's list or null) into out[0..k): k rounds of a warp
// minimum over the list heads; node indices are unique, so exactly the owner of the minimum advances.
__device__ __forceinline__ void warp_merge_lists(const NnEntry* mine, int k, NnEntry* out)
{
	int head = 0;
	for (int round = 0; round < k; round++) {
		const bool have = mine && head < k;
		const double hd = have ? mine[head].d : PPX_INF;
		const int hi = have ? mine[head].idx : kNoNode;
		double d = hd;
		int idx = hi;
		warp_nn_min(d, idx);
		if (idx == hi && hi != kNoNode)
			head++;
		if ((threadIdx.x & 31) == 0) {
			out[round].d = d;
			out[round].idx = idx;
		}
	}
}

// This CTA's part of a neighbour query: scan the nodes it owns among [0, n), k best (ascending) into X.ctaList.
// K >= k is the length of the per-thread candidate list: a compile-time constant, so the list is registers with static
// indexing (sorted insertion = one pass of compare-and-swap with a carried element; pop = shift).  With a run-time
// length the list sat in local memory and one insertion cost ~1 us of dependent L1 round trips -- the part of an
// iteration that grew with the tree.  Ends without a barrier: the caller's cluster barrier publishes the list.
template <int K>
__device__ void cta_scan_slice(const RrtWs& W, const RrtSlice& S, int n, int dim, const double* query, int k, RrtXch& X)
{
	double bd[K];
	int bi[K];
	#pragma unroll
	for (int i = 0; i < K; i++) {
		bd[i] = PPX_INF;
		bi[i] = kNoNode;
	}
	const double q0 = query[0], q1 = query[1], q2 = query[2];
	auto offer = [&](double d, int j) {
		if (!nn_closer(d, j, bd[K - 1], bi[K - 1]))
			return;
		#pragma unroll
		for (int i = 0; i < K; i++) {
			if (nn_closer(d, j, bd[i], bi[i])) {
				const double td = bd[i];
				const int ti = bi[i];
				bd[i] = d;
				bi[i] = j;
				d = td;
				j = ti;
			}
		}
	};
	// flann::L2_Simple over the first `dim` raw coordinates, in its accumulation order
	auto dist2 = [&](double x, double y, double th) {
		const double dx = q0 - x, dy = q1 - y;
		double r = 0.0;
		r += dx * dx;
		r += dy * dy;
		if (dim == 3) {
			const double dt = q2 - th;
			r += dt * dt;
		}
		return r;
	};
	const int stride = S.C << 8; // node index distance between two blocks of this CTA
	if (S.x) {
		int local = threadIdx.x;
		for (int j = (S.rank << 8) + (int)threadIdx.x; j < n; j += stride, local += 256)
			offer(dist2(S.x[local], S.y[local], S.th[local]), j);
	} else {
		// four nodes per thread and turn: their coordinate loads are independent and in flight together
		constexpr int kU = 4;
		for (int j0 = (S.rank << 8) + (int)threadIdx.x; j0 < n; j0 += kU * stride) {
			double dd[kU];
			#pragma unroll
			for (int u = 0; u < kU; u++) {
				const int j = j0 + u * stride;
				const double* c = W.coords + 3 * (size_t)j;
				dd[u] = j < n ? dist2(c[0], c[1], c[2]) : PPX_INF;
			}
			#pragma unroll
			for (int u = 0; u < kU; u++) {
				const int j = j0 + u * stride;
				if (j < n)
					offer(dd[u], j);
			}
		}
	}
	// the 32 lists of a warp -> one list per warp -> the CTA's list
	{
		NnEntry* out = X.warpList[threadIdx.x >> 5];
		for (int round = 0; round < k; round++) {
			double d = bd[0];
			int idx = bi[0];
			warp_nn_min(d, idx);
			if (idx == bi[0] && idx != kNoNode) { // node indices are unique: only the owner pops
				#pragma unroll
				for (int i = 0; i + 1 < K; i++) {
					bd[i] = bd[i + 1];
					bi[i] = bi[i + 1];
				}
				bd[K - 1] = PPX_INF;
				bi[K - 1] = kNoNode;
			}
			if ((threadIdx.x & 31) == 0) {
				out[round].d = d;
				out[round].idx = idx;
			}
		}
	}
	__syncthreads();
	if (threadIdx.x < 32)
		warp_merge_lists(threadIdx.x < kRrtWarps ? X.warpList[threadIdx.x] : nullptr, k, X.ctaList);
}

// Threshold selection, the fast form of a k-nearest query for clusters of >= 4 CTAs (>= 32 warps >= 2 k):
//   pass 1  every warp publishes the nearest node of its threads' nodes (8 * C distinct nodes);
//   ------  cluster barrier; every CTA reads the 8 * C minima and takes their k-th smallest as threshold T: the k nearest
//           nodes overall are certainly not farther than T;
//   pass 2  every thread appends its nodes that are not farther than T to its CTA's candidate list (typically k ... 2 k
//           candidates in the whole cluster);
//   ------  cluster barrier; CTA 0 gathers the candidates and ranks them by counting.
// The list scan above did k rounds of a 5-step shuffle minimum on each of three levels (thread lists -> warp -> CTA ->
// cluster): 24 k - 30 k cycles per query against 4 k for a nearest-node query (PPX_RRT_STATS); it remains the answer when
// a candidate list overflows.
template <typename F>
__device__ __forceinline__ void for_own_nodes(const RrtWs& W, const RrtSlice& S, int n, F&& f)
{
	const int stride = S.C << 8;
	if (S.x) {
		int local = threadIdx.x;
		for (int j = (S.rank << 8) + (int)threadIdx.x; j < n; j += stride, local += 256)
			f(S.x[local], S.y[local], S.th[local], j);
	} else {
		for (int j = (S.rank << 8) + (int)threadIdx.x; j < n; j += stride) {
			const double* c = W.coords + 3 * (size_t)j;
			f(c[0], c[1], c[2], j);
		}
	}
}

__device__ __forceinline__ double query_dist2(const double* q, int dim, double x, double y, double th)
{
	const double dx = q[0] - x, dy = q[1] - y;
	double r = 0.0;
	r += dx * dx;
	r += dy * dy;
	if (dim == 3) {
		const double dt = q[2] - th;
		r += dt * dt;
	}
	return r;
}

__device__ void cta_select_pass1(const RrtWs& W, const RrtSlice& S, int n, int dim, const double* q, RrtXch& X)
{
	double d = PPX_INF;
	int idx = kNoNode;
	for_own_nodes(W, S, n, [&](double x, double y, double th, int j) {
		const double dj = query_dist2(q, dim, x, y, th);
		if (nn_closer(dj, j, d, idx)) {
			d = dj;
			idx = j;
		}
	});
	warp_nn_min(d, idx);
	if ((threadIdx.x & 31) == 0) {
		X.warpMin[threadIdx.x >> 5].d = d;
		X.warpMin[threadIdx.x >> 5].idx = idx;
	}
	if (threadIdx.x == 0)
		X.candCount = 0;
}

// after the cluster barrier that follows pass 1
__device__ void cta_select_pass2(cg::cluster_group& cluster, const RrtWs& W, const RrtSlice& S, int n, int dim, const double* q, int k,
	RrtXch& X)
{
	const int m = S.C * kRrtWarps;
	if ((int)threadIdx.x < m) {
		const NnEntry* remote = cluster.map_shared_rank(X.warpMin, (int)threadIdx.x / kRrtWarps);
		X.allMin[threadIdx.x] = remote[threadIdx.x % kRrtWarps];
	}
	if (threadIdx.x == 0) { // fewer than k nodes among the minima: everything is a candidate
		X.thr.d = PPX_INF;
		X.thr.idx = kNoNode;
	}
	__syncthreads();
	if ((int)threadIdx.x < m) {
		const NnEntry mine = X.allMin[threadIdx.x];
		int rank = 0;
		for (int u = 0; u < m; u++)
			rank += nn_closer(X.allMin[u].d, X.allMin[u].idx, mine.d, mine.idx) ? 1 : 0;
		if (rank == k - 1 && mine.idx != kNoNode)
			X.thr = mine;
	}
	__syncthreads();
	const double td = X.thr.d;
	const int ti = X.thr.idx;
	for_own_nodes(W, S, n, [&](double x, double y, double th, int j) {
		const double dj = query_dist2(q, dim, x, y, th);
		if (!nn_closer(td, ti, dj, j)) { // (dj, j) <= T
			const int pos = atomicAdd(&X.candCount, 1);
			if (pos < kCandCap) {
				X.cand[pos].d = dj;
				X.cand[pos].idx = j;
			}
		}
	});
}

// Every CTA of the cluster, once per command: append the announced node to the slice if this CTA owns it, then scan.
// OP_SELECT has a cluster barrier in its middle: every CTA of the cluster must call this for the same command.
__device__ void cta_nearest_batch(const RrtWs& W, const RrtSlice& S, const RrtCmd& cmd, int dim, RrtXch& X);

__device__ void cta_serve(cg::cluster_group& cluster, const RrtWs& W, const RrtSlice& S, const RrtCmd& cmd, int dim, RrtXch& X)
{
	if (S.x && cmd.nNew > 0) {
		if ((int)threadIdx.x < cmd.nNew) {
			const int newIdx = cmd.newIdx[threadIdx.x];
			if (((newIdx >> 8) % S.C) == S.rank) {
				const int local = (((newIdx >> 8) / S.C) << 8) | (newIdx & 255);
				S.x[local] = cmd.newState[threadIdx.x][0];
				S.y[local] = cmd.newState[threadIdx.x][1];
				S.th[local] = cmd.newState[threadIdx.x][2];
			}
		}
		__syncthreads();
	}
	const double q[3] = { cmd.q[0][0], cmd.q[0][1], cmd.q[0][2] };
	if (cmd.op == OP_NEAREST)
		cta_nearest_batch(W, S, cmd, dim, X);
	else if (cmd.op == OP_SELECT) {
		cta_select_pass1(W, S, cmd.n, dim, q, X);
		cluster_barrier(!S.x);
		cta_select_pass2(cluster, W, S, cmd.n, dim, q, cmd.k, X);
	} else if (cmd.k == 1)
		cta_scan_slice<1>(W, S, cmd.n, dim, q, 1, X);
	else if (cmd.k <= 8)
		cta_scan_slice<8>(W, S, cmd.n, dim, q, cmd.k, X);
	else
		cta_scan_slice<kRrtMaxK>(W, S, cmd.n, dim, q, cmd.k, X);
}

// OP_NEAREST: the nearest node of each of cmd.nq queries in one pass over this CTA's nodes -> X.ctaList[0..nq)
__device__ void cta_nearest_batch(const RrtWs& W, const RrtSlice& S, const RrtCmd& cmd, int dim, RrtXch& X)
{
	const int nq = cmd.nq;
	double bd[kSpec];
	int bi[kSpec];
	#pragma unroll
	for (int b = 0; b < kSpec; b++) {
		bd[b] = PPX_INF;
		bi[b] = kNoNode;
	}
	for_own_nodes(W, S, cmd.n, [&](double x, double y, double th, int j) {
		#pragma unroll
		for (int b = 0; b < kSpec; b++) {
			if (b < nq) {
				const double dj = query_dist2(cmd.q[b], dim, x, y, th);
				if (nn_closer(dj, j, bd[b], bi[b])) {
					bd[b] = dj;
					bi[b] = j;
				}
			}
		}
	});
	#pragma unroll
	for (int b = 0; b < kSpec; b++) {
		if (b < nq) {
			warp_nn_min(bd[b], bi[b]);
			if ((threadIdx.x & 31) == 0) {
				X.warpList[threadIdx.x >> 5][b].d = bd[b];
				X.warpList[threadIdx.x >> 5][b].idx = bi[b];
			}
		}
	}
	__syncthreads();
	if ((int)threadIdx.x < nq) {
		NnEntry best = X.warpList[0][threadIdx.x];
		for (int w = 1; w < kRrtWarps; w++) {
			const NnEntry e = X.warpList[w][threadIdx.x];
			if (nn_closer(e.d, e.idx, best.d, best.idx))
				best = e;
		}
		X.ctaList[threadIdx.x] = best;
	}
}

// thread 0 of CTA 0: the command of the next query; takes (and empties) the list of nodes appended since the last one
__device__ void cmd_fill(RrtCmd& cmd, const RrtWs& W, int op, int k, int n, int nq, const double (*q)[3], int* pendCount, const int* pendIdx)
{
	cmd.op = op;
	cmd.k = k;
	cmd.n = n;
	cmd.nq = nq;
	for (int b = 0; b < nq; b++)
		for (int i = 0; i < 3; i++)
			cmd.q[b][i] = q[b][i];
	const int nNew = *pendCount;
	cmd.nNew = nNew;
	for (int e = 0; e < nNew; e++) {
		cmd.newIdx[e] = pendIdx[e];
		for (int i = 0; i < 3; i++)
			cmd.newState[e][i] = W.coords[3 * (size_t)pendIdx[e] + i];
	}
	*pendCount = 0;
}

// CTA 0: nearest node of each of nq queries among nodes [0, n) -> outIdx / outD2 [nq]
__device__ void cluster_nearest(cg::cluster_group& cluster, const RrtWs& W, const RrtSlice& S, RrtCmd& cmd, RrtXch& X, int n, int dim,
	const double (*q)[3], int nq, int* pendCount, const int* pendIdx, int* outIdx, double* outD2)
{
	if (threadIdx.x == 0)
		cmd_fill(cmd, W, OP_NEAREST, 1, n, nq, q, pendCount, pendIdx);
	cluster_barrier(!S.x);
	cta_serve(cluster, W, S, cmd, dim, X);
	cluster_barrier(!S.x);
	if ((int)threadIdx.x < nq * S.C) { // one DSMEM read per thread, then the minimum over the CTAs from local memory
		const int r = (int)threadIdx.x / nq, b = (int)threadIdx.x % nq;
		X.gather[r][b] = cluster.map_shared_rank(X.ctaList, r)[b];
	}
	__syncthreads();
	if ((int)threadIdx.x < nq) {
		NnEntry best = X.gather[0][threadIdx.x];
		for (int r = 1; r < S.C; r++) {
			const NnEntry e = X.gather[r][threadIdx.x];
			if (nn_closer(e.d, e.idx, best.d, best.idx))
				best = e;
		}
		outIdx[threadIdx.x] = best.idx == kNoNode ? -1 : best.idx;
		outD2[threadIdx.x] = best.d;
	}
	__syncthreads();
}

// CTA 0: k nearest nodes of `query` among nodes [0, n): results (ascending) into s_idx / s_d2 [k], -1 where the tree has
// fewer than k nodes.  `pendingNew` (node appended since the last query, or -1) travels with the command.
__device__ void cluster_knn(cg::cluster_group& cluster, const RrtWs& W, const RrtSlice& S, RrtCmd& cmd, RrtXch& X, int n, int dim,
	const double* query, int k, int* pendCount, const int* pendIdx, int* s_idx, double* s_d2)
{
	bool select = k > 1 && S.C >= 4;
	while (true) {
		if (threadIdx.x == 0)
			cmd_fill(cmd, W, select ? OP_SELECT : OP_SCAN, k, n, 1, (const double (*)[3])query, pendCount, pendIdx);
		cluster_barrier(!S.x); // command published (and CTA 0's global writes of the tree when the CTAs scan W.coords)
		cta_serve(cluster, W, S, cmd, dim, X);
		cluster_barrier(!S.x); // every CTA's list published
		if (!select)
			break;
		// candidates of all CTAs -> X.gather (256 entries), ranked by counting
		NnEntry* flat = &X.gather[0][0];
		if ((int)threadIdx.x < S.C)
			X.counts[threadIdx.x] = *cluster.map_shared_rank(&X.candCount, (int)threadIdx.x);
		__syncthreads();
		int total = 0, mine = -1, myRank = 0;
		bool overflow = false;
		for (int r = 0; r < S.C; r++) {
			const int cnt = X.counts[r];
			overflow |= cnt > kCandCap;
			if (mine < 0 && (int)threadIdx.x < total + cnt) {
				mine = (int)threadIdx.x - total;
				myRank = r;
			}
			total += cnt;
		}
		overflow |= total > kRrtMaxCluster * kRrtMaxK;
		if (overflow) { // uniform over the CTA: answer the query with the list scan
			select = false; // (the appended nodes went out with the first command: the list is empty now)
			continue;
		}
		if (mine >= 0)
			flat[threadIdx.x] = cluster.map_shared_rank(X.cand, myRank)[mine];
		__syncthreads();
		if ((int)threadIdx.x < total) {
			const NnEntry e = flat[threadIdx.x];
			int rank = 0;
			for (int u = 0; u < total; u++)
				rank += nn_closer(flat[u].d, flat[u].idx, e.d, e.idx) ? 1 : 0;
			if (rank < k) {
				s_idx[rank] = e.idx;
				s_d2[rank] = e.d;
			}
		}
		if ((int)threadIdx.x >= total && (int)threadIdx.x < k) { // the tree has fewer than k nodes
			s_idx[threadIdx.x] = -1;
			s_d2[threadIdx.x] = PPX_INF;
		}
		__syncthreads();
		return;
	}
	if ((int)threadIdx.x < S.C * k) {
		const int r = (int)threadIdx.x / k, e = (int)threadIdx.x % k;
		const NnEntry* remote = cluster.map_shared_rank(X.ctaList, r);
		X.gather[r][e] = remote[e];
	}
	__syncthreads();
	if (threadIdx.x < 32) {
		warp_merge_lists((int)threadIdx.x < S.C ? X.gather[threadIdx.x] : nullptr, k, X.warpList[0]);
		__syncwarp();
		if ((int)threadIdx.x < k) {
			const NnEntry e = X.warpList[0][threadIdx.x];
			s_idx[threadIdx.x] = e.idx == kNoNode ? -1 : e.idx;
			s_d2[threadIdx.x] = e.d;
		}
	}
	__syncthreads();
}

// std::equal_to<Point2d> / std::equal_to<Pose2d> (geometry/2dplane.h:107-134): exact comparison of
// the coordinates, the heading through WrapTheta()
__device__ bool state_equal(const double* a, const double* b, int dim)
{
	if (!(a[0] == b[0] && a[1] == b[1]))
		return false;
	return dim == 2 || wrap_theta(a[2]) == wrap_theta(b[2]);
}

__device__ unsigned state_hash(const double* s, int dim)
{
	// +0.0 folds -0.0 (equal under ==) onto one bit pattern
	unsigned long long h = 1469598103934665603ULL;
	for (int i = 0; i < dim; i++) {
		double v = (i == 2 ? wrap_theta(s[i]) : s[i]) + 0.0;
		h = (h ^ (unsigned long long)__double_as_longlong(v)) * 1099511628211ULL;
	}
	return (unsigned)(h ^ (h >> 32));
}

// m_exploredNodeMap.find(state): node index or -1
__device__ int tree_find(const RrtWs& W, const double* s, int dim)
{
	unsigned slot = state_hash(s, dim) & (unsigned)(W.hashCap - 1);
	while (true) {
		int e = W.hash[slot];
		if (e == 0)
			return -1;
		if (state_equal(W.coords + 3 * (size_t)(e - 1), s, dim))
			return e - 1;
		slot = (slot + 1) & (unsigned)(W.hashCap - 1);
	}
}

__device__ void tree_insert_hash(const RrtWs& W, int node, int dim)
{
	unsigned slot = state_hash(W.coords + 3 * (size_t)node, dim) & (unsigned)(W.hashCap - 1);
	while (W.hash[slot] != 0)
		slot = (slot + 1) & (unsigned)(W.hashCap - 1);
	W.hash[slot] = node + 1;
}

// A connection between two states as the validator sees it after an optional Truncate.
struct Conn {
	double from[3], final_[3];
	double length; // Path::m_length (after truncation)
	RSSeg seg; // Reeds-Shepp only (after the Truncate quirk, trap H9)
};

// PathConnection*::Connect(from, to), then Truncate(ratio) when `truncate` (SteerTowards, rrt.h:132-140)
// `pre`: the optimal segment when the caller has solved it cooperatively (rs_solve_warp / rs_solve_quad), else solved here
__device__ void connect(const RrtDev& P, const double* from, const double* to, bool truncate, double distance, Conn& c,
	const RSSolution* pre = nullptr)
{
	for (int i = 0; i < 3; i++)
		c.from[i] = from[i];
	if (P.connection == 0) {
		// PathR2 / PathSE2: length = (to.position - from.position).norm(), m_final = to
		double dx = to[0] - from[0], dy = to[1] - from[1];
		c.length = sqrt(dx * dx + dy * dy);
		c.final_[0] = to[0];
		c.final_[1] = to[1];
		c.final_[2] = to[2];
		if (truncate && c.length > 0) {
			double ratio = distance / c.length; // std::clamp(distance / length, 0.0, 1.0)
			ratio = ratio < 0.0 ? 0.0 : (1.0 < ratio ? 1.0 : ratio);
			// Truncate: m_final = Interpolate(ratio) = (1 - ratio) * m_init + ratio * m_final; m_length *= ratio
			c.final_[0] = (1 - ratio) * from[0] + ratio * to[0];
			c.final_[1] = (1 - ratio) * from[1] + ratio * to[1];
			if (P.dim == 3)
				c.final_[2] = (1 - ratio) * from[2] + ratio * to[2];
			c.length *= ratio;
		}
		return;
	}
	// PathConnectionReedsShepp::Connect (path_reeds_shepp.cpp:174-179), unit multipliers
	Pose s = { from[0], from[1], from[2] }, g = { to[0], to[1], to[2] };
	RSSolution sol;
	if (pre)
		sol = *pre;
	else
		rs_solve(s, g, P.rmin, 1.0f, 1.0f, 0.0f, 0, sol);
	c.seg = sol.seg;
	c.length = sol.mLength * P.rmin;
	// m_final = Interpolate(1.0)
	Pose f = s;
	if (c.length != 0 && !(truncate && c.length > 0)) { // (Truncate below replaces m_final: the full-length pose is not needed then)
		RSWalker w;
		w.init(s);
		f = w.at(c.seg, P.rmin, 1.0 * c.length);
	}
	if (truncate && c.length > 0) {
		double ratio = distance / c.length;
		ratio = ratio < 0.0 ? 0.0 : (1.0 < ratio ? 1.0 : ratio);
		// PathReedsShepp::Truncate (path_reeds_shepp.cpp:49-94)
		const double totalLength = c.length;
		RSWalker w;
		w.init(s);
		f = w.at(c.seg, P.rmin, ratio * totalLength); // `interp` when the loop breaks
		double length = 0;
		for (int i = 0; i < 5; i++) {
			if (!c.seg.valid(i))
				break;
			double motionLength = c.seg.len[i] * P.rmin;
			if (motionLength == 0)
				continue;
			double motionRatio = (ratio * totalLength - length) / motionLength;
			motionRatio = std_min(motionRatio, 1.0);
			length += motionLength;
			if (length >= ratio * totalLength) {
				c.seg.len[i] *= motionRatio;
				if (i + 1 < 5) {
					// trap H9: the loop over ii resets m_motions[i] itself -> motion i becomes Motion()
					c.seg.len[i] = PPX_INF;
					c.seg.code = (c.seg.code & ~(0xFu << (4 * i))) | (0xAu << (4 * i));
				}
				break;
			}
		}
		c.length *= ratio;
	}
	c.final_[0] = f.x;
	c.final_[1] = f.y;
	c.final_[2] = f.th;
}

// IsPathValid of a connection (StateValidatorFree: always true; StateValidatorOccupancyMap:
// state_validator_occupancy_map.cpp:28-71 with Path::GetLength() = the stored m_length)
__device__ bool conn_valid(const RrtDev& P, const MapDev& map, const Conn& c)
{
	if (P.validator == 0)
		return true;
	const double pathLength = c.length;
	float dist = 0.0f;
	Pose a = { c.from[0], c.from[1], c.from[2] };
	if (pathLength == 0.0)
		return is_state_valid(map, a.x, a.y, a.th, P.minSafeRadius, &dist);
	RSWalker w;
	if (P.connection == 1)
		w.init(a);
	// PathReedsShepp::Interpolate uses m_length as totalLength: the truncated length with the
	// truncated motions (trap H9), exactly what the reference's samples see
	double length = 0.0;
	while (length < pathLength) {
		double ratio = length / pathLength;
		Pose p;
		if (P.connection == 0) {
			p.x = (1 - ratio) * c.from[0] + ratio * c.final_[0];
			p.y = (1 - ratio) * c.from[1] + ratio * c.final_[1];
			p.th = (1 - ratio) * c.from[2] + ratio * c.final_[2];
		} else {
			p = w.at(c.seg, P.rmin, ratio * pathLength);
		}
		if (!is_state_valid(map, p.x, p.y, p.th, P.minSafeRadius, &dist))
			return false;
		length += next_step(map, p.x, p.y, dist, P.minSafeRadius, P.minStep);
	}
	return true;
}

}

namespace {

// IsPathValid of an UNTRUNCATED connection (SteerExactly: the parent candidates of RRT*).  A whole Reeds-Shepp path with its own
// length is exactly what the search kernel's shots are: the certified FP32 cell filter of ppx_fastlane.cuh gives the exact
// loop's verdict (same cells, same steps) at a fraction of the FP64 walker's cost per sample.  Anything else -- straight
// connections, a stored length that is not the segment's own (truncation, trap H9) -- takes the exact loop.
__device__ bool conn_valid_exactly(const RrtDev& P, const MapDev& map, const FastMap& fm, const Conn& c)
{
	if (P.validator == 1 && P.connection == 1 && c.length > 0.0 && rs_total_length(c.seg) * P.rmin == c.length) {
		const Pose a = { c.from[0], c.from[1], c.from[2] };
		float last;
		return rs_is_path_valid_filtered(map, fm, a, c.seg, P.rmin, P.minSafeRadius, P.minStep, &last, nullptr);
	}
	return conn_valid(P, map, c);
}

// Reeds-Shepp solves spread over the CTA: warp w evaluates base formulas w and w + 8 (of 12) of up to eight pairs, one pair
// per quad of lanes (rs_quad_part: the four mirrored goals on the quad's lanes, no divergence on the formula switch); the
// per-warp partial minima of a pair are combined by the same lexicographic (float cost, word) rule.  One warp walking
// all 12 formulas took 17 k cycles per steering solve (PPX_RRT_STATS).
struct RsPart {
	float cost;
	int word;
	double t, u, v;
};

__device__ void rs_finish(const RsPart* parts /* [kRrtWarps] */, double r, RSSolution& sol)
{
	const float INF = __int_as_float(0x7f800000);
	float bestCost = INF;
	int bestWord = 64;
	double t = 0, u = 0, v = 0;
	for (int w = 0; w < kRrtWarps; w++) {
		const RsPart p = parts[w];
		if (p.cost < bestCost || (p.cost == bestCost && p.word < bestWord)) {
			bestCost = p.cost;
			bestWord = p.word;
			t = p.t; u = p.u; v = p.v;
		}
	}
	if (!(bestCost < INF) || bestWord >= 48) {
		sol.word = -1;
		for (int m = 0; m < 5; m++)
			sol.seg.len[m] = PPX_INF;
		sol.seg.code = 0xAAAAA;
		sol.mLength = 0.0;
		sol.cost = INF;
		return;
	}
	sol.word = bestWord;
	sol.mLength = rs_build_segment(bestWord, t, u, v, sol.seg);
	sol.cost = rs_compute_cost(sol.seg, sol.mLength, r, 1.0f, 1.0f, 0.0f);
}

// every thread of the CTA, converged: the part of (from -> to) that this warp's formulas give, pair = quad of lanes
__device__ __forceinline__ void rs_part_store(const double* from, const double* to, double r, bool store, RsPart* parts /* [kRrtWarps] of the pair */)
{
	const Pose ps = { from[0], from[1], from[2] }, pg = { to[0], to[1], to[2] };
	RsPart p;
	double len;
	rs_quad_part(ps, pg, r, 1.0f, 1.0f, 0.0f, 0, (int)(threadIdx.x >> 5), kRrtWarps, p.cost, len, p.word, p.t, p.u, p.v);
	if (store && (threadIdx.x & 3) == 0)
		parts[threadIdx.x >> 5] = p;
}

enum { CTL_RUN = 0, CTL_SKIP = 1, CTL_DONE = 2 };

__global__ void __launch_bounds__(kRrtThreads) k_rrt(RrtDev P, MapDev map, FastMap fm, int64_t ntrees, const double* __restrict__ starts,
	const double* __restrict__ goals, const unsigned long long* __restrict__ seeds, char* __restrict__ slabs, size_t slabBytes,
	int cap, int hashCap, int sliceCap, ppx_rrt_result* __restrict__ results, int64_t nodeCapacity, double* __restrict__ outNodes,
	int32_t* __restrict__ outParents, double* __restrict__ outCosts, unsigned long long* __restrict__ queue,
	unsigned long long* __restrict__ stats)
{
	extern __shared__ __align__(16) double s_sliceMem[]; // [3][sliceCap] when the slices live in shared memory
	__shared__ double s_goal[3];
	__shared__ int s_nearIdx[kRrtMaxK];
	__shared__ double s_nearD2[kRrtMaxK];
	__shared__ double s_candCost[kRrtMaxK];
	__shared__ int s_candValid[kRrtMaxK];
	__shared__ int s_n, s_done, s_inserted, s_pendCount, s_pendIdx[kSpec];
	__shared__ double s_bq[kSpec][3], s_bNew[kSpec][3], s_bNearD2[kSpec];
	__shared__ int s_bNear[kSpec], s_bOk[kSpec], s_bStale[kSpec];
	__shared__ long long s_tree;
	__shared__ RrtCmd s_cmd;
	__shared__ RrtXch s_xch;
	__shared__ RsPart s_part[kRrtMaxK][kRrtWarps];
	// PPX_RRT_STATS: cycles of thread 0 of CTA 0 per phase (nearest query, steering solve, connect, its collision check,
	// neighbourhood query, parent candidates, commit) and the number of iterations / accepted iterations
	__shared__ unsigned long long s_prof[10];
	if (threadIdx.x < 10)
		s_prof[threadIdx.x] = 0;
	long long tPhase = 0;
	#define RRT_PHASE(i) \
		if (stats && threadIdx.x == 0) { \
			const long long now = clock64(); \
			s_prof[i] += (unsigned long long)(now - tPhase); \
			tPhase = now; \
		}

	cg::cluster_group cluster = cg::this_cluster();
	RrtSlice S;
	S.rank = (int)cluster.block_rank();
	S.C = (int)cluster.num_blocks();
	S.x = sliceCap > 0 ? s_sliceMem : nullptr;
	S.y = S.x + sliceCap;
	S.th = S.y + sliceCap;

	RrtWs W;
	{
		char* p = slabs + (size_t)(blockIdx.x / S.C) * slabBytes;
		W.coords = (double*)p; p += (size_t)cap * 24;
		W.cost = (double*)p; p += (size_t)cap * 8;
		W.mt = (unsigned long long*)p; p += 312 * 8;
		W.parent = (int*)p; p += (size_t)cap * 4;
		W.hash = (int*)p;
		W.cap = cap;
		W.hashCap = hashCap;
	}

	if (S.rank != 0) {
		// scan servers: one command per neighbour query of CTA 0 until it says the cluster's trees are done
		const RrtCmd* remote = cluster.map_shared_rank(&s_cmd, 0);
		while (true) {
			cluster_barrier(!S.x);
			for (int i = threadIdx.x; i < (int)(sizeof(RrtCmd) / 8); i += blockDim.x)
				((unsigned long long*)&s_cmd)[i] = ((const unsigned long long*)remote)[i];
			__syncthreads();
			const int op = s_cmd.op;
			if (op != OP_EXIT)
				cta_serve(cluster, W, S, s_cmd, P.dim, s_xch);
			cluster_barrier(!S.x); // CTA 0 reads the lists after this one (and does not exit before it)
			if (op == OP_EXIT)
				return;
		}
	}

	while (true) {
		if (threadIdx.x == 0)
			s_tree = (long long)atomicAdd(queue, 1ull);
		__syncthreads();
		const long long t = s_tree;
		if (t >= ntrees) {
			if (threadIdx.x == 0)
				s_cmd.op = OP_EXIT;
			cluster_barrier(!S.x);
			cluster_barrier(!S.x);
			return;
		}
		for (int i = threadIdx.x; i < hashCap; i += blockDim.x)
			W.hash[i] = 0;
		__syncthreads();

		// state of the reference's loop: count / n / nEdges / status in every thread (uniform), RNG and solution on thread 0
		RrtRng rng { W.mt, 0 };
		int count = -1, status = 1, solution = -1, nEdges = 0;
		int n = 1;
		if (threadIdx.x == 0) {
			rng.seed(seeds[t]);
			for (int i = 0; i < 3; i++) {
				s_goal[i] = goals[3 * t + i];
				W.coords[i] = starts[3 * t + i];
			}
			if (P.dim == 2) {
				W.coords[2] = 0.0;
				s_goal[2] = 0.0;
			}
			W.parent[0] = -1;
			W.cost[0] = 0.0; // NodeInfo::costFromGoal = 0, rrt_star.h:36
			tree_insert_hash(W, 0, P.dim); // Tree::CreateRootNode
			s_pendCount = 1; // the nodes the next command announces to the cluster
			s_pendIdx[0] = 0;
		}
		__syncthreads();

		// SPECULATION OVER REJECTED ITERATIONS.  An iteration whose steering connection is invalid changes nothing (rrt.h:75-77,
		// rrt_star.h:78-79: `continue`), and the random numbers an iteration draws do not depend on the tree.  So kSpec
		// samples are drawn ahead, their nearest nodes come from ONE pass over the tree, their steering connections are
		// solved and checked in parallel (Reeds-Shepp words over the warps by base formula, one sample per warp for the
		// rest), and the batch is then resolved IN ORDER: a rejected sample only counts; an accepted one runs the
		// reference's statements (neighbourhood, best parent, Extend, goal test) against the tree as it is then; after an
		// insertion the later samples of the batch compare their nearest node with the new one and are re-steered if it
		// is nearer.  Iteration counts, edge counts and the tree are those of the one-sample-at-a-time loop.
		auto steer_one = [&](int b, const RsPart* parts) { // one thread: SteerTowards + IsPathValid of sample b
			RSSolution sol;
			if (P.connection == 1)
				rs_finish(parts, P.rmin, sol);
			Conn c;
			connect(P, W.coords + 3 * (size_t)s_bNear[b], s_bq[b], true, P.maxConnectionDistance, c, P.connection == 1 ? &sol : nullptr);
			const bool ok = conn_valid(P, map, c);
			s_bOk[b] = ok ? 1 : 0;
			s_bNew[b][0] = c.final_[0];
			s_bNew[b][1] = c.final_[1];
			s_bNew[b][2] = P.dim == 3 ? c.final_[2] : 0.0;
		};
		bool done = false;
		while (!done) {
			if (threadIdx.x == 0) {
				for (int b = 0; b < kSpec; b++) {
					// Random<double>::SampleUniform(0, 1) < goalBias ? goal : m_stateSpace->SampleUniform()
					if (rng.uniform(0.0, 1.0) < P.goalBias) {
						s_bq[b][0] = s_goal[0];
						s_bq[b][1] = s_goal[1];
						s_bq[b][2] = s_goal[2];
					} else {
						s_bq[b][0] = rng.uniform(P.lb[0], P.ub[0]);
						s_bq[b][1] = rng.uniform(P.lb[1], P.ub[1]);
						s_bq[b][2] = P.dim == 3 ? rng.uniform(P.lb[2], P.ub[2]) : 0.0;
					}
					s_bStale[b] = 0;
				}
				if (stats)
					tPhase = clock64();
			}
			__syncthreads();
			// Tree::GetNearestNode(randomState) for the whole batch
			cluster_nearest(cluster, W, S, s_cmd, s_xch, n, P.dim, s_bq, kSpec, &s_pendCount, s_pendIdx, s_bNear, s_bNearD2);
			RRT_PHASE(0)
			if (P.connection == 1) {
				const int g = (int)((threadIdx.x & 31) >> 2); // sample g on quad g of every warp
				rs_part_store(W.coords + 3 * (size_t)s_bNear[g], s_bq[g], P.rmin, true, s_part[g]);
				__syncthreads();
			}
			RRT_PHASE(1)
			if ((threadIdx.x & 31) == 0)
				steer_one((int)(threadIdx.x >> 5), s_part[threadIdx.x >> 5]);
			__syncthreads();
			RRT_PHASE(3)
			for (int b = 0; b < kSpec; b++) {
				count++; // every thread keeps the counters: the limits are checked without a broadcast
				if (count > (long long)P.maxIteration || (unsigned)n > P.maxNodes) { // rrt.h:64-69
					status = -1;
					done = true;
					break;
				}
				if (stats && threadIdx.x == 0)
					s_prof[8]++;
				if (s_bStale[b]) { // a node inserted earlier in this batch is nearer: steer from it
					if (P.connection == 1) {
						rs_part_store(W.coords + 3 * (size_t)s_bNear[b], s_bq[b], P.rmin, (threadIdx.x & 31) < 4, s_part[0]);
						__syncthreads();
					}
					if (threadIdx.x == 0) {
						steer_one(b, s_part[0]);
						if (stats)
							s_prof[2]++;
					}
					__syncthreads();
				}
				nEdges++;
				if (!s_bOk[b])
					continue;
				const double* newState = s_bNew[b];
				const int nearest = s_bNear[b];
				const int size = n;
				if (threadIdx.x == 0) {
					s_done = 0;
					s_inserted = -1;
				}
				if (!P.star) {
					if (threadIdx.x == 0) {
						// Tree::Extend(newState, nearestNode), tree.h:124-146
						int node = tree_find(W, newState, P.dim);
						if (node < 0) {
							if (n >= W.cap) {
								status = PPX_ERR_CAPACITY;
								s_done = 1;
							} else {
								node = n;
								W.coords[3 * (size_t)node] = newState[0];
								W.coords[3 * (size_t)node + 1] = newState[1];
								W.coords[3 * (size_t)node + 2] = newState[2];
								W.parent[node] = nearest;
								W.cost[node] = 0.0;
								tree_insert_hash(W, node, P.dim);
								s_pendIdx[s_pendCount++] = node;
								s_inserted = node;
							}
						}
						// RRT::IsSolution: (state - goal).norm() < 1, rrt.h:125-128
						double dx = newState[0] - s_goal[0], dy = newState[1] - s_goal[1];
						if (!s_done && sqrt(dx * dx + dy * dy) < 1) {
							solution = node;
							status = PPX_OK;
							s_done = 1;
						}
					}
				} else {
					// RRT*: best parent among the nn nearest nodes of newState, rrt_star.h:84-97
					unsigned nnu = (unsigned)log((double)size); // std::max<unsigned int>(1, log(m_tree.Size()))
					if (nnu < 1u)
						nnu = 1u;
					const int nn = (int)(nnu > (unsigned)kRrtMaxK ? (unsigned)kRrtMaxK : nnu);
					if (stats && threadIdx.x == 0) {
						tPhase = clock64();
						s_prof[9]++;
					}
					cluster_knn(cluster, W, S, s_cmd, s_xch, size, P.dim, newState, nn, &s_pendCount, s_pendIdx, s_nearIdx, s_nearD2);
					RRT_PHASE(4)
					if (P.connection == 1) {
						// candidate j on quad j % 8 of every warp (warp w: base formulas w, w + 8), then thread j combines the
						// eight parts, builds and checks the connection
						for (int jb = 0; jb < nn; jb += 8) {
							const int j = jb + (int)((threadIdx.x & 31) >> 2);
							const bool have = j < nn && s_nearIdx[j] >= 0;
							rs_part_store(W.coords + 3 * (size_t)(have ? s_nearIdx[j] : 0), newState, P.rmin, have, s_part[have ? j : 0]);
						}
						__syncthreads();
						// candidate j on lane j / 8 of warp j % 8: the connections of different candidates take different branches
						// (segment types, sample counts); side by side in one warp they would run one after the other
						const int j = (int)(threadIdx.x & 31) * kRrtWarps + (int)(threadIdx.x >> 5);
						if ((threadIdx.x & 31) < 2 && j < nn && s_nearIdx[j] >= 0) {
							const int node = s_nearIdx[j];
							RSSolution sol;
							rs_finish(s_part[j], P.rmin, sol);
							Conn c;
							connect(P, W.coords + 3 * (size_t)node, newState, false, 0.0, c, &sol); // SteerExactly
							s_candCost[j] = W.cost[node] + c.length;
							s_candValid[j] = conn_valid_exactly(P, map, fm, c) ? 1 : 0;
						}
					} else {
						const int j = (int)(threadIdx.x & 31) * kRrtWarps + (int)(threadIdx.x >> 5); // as above
						if ((threadIdx.x & 31) < 2 && j < nn && s_nearIdx[j] >= 0) {
							const int node = s_nearIdx[j];
							Conn c;
							connect(P, W.coords + 3 * (size_t)node, newState, false, 0.0, c); // SteerExactly
							s_candCost[j] = W.cost[node] + c.length;
							s_candValid[j] = conn_valid(P, map, c) ? 1 : 0;
						}
					}
					__syncthreads();
					RRT_PHASE(5)
					if (threadIdx.x == 0) {
						int bestParent = -1;
						double bestCost = PPX_INF;
						for (int j = 0; j < nn; j++) {
							if (s_nearIdx[j] < 0)
								break;
							if (s_candCost[j] < bestCost) {
								nEdges++; // the reference evaluates IsPathValid only here (short-circuit &&)
								if (s_candValid[j]) {
									bestParent = s_nearIdx[j];
									bestCost = s_candCost[j];
								}
							}
						}
						// Tree::Extend(newState, bestParentNode): an existing state returns its node unchanged;
						// a null parent falls back to the nearest node of the target (tree.h:127-135)
						int node = tree_find(W, newState, P.dim);
						if (node < 0) {
							if (n >= W.cap) {
								status = PPX_ERR_CAPACITY;
								s_done = 1;
							} else {
								node = n;
								W.coords[3 * (size_t)node] = newState[0];
								W.coords[3 * (size_t)node + 1] = newState[1];
								W.coords[3 * (size_t)node + 2] = newState[2];
								W.parent[node] = bestParent >= 0 ? bestParent : s_nearIdx[0];
								tree_insert_hash(W, node, P.dim);
								s_pendIdx[s_pendCount++] = node;
								s_inserted = node;
							}
						}
						if (!s_done) {
							W.cost[node] = bestCost; // (*newNode)->info.costFromGoal = bestCost, also for an existing node
							if (state_equal(newState, s_goal, P.dim)) { // RRTStar::IsSolution, rrt_star.h:136-139
								solution = node;
								status = PPX_OK;
								s_done = 1;
							}
						}
					}
					RRT_PHASE(6)
				}
				__syncthreads();
				const int ins = s_inserted;
				if (ins >= 0)
					n = ins + 1;
				if (s_done) {
					done = true;
					break;
				}
				// the later samples of the batch: is the new node nearer than the nearest node they were steered from?
				if (ins >= 0 && (int)threadIdx.x > b && (int)threadIdx.x < kSpec) {
					const double d = state_dist2(s_bq[threadIdx.x], W.coords + 3 * (size_t)ins, P.dim);
					if (d < s_bNearD2[threadIdx.x]) { // the new node has the largest index: it wins strictly only
						s_bNear[threadIdx.x] = ins;
						s_bNearD2[threadIdx.x] = d;
						s_bStale[threadIdx.x] = 1;
					}
				}
				__syncthreads();
			}
		}
		if (stats && threadIdx.x == 0)
			for (int i = 0; i < 10; i++)
				atomicAdd(&stats[i], s_prof[i]);

		if (threadIdx.x == 0) {
			ppx_rrt_result r;
			r.status = status == 1 ? -1 : status;
			r.iterations = count;
			r.n_nodes = n;
			r.solution = solution;
			r.path_len = 0;
			for (int m = solution; m >= 0; m = W.parent[m])
				r.path_len++;
			r.n_edges_checked = nEdges;
			r.solution_cost = solution >= 0 ? W.cost[solution] : PPX_INF;
			results[t] = r;
			s_n = n;
		}
		__syncthreads();
		const int total = s_n < nodeCapacity ? s_n : (int)nodeCapacity;
		for (int i = threadIdx.x; i < total; i += blockDim.x) {
			if (outNodes) {
				outNodes[((size_t)t * nodeCapacity + i) * 3] = W.coords[3 * (size_t)i];
				outNodes[((size_t)t * nodeCapacity + i) * 3 + 1] = W.coords[3 * (size_t)i + 1];
				outNodes[((size_t)t * nodeCapacity + i) * 3 + 2] = W.coords[3 * (size_t)i + 2];
			}
			if (outParents)
				outParents[(size_t)t * nodeCapacity + i] = W.parent[i];
			if (outCosts)
				outCosts[(size_t)t * nodeCapacity + i] = W.cost[i];
		}
		__syncthreads();
	}
}

}

extern "C" {

int ppx_rrt_run(ppx_ctx* ctx, const ppx_rrt_params* params, const ppx_validator_params* vp, int64_t ntrees,
	const double* starts, const double* goals, const uint64_t* seeds, ppx_rrt_result* results, int64_t node_capacity,
	double* nodes, int32_t* parents, double* costs)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	PPX_REQUIRE(ctx, params && ntrees >= 0 && (ntrees == 0 || (starts && goals && seeds && results)), "bad arguments");
	PPX_REQUIRE(ctx, params->dim == 2 || params->dim == 3, "rrt: dim must be 2 or 3");
	PPX_REQUIRE(ctx, params->connection == 0 || (params->connection == 1 && params->dim == 3), "rrt: Reeds-Shepp connection needs dim 3");
	PPX_REQUIRE(ctx, params->star || params->dim == 2, "rrt: RRT (non-star) exists for R^2 only (rrt.h:125-128,154-155)");
	PPX_REQUIRE(ctx, params->validator == 0 || (params->dim == 3 && vp), "rrt: the occupancy-map validator needs dim 3 and validator parameters");
	PPX_REQUIRE(ctx, params->max_nodes < (1u << 28), "rrt: max_nodes too large");
	PPX_REQUIRE(ctx, !(nodes || parents || costs) || node_capacity > 0, "rrt: node outputs need node_capacity");
	if (params->validator == 1 && !ctx->has_map)
		return ppx_fail(ctx, PPX_ERR_STATE, "no map uploaded (ppx_map_upload)");
	if (ntrees == 0)
		return PPX_OK;
	RrtDev P {};
	P.dim = params->dim;
	P.connection = params->connection;
	P.star = params->star;
	P.validator = params->validator;
	P.maxIteration = params->max_iteration;
	P.maxNodes = params->max_nodes;
	P.maxConnectionDistance = params->max_connection_distance;
	P.goalBias = params->goal_bias;
	P.rmin = params->min_turning_radius;
	for (int i = 0; i < 3; i++) {
		P.lb[i] = params->lb[i];
		P.ub[i] = params->ub[i];
	}
	P.minSafeRadius = vp ? vp->min_safe_radius : 0.0f;
	P.minStep = vp ? vp->min_path_interpolation_distance : 0.1f;
	const int cap = (int)params->max_nodes + 2;
	int hashCap = 1024;
	while (hashCap < 2 * cap)
		hashCap <<= 1;
	size_t slab = ((size_t)cap * 36 + 312 * 8 + (size_t)hashCap * 4 + 255) & ~(size_t)255;
	// Cluster size: as many SMs per tree as the batch leaves free (option "rrt_cluster" / PPX_RRT_CLUSTER overrides: 1..16).
	int C = 1;
	while (C < kRrtMaxCluster && ntrees * (2 * C) <= (int64_t)ctx->num_sms)
		C *= 2;
	if (const char* e = getenv("PPX_RRT_CLUSTER")) {
		const int v = atoi(e);
		if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16)
			C = v;
	}
	PPX_CUDA(ctx, cudaFuncSetAttribute(k_rrt, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
	cudaLaunchConfig_t cfg {};
	cudaLaunchAttribute attr[1];
	int sliceCap = 0, clusters = 0;
	size_t smem = 0;
	for (;; C /= 2) {
		// slices in shared memory when they fit (200 KB per CTA; 96 KB without a cluster so that two trees share an SM)
		sliceCap = ((cap + 256 * C - 1) / (256 * C)) * 256;
		smem = (size_t)sliceCap * 24;
		if (smem > (C == 1 ? (size_t)96 : (size_t)200) * 1024) {
			sliceCap = 0;
			smem = 0;
		}
		PPX_CUDA(ctx, cudaFuncSetAttribute(k_rrt, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		cfg.gridDim = dim3((unsigned)C, 1, 1);
		cfg.blockDim = dim3(kRrtThreads, 1, 1);
		cfg.dynamicSmemBytes = smem;
		cfg.stream = ctx->stream;
		attr[0].id = cudaLaunchAttributeClusterDimension;
		attr[0].val.clusterDim.x = (unsigned)C;
		attr[0].val.clusterDim.y = 1;
		attr[0].val.clusterDim.z = 1;
		cfg.attrs = attr;
		cfg.numAttrs = 1;
		int maxClusters = 0;
		cudaError_t err = cudaOccupancyMaxActiveClusters(&maxClusters, k_rrt, &cfg);
		if (err == cudaSuccess && maxClusters > 0) {
			clusters = (int)std::min<int64_t>(ntrees, (int64_t)maxClusters);
			break;
		}
		cudaGetLastError();
		if (C == 1)
			return ppx_fail(ctx, PPX_ERR_CUDA, "rrt: no launch configuration (%s)", cudaGetErrorString(err));
	}
	cfg.gridDim = dim3((unsigned)(clusters * C), 1, 1);
	size_t off = 0;
	auto carve = [&](size_t bytes) {
		size_t o = off;
		off += (bytes + 255) & ~(size_t)255;
		return o;
	};
	const bool wantNodes = nodes || parents || costs;
	size_t oSlabs = carve((size_t)clusters * slab), oStarts = carve(ntrees * 24), oGoals = carve(ntrees * 24), oSeeds = carve(ntrees * 8);
	size_t oRes = carve(ntrees * sizeof(ppx_rrt_result));
	size_t oNodes = carve(nodes ? (size_t)ntrees * node_capacity * 24 : 0), oPar = carve(parents ? (size_t)ntrees * node_capacity * 4 : 0);
	size_t oCost = carve(costs ? (size_t)ntrees * node_capacity * 8 : 0);
	char* base;
	PPX_CHECK(ppx_scratch(ctx, 6, off, (void**)&base));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oStarts, starts, ntrees * 24, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oGoals, goals, ntrees * 24, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oSeeds, seeds, ntrees * 8, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemsetAsync(ctx->d_queue, 0, sizeof(unsigned long long), ctx->stream));
	MapDev map = ctx->has_map ? ctx->map_dev() : MapDev {};
	static const bool statsOn = getenv("PPX_RRT_STATS") != nullptr; // phase cycles to stderr (tools/time_rrt_single.py)
	static unsigned long long* d_stats = nullptr;
	if (statsOn && !d_stats)
		PPX_CUDA(ctx, cudaMalloc(&d_stats, 10 * sizeof(unsigned long long)));
	if (statsOn)
		PPX_CUDA(ctx, cudaMemsetAsync(d_stats, 0, 10 * sizeof(unsigned long long), ctx->stream));
	FastMap fm = ctx->has_map ? fast_map_derive(ctx->desc) : FastMap {};
	if (!ctx->rs_filter)
		fm.enabled = 0;
	PPX_CUDA(ctx, cudaLaunchKernelEx(&cfg, k_rrt, P, map, fm, (int64_t)ntrees, (const double*)(base + oStarts), (const double*)(base + oGoals),
		(const unsigned long long*)(base + oSeeds), base + oSlabs, slab, cap, hashCap, sliceCap, (ppx_rrt_result*)(base + oRes),
		(int64_t)(wantNodes ? node_capacity : 0), nodes ? (double*)(base + oNodes) : (double*)nullptr,
		parents ? (int32_t*)(base + oPar) : (int32_t*)nullptr, costs ? (double*)(base + oCost) : (double*)nullptr, ctx->d_queue,
		d_stats));
	if (d_stats) {
		unsigned long long st[10];
		cudaStreamSynchronize(ctx->stream);
		cudaMemcpy(st, d_stats, sizeof(st), cudaMemcpyDeviceToHost);
		const double it = (double)std::max<unsigned long long>(1, st[8]), acc = (double)std::max<unsigned long long>(1, st[9]);
		fprintf(stderr, "rrt stats (cluster %d, slices in %s, batches of %d samples): %llu iterations, %llu accepted, %llu re-steered | cycles "
			"per iteration (batch phases shared by the samples of a batch): nearest %.0f, steer solve %.0f, connect + check %.0f | per accepted "
			"iteration: neighbourhood %.0f, candidates %.0f, commit %.0f\n",
			C, sliceCap ? "shared memory" : "global memory", kSpec, st[8], st[9], st[2], st[0] / it, st[1] / it, st[3] / it, st[4] / acc,
			st[5] / acc, st[6] / acc);
	}
	PPX_LAUNCHED(ctx);
	PPX_CUDA(ctx, cudaMemcpyAsync(results, base + oRes, ntrees * sizeof(ppx_rrt_result), cudaMemcpyDeviceToHost, ctx->stream));
	if (nodes)
		PPX_CUDA(ctx, cudaMemcpyAsync(nodes, base + oNodes, (size_t)ntrees * node_capacity * 24, cudaMemcpyDeviceToHost, ctx->stream));
	if (parents)
		PPX_CUDA(ctx, cudaMemcpyAsync(parents, base + oPar, (size_t)ntrees * node_capacity * 4, cudaMemcpyDeviceToHost, ctx->stream));
	if (costs)
		PPX_CUDA(ctx, cudaMemcpyAsync(costs, base + oCost, (size_t)ntrees * node_capacity * 8, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

} // extern "C"
