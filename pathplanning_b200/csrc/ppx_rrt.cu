// libppx.so: RRT and RRT* (A11) with the brute-force nearest-neighbour scan (A10) in the loop,
// one tree per CTA, batched over trees (algo/rrt.h:55-95, algo/rrt_star.h:53-112, utils/tree.h).
//
// A tree is a chain of dependent iterations, each dominated by one or two nearest-neighbour
// queries over all current nodes.  The CTA's 256 threads scan the node array in parallel (exact
// FP64 squared distances, ties to the lower insertion index like the oracle's flann stand-in) and
// merge with a k-round block argmin; the RRT* parent candidates (k <= 16 Reeds-Shepp connections +
// collision checks) are evaluated one per thread; the scalar control flow -- RNG draws, steering,
// duplicate-state rule, goal test -- runs on thread 0 in the reference's statement order.
// Independent trees (different seeds / queries) fill the other SMs.
#include <algorithm>
#include "ppx_ctx.h"

using namespace ppx;

namespace {

constexpr int kRrtThreads = 256;
constexpr int kRrtMaxK = 16;

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

// k nearest nodes of `query` among nodes [0, n): results (ascending) into s_idx / s_d2 [k].
__device__ void cta_knn(const RrtWs& W, int n, int dim, const double* query, int k, int* s_idx, double* s_d2,
	double* s_wd, int* s_wi, int* s_wo, int* s_winner)
{
	double bd[kRrtMaxK];
	int bi[kRrtMaxK];
	for (int i = 0; i < k; i++) {
		bd[i] = PPX_INF;
		bi[i] = 0x7fffffff;
	}
	double q[3] = { query[0], query[1], query[2] };
	// four nodes per thread and turn: their coordinate loads are independent and in flight together (the scan is bound by
	// L2 latency, not bandwidth, when every turn waits for its own load); candidates are taken in ascending index order
	constexpr int kU = 4;
	for (int j0 = threadIdx.x; j0 < n; j0 += kU * blockDim.x) {
		double dd[kU];
		#pragma unroll
		for (int u = 0; u < kU; u++) {
			const int j = j0 + u * (int)blockDim.x;
			dd[u] = j < n ? state_dist2(q, W.coords + 3 * (size_t)j, dim) : PPX_INF;
		}
		#pragma unroll
		for (int u = 0; u < kU; u++) {
			const int j = j0 + u * (int)blockDim.x;
			const double d = dd[u];
			if (j < n && nn_closer(d, j, bd[k - 1], bi[k - 1])) {
				int p = k - 1;
				while (p > 0 && nn_closer(d, j, bd[p - 1], bi[p - 1])) {
					bd[p] = bd[p - 1];
					bi[p] = bi[p - 1];
					p--;
				}
				bd[p] = d;
				bi[p] = j;
			}
		}
	}
	int head = 0;
	for (int round = 0; round < k; round++) {
		double d = head < k ? bd[head] : PPX_INF;
		int idx = head < k ? bi[head] : 0x7fffffff;
		int owner = threadIdx.x;
		#pragma unroll
		for (int o = 16; o > 0; o >>= 1) {
			double od = __shfl_down_sync(0xffffffffu, d, o);
			int oi = __shfl_down_sync(0xffffffffu, idx, o);
			int oo = __shfl_down_sync(0xffffffffu, owner, o);
			if (nn_closer(od, oi, d, idx)) {
				d = od;
				idx = oi;
				owner = oo;
			}
		}
		if ((threadIdx.x & 31) == 0) {
			s_wd[threadIdx.x >> 5] = d;
			s_wi[threadIdx.x >> 5] = idx;
			s_wo[threadIdx.x >> 5] = owner;
		}
		__syncthreads();
		if (threadIdx.x == 0) {
			int best = 0;
			for (int w = 1; w < (int)(blockDim.x >> 5); w++)
				if (nn_closer(s_wd[w], s_wi[w], s_wd[best], s_wi[best]))
					best = w;
			bool have = s_wi[best] != 0x7fffffff;
			s_idx[round] = have ? s_wi[best] : -1;
			s_d2[round] = s_wd[best];
			*s_winner = have ? s_wo[best] : -1;
		}
		__syncthreads();
		if ((int)threadIdx.x == *s_winner)
			head++;
		__syncthreads();
	}
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
	if (c.length != 0) {
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

enum { CTL_RUN = 0, CTL_SKIP = 1, CTL_DONE = 2 };

__global__ void __launch_bounds__(kRrtThreads) k_rrt(RrtDev P, MapDev map, int64_t ntrees, const double* __restrict__ starts,
	const double* __restrict__ goals, const unsigned long long* __restrict__ seeds, char* __restrict__ slabs, size_t slabBytes,
	int cap, int hashCap, ppx_rrt_result* __restrict__ results, int64_t nodeCapacity, double* __restrict__ outNodes,
	int32_t* __restrict__ outParents, double* __restrict__ outCosts, unsigned long long* __restrict__ queue)
{
	__shared__ double s_query[3], s_new[3], s_goal[3];
	__shared__ int s_nearIdx[kRrtMaxK];
	__shared__ double s_nearD2[kRrtMaxK];
	__shared__ double s_candCost[kRrtMaxK];
	__shared__ int s_candValid[kRrtMaxK];
	__shared__ double s_wd[8];
	__shared__ int s_wi[8], s_wo[8];
	__shared__ int s_winner, s_ctlbuf[2], s_n, s_nn;
	__shared__ long long s_tree;

	RrtWs W;
	{
		char* p = slabs + (size_t)blockIdx.x * slabBytes;
		W.coords = (double*)p; p += (size_t)cap * 24;
		W.cost = (double*)p; p += (size_t)cap * 8;
		W.mt = (unsigned long long*)p; p += 312 * 8;
		W.parent = (int*)p; p += (size_t)cap * 4;
		W.hash = (int*)p;
		W.cap = cap;
		W.hashCap = hashCap;
	}

	while (true) {
		if (threadIdx.x == 0)
			s_tree = (long long)atomicAdd(queue, 1ull);
		__syncthreads();
		const long long t = s_tree;
		if (t >= ntrees)
			return;
		for (int i = threadIdx.x; i < hashCap; i += blockDim.x)
			W.hash[i] = 0;
		__syncthreads();

		// thread-0 state of the reference's loop
		RrtRng rng { W.mt, 0 };
		int count = -1, status = 1, solution = -1, nEdges = 0;
		int n = 0;
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
			n = 1;
		}
		__syncthreads();

		// The control word is double-buffered by iteration parity: thread 0 may already be writing the next
		// iteration's word (after its last barrier of this iteration) while a late warp still reads this one.
		unsigned it = 0;
		while (true) {
			it++;
			volatile int& s_ctl = s_ctlbuf[it & 1];
			if (threadIdx.x == 0) {
				count++;
				if (count > (long long)P.maxIteration || (unsigned)n > P.maxNodes) { // rrt.h:64-69
					status = -1;
					s_ctl = CTL_DONE;
				} else {
					// Random<double>::SampleUniform(0, 1) < goalBias ? goal : m_stateSpace->SampleUniform()
					if (rng.uniform(0.0, 1.0) < P.goalBias) {
						s_query[0] = s_goal[0];
						s_query[1] = s_goal[1];
						s_query[2] = s_goal[2];
					} else {
						s_query[0] = rng.uniform(P.lb[0], P.ub[0]);
						s_query[1] = rng.uniform(P.lb[1], P.ub[1]);
						s_query[2] = P.dim == 3 ? rng.uniform(P.lb[2], P.ub[2]) : 0.0;
					}
					s_ctl = CTL_RUN;
					s_n = n;
				}
			}
			__syncthreads();
			if (s_ctl == CTL_DONE)
				break;
			const int size = s_n;
			// Tree::GetNearestNode(randomState)
			cta_knn(W, size, P.dim, s_query, 1, s_nearIdx, s_nearD2, s_wd, s_wi, s_wo, &s_winner);
			int nearest = -1;
			RSSolution steerSol;
			if (P.connection == 1 && threadIdx.x < 32) { // the 48 words of the steering connection on the lanes of warp 0
				const double* f = W.coords + 3 * (size_t)s_nearIdx[0];
				const Pose ps = { f[0], f[1], f[2] }, pg = { s_query[0], s_query[1], s_query[2] };
				rs_solve_warp(ps, pg, P.rmin, 1.0f, 1.0f, 0.0f, steerSol);
			}
			if (threadIdx.x == 0) {
				nearest = s_nearIdx[0];
				Conn c;
				connect(P, W.coords + 3 * (size_t)nearest, s_query, true, P.maxConnectionDistance, c,
					P.connection == 1 ? &steerSol : nullptr); // SteerTowards
				nEdges++;
				if (!conn_valid(P, map, c)) {
					s_ctl = CTL_SKIP;
				} else {
					s_new[0] = c.final_[0];
					s_new[1] = c.final_[1];
					s_new[2] = P.dim == 3 ? c.final_[2] : 0.0;
					if (P.star) {
						unsigned nn = (unsigned)log((double)size); // std::max<unsigned int>(1, log(m_tree.Size()))
						if (nn < 1u)
							nn = 1u;
						s_nn = (int)(nn > (unsigned)kRrtMaxK ? (unsigned)kRrtMaxK : nn);
					}
				}
			}
			__syncthreads();
			if (s_ctl == CTL_SKIP)
				continue;
			if (!P.star) {
				if (threadIdx.x == 0) {
					// Tree::Extend(newState, nearestNode), tree.h:124-146
					int node = tree_find(W, s_new, P.dim);
					if (node < 0) {
						if (n >= W.cap) {
							status = PPX_ERR_CAPACITY;
							s_ctl = CTL_DONE;
						} else {
							node = n++;
							W.coords[3 * (size_t)node] = s_new[0];
							W.coords[3 * (size_t)node + 1] = s_new[1];
							W.coords[3 * (size_t)node + 2] = s_new[2];
							W.parent[node] = nearest;
							W.cost[node] = 0.0;
							tree_insert_hash(W, node, P.dim);
						}
					}
					// RRT::IsSolution: (state - goal).norm() < 1, rrt.h:125-128
					double dx = s_new[0] - s_goal[0], dy = s_new[1] - s_goal[1];
					if (s_ctl != CTL_DONE && sqrt(dx * dx + dy * dy) < 1) {
						solution = node;
						status = PPX_OK;
						s_ctl = CTL_DONE;
					}
				}
				__syncthreads();
				if (s_ctl == CTL_DONE)
					break;
				continue;
			}
			// RRT*: best parent among the nn nearest nodes of newState, rrt_star.h:84-97
			const int nn = s_nn;
			cta_knn(W, size, P.dim, s_new, nn, s_nearIdx, s_nearD2, s_wd, s_wi, s_wo, &s_winner);
			if (P.connection == 1) {
				// candidate j on the quad of lanes 4j..4j+3 (rs_solve_quad: the 4 mirrored goals on the lanes, the 12 base
				// formulas in a loop); lane 4j then builds and checks the connection
				if (threadIdx.x < 4 * kRrtMaxK) {
					const int j = (int)threadIdx.x >> 2;
					const bool have = j < nn && s_nearIdx[j] >= 0;
					const int node = have ? s_nearIdx[j] : 0;
					const double* f = W.coords + 3 * (size_t)node;
					const Pose ps = { f[0], f[1], f[2] }, pg = { s_new[0], s_new[1], s_new[2] };
					int w;
					double t, u, v;
					rs_solve_quad(ps, pg, P.rmin, 1.0f, 1.0f, 0.0f, 0, w, t, u, v);
					if (have && (threadIdx.x & 3) == 0) {
						RSSolution sol;
						sol.word = w;
						if (w >= 0) {
							sol.mLength = rs_build_segment(w, t, u, v, sol.seg);
							sol.cost = 0.0f;
						} else {
							for (int m = 0; m < 5; m++)
								sol.seg.len[m] = PPX_INF;
							sol.seg.code = 0xAAAAA;
							sol.mLength = 0.0;
							sol.cost = __int_as_float(0x7f800000);
						}
						Conn c;
						connect(P, f, s_new, false, 0.0, c, &sol); // SteerExactly
						s_candCost[j] = W.cost[node] + c.length;
						s_candValid[j] = conn_valid(P, map, c) ? 1 : 0;
					}
				}
			} else if ((int)threadIdx.x < nn && s_nearIdx[threadIdx.x] >= 0) {
				const int node = s_nearIdx[threadIdx.x];
				Conn c;
				connect(P, W.coords + 3 * (size_t)node, s_new, false, 0.0, c); // SteerExactly
				s_candCost[threadIdx.x] = W.cost[node] + c.length;
				s_candValid[threadIdx.x] = conn_valid(P, map, c) ? 1 : 0;
			}
			__syncthreads();
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
				int node = tree_find(W, s_new, P.dim);
				if (node < 0) {
					if (n >= W.cap) {
						status = PPX_ERR_CAPACITY;
						s_ctl = CTL_DONE;
					} else {
						node = n++;
						W.coords[3 * (size_t)node] = s_new[0];
						W.coords[3 * (size_t)node + 1] = s_new[1];
						W.coords[3 * (size_t)node + 2] = s_new[2];
						W.parent[node] = bestParent >= 0 ? bestParent : s_nearIdx[0];
						tree_insert_hash(W, node, P.dim);
					}
				}
				if (s_ctl != CTL_DONE) {
					W.cost[node] = bestCost; // (*newNode)->info.costFromGoal = bestCost, also for an existing node
					if (state_equal(s_new, s_goal, P.dim)) { // RRTStar::IsSolution, rrt_star.h:136-139
						solution = node;
						status = PPX_OK;
						s_ctl = CTL_DONE;
					}
				}
			}
			__syncthreads();
			if (s_ctl == CTL_DONE)
				break;
		}

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
	const int ctas = (int)std::min<int64_t>(ntrees, (int64_t)ctx->num_sms * 2);
	size_t off = 0;
	auto carve = [&](size_t bytes) {
		size_t o = off;
		off += (bytes + 255) & ~(size_t)255;
		return o;
	};
	const bool wantNodes = nodes || parents || costs;
	size_t oSlabs = carve((size_t)ctas * slab), oStarts = carve(ntrees * 24), oGoals = carve(ntrees * 24), oSeeds = carve(ntrees * 8);
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
	k_rrt<<<ctas, kRrtThreads, 0, ctx->stream>>>(P, map, ntrees, (const double*)(base + oStarts), (const double*)(base + oGoals),
		(const unsigned long long*)(base + oSeeds), base + oSlabs, slab, cap, hashCap, (ppx_rrt_result*)(base + oRes),
		wantNodes ? node_capacity : 0, nodes ? (double*)(base + oNodes) : nullptr, parents ? (int32_t*)(base + oPar) : nullptr,
		costs ? (double*)(base + oCost) : nullptr, ctx->d_queue);
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
