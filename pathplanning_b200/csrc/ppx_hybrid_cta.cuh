// The graph search of one query per CTA (included by ppx_hybrid.cu inside its anonymous namespace).
//
// A query is a chain of dependent expansions (algo/a_star.h:326-434): pop -> children -> frontier updates.  What is
// inherently sequential in an expansion is small -- the pop, the coin flip for the Reeds-Shepp shot (the RNG stream is
// consumed in expansion order, hybrid_a_star.cpp:81-82) and the ORDER in which the children meet the frontier / explored
// set (a_star.h:385-402).  What is expensive -- the ten constant-steer children of a node with their adaptive collision
// checks, Voronoi terms and heuristics -- is a pure function of the node and can be computed before the node is popped.
//
//   warp 0 (master)   pop (32-ary heap, top 3 levels in shared memory, warp-wide arg-min per level), RNG (state in
//                     shared memory), Reeds-Shepp shot when wanted, then the COMMIT of the <= 11 children by the
//                     lanes in parallel: key-table probes in one round trip, decisions in the reference's order
//                     (children of one parent that share a discrete cell are resolved through shuffles), node numbers
//                     and push sequence numbers by prefix sum, node records written by their lanes.
//   warps 1.. (workers) watch the top of the master's heap and generate the children of the best frontier nodes that
//                     have none yet (up to three nodes = 30 lanes per pass) into per-node records in global memory,
//                     claim by CAS, publish by release store.  The master generates a popped node's children itself
//                     only when no worker has claimed it yet.
//
// The results (expansion sequence, node numbering, costs) are those of the one-warp kernel k_hybrid_search, which the
// parity tests compare entry by entry with the reference.
#pragma once

// heap entries kept in shared memory: levels 0-2 of the 32-ary heap (latency mode) or levels 0-1 (throughput mode)
constexpr int kCtaHeapSmemBig = 1 + 32 + 1024, kCtaHeapSmemSmall = 1 + 32;
constexpr int kCtaMaxWarps = 8;

struct KidRec { // one constant-steer child of a node, written by whoever generated it
	double x, y, th; // continuousPose
	double cost; // transition cost
	double h; // heuristic of the child state
	double length; // action path length (after truncation)
	float last; // lastValidRatio
	int ki; // key-table index of the discrete pose; -1 outside the table
	int ok; // child exists
	int dir;
};

enum { GEN_PENDING = 0, GEN_CLAIMED = 1, GEN_READY = 2, GEN_NA = 3 };

struct CtaWs { // per CTA (query slot), carved from the slab
	SearchWs W;
	KidRec* kids; // [maxNodes][nCs] constant-steer children
	KidRec* rsKid; // [maxNodes] Reeds-Shepp child (nodes whose heuristic is below 10: the shot is certain, hybrid_a_star.cpp:82)
	int* gstate; // [maxNodes] state of kids
	int* rstate; // [maxNodes] state of rsKid (GEN_NA: the shot depends on the coin flip, made by the master on demand)
};

__host__ __device__ inline size_t cta_ws_bytes(int maxNodes, int tableSize, int nCs)
{
	return ws_bytes(maxNodes, tableSize) + align16((size_t)maxNodes * (nCs + 1) * sizeof(KidRec)) + 2 * align16((size_t)maxNodes * 4);
}

__device__ inline CtaWs cta_ws_carve(char* base, const WsLayout& L, int nCs)
{
	CtaWs C;
	C.W = ws_carve(base, L);
	char* p = base + ws_bytes(L.maxNodes, L.tableSize);
	C.kids = (KidRec*)p;
	p += (size_t)L.maxNodes * nCs * sizeof(KidRec);
	C.rsKid = (KidRec*)p;
	p += align16((size_t)L.maxNodes * sizeof(KidRec));
	C.gstate = (int*)p;
	p += align16((size_t)L.maxNodes * 4);
	C.rstate = (int*)p;
	return C;
}

__device__ __forceinline__ int ld_acquire_cta(const int* p)
{
	int v;
	asm volatile("ld.acquire.cta.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_release_cta(int* p, int v)
{
	asm volatile("st.release.cta.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// One constant-steer child of `node` (slot cslot) with its heuristic and table index, into the node's record.
__device__ __noinline__ void cta_gen_kid(const SearchConsts* consts, const float* wf, const Pose& goal, const CtaWs& C, int nCs,
	int node, int cslot)
{
	const HybridDev& P = consts->P;
	const MapDev& map = consts->map;
	const SearchWs& W = C.W;
	const Pose par = { W.pose[3 * node], W.pose[3 * node + 1], W.pose[3 * node + 2] };
	const Key3 key = discretize_pose(P, par);
	Child c;
	csteer_child(P, map, par, key, cslot >> 1, cslot & 1, c, &consts->fm);
	KidRec r;
	r.ok = c.ok ? 1 : 0;
	r.ki = -1;
	r.h = 0.0;
	if (c.ok) {
		r.h = hy_heuristic(P, map, wf, goal, c.pose);
		r.ki = key_table_index(W, c.key);
		r.x = c.pose.x;
		r.y = c.pose.y;
		r.th = c.pose.th;
		r.cost = c.cost;
		r.length = c.length;
		r.last = c.last;
		r.dir = c.dir;
	}
	C.kids[(size_t)node * nCs + cslot] = r;
}

// The Reeds-Shepp child of `node` (GetReedsSheppChild, hybrid_a_star.cpp:149-173): solved by the whole warp, built and
// checked by lane 0, into *out (global record or the master's shared one).
__device__ __noinline__ void cta_gen_rs(const SearchConsts* consts, const float* wf, const Pose& goal, const SearchWs& W, const Pose& par,
	KidRec* out)
{
	const HybridDev& P = consts->P;
	RSSolution sol;
	hy_rs_solve_warp(P, par, goal, sol);
	if ((threadIdx.x & 31) == 0) {
		Child c;
		hy_rs_child(P, consts->map, par, sol, c, &consts->fm);
		KidRec r;
		r.ok = c.ok ? 1 : 0;
		r.ki = -1;
		r.h = 0.0;
		if (c.ok) {
			r.h = hy_heuristic(P, consts->map, wf, goal, c.pose);
			r.ki = key_table_index(W, c.key);
			r.x = c.pose.x;
			r.y = c.pose.y;
			r.th = c.pose.th;
			r.cost = c.cost;
			r.length = c.length;
			r.last = c.last;
			r.dir = c.dir;
		}
		*out = r;
	}
	__syncwarp();
}

// Key table (discrete pose -> 0 absent / 1 explored / n + 2 frontier node n): 16-bit entries in shared memory when the
// table fits, else the 32-bit table of the slab.
template <bool kSmem>
struct CtaTable {
	uint16_t* s;
	uint32_t* g;
	__device__ __forceinline__ uint32_t get(int i) const { return kSmem ? (uint32_t)s[i] : g[i]; }
	__device__ __forceinline__ void set(int i, uint32_t v) const
	{
		if (kSmem)
			s[i] = (uint16_t)v;
		else
			g[i] = v;
	}
};

// ---- frontier: 32-ary min-heap on (totalCost asc, push sequence desc), first kCtaHeapSmem entries in shared memory ----
template <int kCtaHeapSmem>
struct CtaHeap {
	double* sf;
	uint32_t* ss;
	uint32_t* sn;
	double* gf;
	uint32_t* gs;
	uint32_t* gn;
	__device__ __forceinline__ double f(int i) const { return i < kCtaHeapSmem ? sf[i] : gf[i]; }
	__device__ __forceinline__ uint32_t seq(int i) const { return i < kCtaHeapSmem ? ss[i] : gs[i]; }
	__device__ __forceinline__ uint32_t node(int i) const { return i < kCtaHeapSmem ? sn[i] : gn[i]; }
	__device__ __forceinline__ void set(int i, double f_, uint32_t s_, uint32_t n_) const
	{
		if (i < kCtaHeapSmem) {
			sf[i] = f_;
			ss[i] = s_;
			sn[i] = n_;
		} else {
			gf[i] = f_;
			gs[i] = s_;
			gn[i] = n_;
		}
	}
};

// lane 0 only
template <int kHS>
__device__ __forceinline__ void cta_heap_push(const CtaHeap<kHS>& H, int& size, double f, uint32_t seq, uint32_t node)
{
	int i = size++;
	while (i > 0) {
		int p = (i - 1) >> 5;
		double pf = H.f(p);
		uint32_t ps = H.seq(p);
		if (!heap_before(f, seq, pf, ps))
			break;
		H.set(i, pf, ps, H.node(p));
		i = p;
	}
	H.set(i, f, seq, node);
}

// whole warp; returns the popped node (size > 0 on entry), warp-uniform
template <int kHS>
__device__ __forceinline__ int cta_heap_pop(const CtaHeap<kHS>& H, int& size, unsigned lane)
{
	const int top = (int)H.node(0);
	const int n = --size;
	if (n == 0)
		return top;
	const double f = H.f(n);
	const uint32_t seq = H.seq(n), node = H.node(n);
	int i = 0;
	while (true) {
		const int c0 = 32 * i + 1;
		if (c0 >= n)
			break;
		const int idx = c0 + (int)lane;
		const bool have = idx < n;
		double cf = 0.0;
		uint32_t cs = 0;
		if (have) {
			cf = H.f(idx);
			cs = H.seq(idx);
		}
		// arg-min over the lanes by (f asc, seq desc); f >= 0, so its bit pattern orders like the value
		const unsigned long long fb = (unsigned long long)__double_as_longlong(cf + 0.0);
		unsigned k = have ? (unsigned)(fb >> 32) : 0xffffffffu;
		unsigned m = __reduce_min_sync(0xffffffffu, k);
		bool cand = have && k == m;
		k = cand ? (unsigned)fb : 0xffffffffu;
		m = __reduce_min_sync(0xffffffffu, k);
		cand = cand && k == m;
		k = cand ? ~cs : 0xffffffffu;
		m = __reduce_min_sync(0xffffffffu, k);
		cand = cand && k == m;
		const int bl = __ffs(__ballot_sync(0xffffffffu, cand)) - 1;
		const double bf = __shfl_sync(0xffffffffu, cf, bl);
		const uint32_t bs = __shfl_sync(0xffffffffu, cs, bl);
		if (!heap_before(bf, bs, f, seq))
			break;
		if (lane == 0)
			H.set(i, bf, bs, H.node(c0 + bl));
		__syncwarp();
		i = c0 + bl;
	}
	if (lane == 0)
		H.set(i, f, seq, node);
	__syncwarp();
	return top;
}

template <int kWarps, bool kSmemTable, int kCtaHeapSmem>
__global__ void __launch_bounds__(kWarps * 32, kWarps == 2 ? 8 : 3) k_hybrid_search_cta(const SearchConsts* __restrict__ consts, int64_t nq,
	const double* __restrict__ starts, const double* __restrict__ goals, const unsigned long long* __restrict__ seeds,
	const int* __restrict__ wfSlot, const float* __restrict__ wfCost, WsLayout L, size_t slabBytes, char* __restrict__ slabs,
	ppx_hybrid_limits limits, ppx_hybrid_result* __restrict__ results, ppx_hybrid_path_state* __restrict__ paths,
	ppx_hybrid_log_entry* __restrict__ logs, unsigned long long* __restrict__ queue, const int* __restrict__ order,
	unsigned long long* __restrict__ stats /* nullptr: off; see launch_search */)
{
	// dynamic shared memory (kSmemTable): the key table (16-bit entries), then the "ready" bitmap -- two bits per node, set
	// by whoever finished the node's constant-steer children (bit 0) / Reeds-Shepp child (bit 1), so that the master
	// learns without a global round trip that the records can be loaded
	extern __shared__ uint16_t s_table[];
	uint32_t* const s_ready = kSmemTable ? (uint32_t*)(s_table + ((L.tableSize + 7) & ~7)) : nullptr;
	__shared__ double s_hf[kCtaHeapSmem];
	__shared__ uint32_t s_hs[kCtaHeapSmem];
	__shared__ uint32_t s_hn[kCtaHeapSmem];
	__shared__ unsigned long long s_mt[312];
	__shared__ KidRec s_rsKid;
	__shared__ long long s_q;
	__shared__ volatile int s_run, s_heapSize, s_epoch, s_numNodes;
	__shared__ int s_sweepGen, s_sweepRs; // latency mode: cursors of the workers' sweeps over all nodes

	const HybridDev& P = consts->P;
	const MapDev& map = consts->map;
	const unsigned lane = threadIdx.x & 31;
	const int wib = threadIdx.x >> 5;
	const int nCs = 2 * P.numDeltas;
	const int perPass = min(3, 32 / nCs);
	const CtaWs C = cta_ws_carve(slabs + (size_t)blockIdx.x * slabBytes, L, nCs);
	const SearchWs& W = C.W;
	const size_t cells = (size_t)map.rows * map.cols;
	const CtaHeap<kCtaHeapSmem> H { s_hf, s_hs, s_hn, W.heapF, W.heapSeq, W.heapNode };
	const CtaTable<kSmemTable> T { s_table, W.table };

	while (true) {
		if (threadIdx.x == 0) {
			long long q = (long long)atomicAdd(queue, 1ull);
			if (q < nq && order)
				q = order[q]; // longest-expected-first schedule (k_query_order); results do not depend on it
			s_q = q;
			s_run = 0;
			s_heapSize = 0;
			s_epoch = 0;
			s_numNodes = 0;
			s_sweepGen = 0;
			s_sweepRs = 0;
		}
		__syncthreads();
		const long long q = s_q;
		if (q >= nq)
			break;
		for (int i = threadIdx.x; i < L.tableSize; i += kWarps * 32)
			T.set(i, 0u);
		if (kSmemTable)
			for (int i = threadIdx.x; i < (L.maxNodes + 15) / 16; i += kWarps * 32)
				s_ready[i] = 0u;
		const Pose goal = { goals[3 * q], goals[3 * q + 1], goals[3 * q + 2] };
		const float* wf = wfCost + (size_t)(wfSlot ? wfSlot[q] : q) * cells;
		__syncthreads();

		if (wib != 0) {
			// ---------------------------------------------------------------- worker
			while (s_run == 0)
				__nanosleep(100);
			unsigned long long wPass = 0, wNodes = 0, wCycles = 0, wIdle = 0, wRs = 0;
			unsigned swMask = 0u, rsMask = 0u; // the sweeps' current chunks of 32 nodes: which still need work
			int swBase = 0, rsBase = 0;
			while (s_run == 1) {
				const int epoch = s_epoch;
				const int hs = s_heapSize;
				int cand = -1, candR = -1;
				float cf = 0.0f;
				if ((int)lane < hs) {
					cand = (int)((volatile uint32_t*)s_hn)[lane];
					cf = (float)((volatile double*)s_hf)[lane];
				}
				__threadfence_block();
				if (cand >= 0) {
					if (W.meta[cand] & 0x100u)
						cand = -1; // removed from the frontier
					else {
						if (ld_acquire_cta(C.rstate + cand) == GEN_PENDING)
							candR = cand;
						if (ld_acquire_cta(C.gstate + cand) != GEN_PENDING)
							cand = -1; // has its children already (or someone is at it)
					}
				}
				int picked[3] = { -1, -1, -1 };
				int np = 0;
				for (int tries = 0; tries < 6 && np < perPass; tries++) {
					unsigned key = cand >= 0 ? ((__float_as_uint(cf) & ~31u) | lane) : 0xffffffffu;
					unsigned m = __reduce_min_sync(0xffffffffu, key);
					if (m == 0xffffffffu)
						break;
					const int wl = (int)(m & 31u);
					const int node = __shfl_sync(0xffffffffu, cand, wl);
					int won = 0;
					if ((int)lane == wl) {
						won = atomicCAS(C.gstate + node, GEN_PENDING, GEN_CLAIMED) == GEN_PENDING;
						cand = -1;
					}
					won = __shfl_sync(0xffffffffu, won, wl);
					// the same node may sit in two slots of a heap that is being rearranged: drop the duplicates
					if (cand == node)
						cand = -1;
					if (won) {
						if (np == 0) picked[0] = node;
						else if (np == 1) picked[1] = node;
						else picked[2] = node;
						np++;
					}
				}
				if (np > 0) {
					const long long wt0 = stats ? clock64() : 0;
					const int t = (int)lane / nCs, cslot = (int)lane % nCs;
					if (t < np) {
						const int node = t == 0 ? picked[0] : (t == 1 ? picked[1] : picked[2]);
						cta_gen_kid(consts, wf, goal, C, nCs, node, cslot);
					}
					__syncwarp();
					__threadfence_block();
					if ((int)lane < np) {
						const int node = lane == 0 ? picked[0] : (lane == 1 ? picked[1] : picked[2]);
						st_release_cta(C.gstate + node, GEN_READY);
						if (kSmemTable)
							atomicOr(s_ready + (node >> 4), 1u << ((node & 15) * 2));
					}
					if (stats) {
						wPass++;
						wNodes += np;
						wCycles += (unsigned long long)(clock64() - wt0);
					}
					continue;
				}
				// no constant-steer work: one certain Reeds-Shepp shot, best frontier node first
				int rsNode = -1;
				for (int tries = 0; tries < 4 && rsNode < 0; tries++) {
					unsigned key = candR >= 0 ? ((__float_as_uint(cf) & ~31u) | lane) : 0xffffffffu;
					unsigned m = __reduce_min_sync(0xffffffffu, key);
					if (m == 0xffffffffu)
						break;
					const int wl = (int)(m & 31u);
					const int node = __shfl_sync(0xffffffffu, candR, wl);
					int won = 0;
					if ((int)lane == wl) {
						won = atomicCAS(C.rstate + node, GEN_PENDING, GEN_CLAIMED) == GEN_PENDING;
						candR = -1;
					}
					won = __shfl_sync(0xffffffffu, won, wl);
					if (candR == node)
						candR = -1;
					if (won)
						rsNode = node;
				}
				if (rsNode >= 0) {
					const Pose par = { W.pose[3 * rsNode], W.pose[3 * rsNode + 1], W.pose[3 * rsNode + 2] };
					cta_gen_rs(consts, wf, goal, W, par, C.rsKid + rsNode);
					__threadfence_block();
					if (lane == 0) {
						st_release_cta(C.rstate + rsNode, GEN_READY);
						if (kSmemTable)
							atomicOr(s_ready + (rsNode >> 4), 2u << ((rsNode & 15) * 2));
					}
					wRs++;
					continue;
				}
				if (kWarps >= 4) {
					// Latency mode (the GPU holds the whole batch): nothing near the top of the frontier needs work, so
					// sweep over ALL nodes in creation order, one pass per turn of this loop (the top of the frontier is
					// looked at again in between) -- a long query expands most of what it creates.
					if (swMask == 0u) {
						const int nn = s_numNodes;
						int base = -1, take = 0;
						if (lane == 0) {
							const int cur = *(volatile int*)&s_sweepGen;
							take = min(32, nn - cur);
							if (take > 0 && atomicCAS(&s_sweepGen, cur, cur + take) == cur)
								base = cur;
						}
						base = __shfl_sync(0xffffffffu, base, 0);
						take = __shfl_sync(0xffffffffu, take, 0);
						if (base >= 0) {
							__threadfence_block();
							const int mine = base + (int)lane;
							const bool need = (int)lane < take && !(W.meta[mine] & 0x100u) && ld_acquire_cta(C.gstate + mine) == GEN_PENDING;
							swMask = __ballot_sync(0xffffffffu, need);
							swBase = base;
						}
					}
					if (swMask) {
						int ids[3] = { -1, -1, -1 };
						int cnt = 0;
						for (int t = 0; t < perPass && swMask; t++) {
							const int b = __ffs(swMask) - 1;
							swMask &= swMask - 1;
							int won = 0;
							if (lane == 0)
								won = atomicCAS(C.gstate + swBase + b, GEN_PENDING, GEN_CLAIMED) == GEN_PENDING;
							won = __shfl_sync(0xffffffffu, won, 0);
							if (won) {
								if (cnt == 0) ids[0] = swBase + b;
								else if (cnt == 1) ids[1] = swBase + b;
								else ids[2] = swBase + b;
								cnt++;
							}
						}
						if (cnt > 0) {
							const int t = (int)lane / nCs, cslot = (int)lane % nCs;
							if (t < cnt)
								cta_gen_kid(consts, wf, goal, C, nCs, t == 0 ? ids[0] : (t == 1 ? ids[1] : ids[2]), cslot);
							__syncwarp();
							__threadfence_block();
							if ((int)lane < cnt) {
								const int node = lane == 0 ? ids[0] : (lane == 1 ? ids[1] : ids[2]);
								st_release_cta(C.gstate + node, GEN_READY);
								if (kSmemTable)
									atomicOr(s_ready + (node >> 4), 1u << ((node & 15) * 2));
							}
							wNodes += cnt;
							wPass++;
						}
						continue;
					}
					if (rsMask == 0u) {
						const int nn = s_numNodes;
						int base = -1, take = 0;
						if (lane == 0) {
							const int cur = *(volatile int*)&s_sweepRs;
							take = min(32, nn - cur);
							if (take > 0 && atomicCAS(&s_sweepRs, cur, cur + take) == cur)
								base = cur;
						}
						base = __shfl_sync(0xffffffffu, base, 0);
						take = __shfl_sync(0xffffffffu, take, 0);
						if (base >= 0) {
							__threadfence_block();
							const int mine = base + (int)lane;
							const bool need = (int)lane < take && !(W.meta[mine] & 0x100u) && ld_acquire_cta(C.rstate + mine) == GEN_PENDING;
							rsMask = __ballot_sync(0xffffffffu, need);
							rsBase = base;
						}
					}
					if (rsMask) {
						const int node = rsBase + __ffs(rsMask) - 1;
						rsMask &= rsMask - 1;
						int won = 0;
						if (lane == 0)
							won = atomicCAS(C.rstate + node, GEN_PENDING, GEN_CLAIMED) == GEN_PENDING;
						won = __shfl_sync(0xffffffffu, won, 0);
						if (won) {
							const Pose par = { W.pose[3 * node], W.pose[3 * node + 1], W.pose[3 * node + 2] };
							cta_gen_rs(consts, wf, goal, W, par, C.rsKid + node);
							__threadfence_block();
							if (lane == 0) {
								st_release_cta(C.rstate + node, GEN_READY);
								if (kSmemTable)
									atomicOr(s_ready + (node >> 4), 2u << ((node & 15) * 2));
							}
							wRs++;
						}
						continue;
					}
				}
				wIdle++;
				while (s_epoch == epoch && s_run == 1) // nothing to do until the master changes the frontier
					__nanosleep(100);
			}
			if (stats && lane == 0) {
				atomicAdd(stats + 16, wPass);
				atomicAdd(stats + 17, wNodes);
				atomicAdd(stats + 18, wCycles);
				atomicAdd(stats + 19, wIdle);
				atomicAdd(stats + 20, wRs);
			}
			__syncthreads(); // end of the query (pairs with the master's)
			continue;
		}

		// -------------------------------------------------------------------- master (warp 0)
		const Pose start = { starts[3 * q], starts[3 * q + 1], starts[3 * q + 2] };
		int mtIdx = 0;
		Mt64 rng { s_mt, &mtIdx };
		// warp-uniform search state (every lane keeps a copy; lane 0 does the scalar work)
		int numNodes = 0, heapSize = 0, status = PPX_SEARCH_RUNNING, solution = -1;
		int nExpanded = 0, nChildren = 0, nRsAttempts = 0, epoch = 0;
		uint32_t seqNext = 0;
		unsigned long long hash = 1469598103934665603ULL;
		{
			// AStar::InitializeSearch, a_star.h:350-364
			int ki = -1;
			if (lane == 0) {
				rng.seed(seeds ? seeds[q] : 0ull);
				const double hStart = hy_heuristic(P, map, wf, goal, start);
				const Key3 k = discretize_pose(P, start);
				ki = key_table_index(W, k);
				if (ki >= 0 && W.maxNodes > 0) {
					W.pose[0] = start.x;
					W.pose[1] = start.y;
					W.pose[2] = start.th;
					W.g[0] = 0.0;
					W.f[0] = 0.0;
					W.h[0] = hStart;
					W.parent[0] = -1;
					W.keyIdx[0] = ki;
					W.meta[0] = 2u;
					W.last[0] = 1.0f;
					W.length[0] = 0.0;
					C.gstate[0] = GEN_PENDING;
					C.rstate[0] = hStart < 10.0 ? GEN_PENDING : GEN_NA;
					T.set(ki, 2u);
					H.set(0, 0.0, 0u, 0u);
				}
			}
			ki = __shfl_sync(0xffffffffu, ki, 0);
			if (ki < 0 || W.maxNodes <= 0)
				status = PPX_ERR_CAPACITY;
			else {
				numNodes = 1;
				heapSize = 1;
				seqNext = 1;
			}
			if (lane == 0) {
				__threadfence_block();
				s_heapSize = heapSize;
				s_run = 1;
			}
			__syncwarp();
		}
		ppx_hybrid_log_entry* log = (logs && limits.max_log > 0) ? logs + (size_t)q * limits.max_log : nullptr;
		unsigned long long sc[8] = { 0, 0, 0, 0, 0, 0, 0, 0 }, sn[6] = { 0, 0, 0, 0, 0, 0 };
#define PPX_TICK(i) do { if (stats) { long long now_ = clock64(); sc[i] += (unsigned long long)(now_ - tick_); tick_ = now_; } } while (0)

		while (status == PPX_SEARCH_RUNNING) {
			long long tick_ = stats ? clock64() : 0;
			if (limits.max_expansions > 0 && nExpanded >= limits.max_expansions) {
				status = PPX_ERR_LIMIT;
				break;
			}
			// ---- pop (AStar::SearchPath loop, a_star.h:336-344, up to m_explored.Insert)
			int n = -1, pki = 0, st = 0, rst = 0;
			uint32_t pmeta = 0, rbits = 0;
			double hCost = 0, parentG = 0;
			Pose parent = { 0, 0, 0 };
			// the children's records are loaded in the same round trip as the node itself when the bitmap says they exist
			bool ok = false, kidsEarly = false, rsEarly = false;
			int ki = -1, dir = 0;
			double cx = 0, cy = 0, cth = 0, ccost = 0, ch = 0, clen = 0;
			float clast = 1.0f;
			while (heapSize > 0) {
				const int cand = cta_heap_pop(H, heapSize, lane);
				if (kSmemTable)
					rbits = (((volatile uint32_t*)s_ready)[cand >> 4] >> ((cand & 15) * 2)) & 3u;
				__threadfence_block();
				// everything the expansion reads, issued together
				pmeta = W.meta[cand];
				parent = { W.pose[3 * cand], W.pose[3 * cand + 1], W.pose[3 * cand + 2] };
				pki = W.keyIdx[cand];
				hCost = W.h[cand];
				parentG = W.g[cand];
				kidsEarly = (rbits & 1u) != 0u;
				rsEarly = (rbits & 2u) != 0u;
				const KidRec* r = nullptr;
				if (kidsEarly && (int)lane < nCs)
					r = C.kids + ((size_t)cand * nCs + lane);
				if (rsEarly && (int)lane == nCs)
					r = C.rsKid + cand;
				int rok = 0;
				if (r) {
					rok = r->ok; ki = r->ki; dir = r->dir;
					cx = r->x; cy = r->y; cth = r->th; ccost = r->cost; ch = r->h; clen = r->length; clast = r->last;
				}
				st = kidsEarly ? GEN_READY : ld_acquire_cta(C.gstate + cand);
				rst = rsEarly ? GEN_READY : ld_acquire_cta(C.rstate + cand);
				ok = rok != 0;
				if (pmeta & 0x100u)
					continue; // removed from the frontier by ProcessPossibleShortcut
				n = cand;
				break;
			}
			if (lane == 0) {
				s_heapSize = heapSize;
				s_epoch = ++epoch;
			}
			if (n < 0) {
				status = -1; // Status::Failure: frontier exhausted
				break;
			}
			if (identical_poses(parent, goal)) { // GraphSearch::IsSolution, hybrid_a_star.h:193-196
				solution = n;
				status = PPX_OK;
				break;
			}
			auto loadRec = [&](const KidRec* r) {
				ok = r->ok != 0;
				if (ok) {
					ki = r->ki; dir = r->dir;
					cx = r->x; cy = r->y; cth = r->th; ccost = r->cost; ch = r->h; clen = r->length; clast = r->last;
				}
			};
			PPX_TICK(0); // pop
			nExpanded++;
			hash = (hash ^ (unsigned long long)(uint32_t)pki) * 1099511628211ULL;
			int want = 0;
			if (lane == 0) {
				T.set(pki, 1u); // explored (and no longer in the frontier)
				if (log && nExpanded <= limits.max_log) {
					const Key3 pk = discretize_pose(P, parent);
					ppx_hybrid_log_entry& e = log[nExpanded - 1];
					e.key[0] = pk.x;
					e.key[1] = pk.y;
					e.key[2] = pk.k;
					e.direction = (int)(pmeta & 3u);
					e.pose[0] = parent.x;
					e.pose[1] = parent.y;
					e.pose[2] = parent.th;
				}
				want = rs_shot_wanted(rng, hCost) ? 1 : 0;
			}
			want = __shfl_sync(0xffffffffu, want, 0);
			nRsAttempts += want;
			PPX_TICK(1); // bookkeeping + RNG
			// ---- the children of n: from a worker, or generated here when no worker has taken the node yet
			if (stats)
				sn[st == GEN_READY ? 0 : (st == GEN_CLAIMED ? 1 : 2)]++;
			if (st == GEN_PENDING) {
				int won = 0;
				if (lane == 0)
					won = atomicCAS(C.gstate + n, GEN_PENDING, GEN_CLAIMED) == GEN_PENDING;
				won = __shfl_sync(0xffffffffu, won, 0);
				if (won) {
					if ((int)lane < nCs)
						cta_gen_kid(consts, wf, goal, C, nCs, n, (int)lane);
					__syncwarp();
					st = GEN_READY;
				} else
					st = GEN_CLAIMED;
			}
			PPX_TICK(2); // own generation
			// ---- the Reeds-Shepp shot (hybrid_a_star.cpp:81-88): from a worker when it was certain (h < 10), else made here
			const KidRec* rsRec = nullptr;
			if (want) {
				bool own = rst == GEN_NA;
				if (rst == GEN_PENDING) {
					int won = 0;
					if (lane == 0)
						won = atomicCAS(C.rstate + n, GEN_PENDING, GEN_CLAIMED) == GEN_PENDING;
					own = __shfl_sync(0xffffffffu, won, 0) != 0;
				}
				if (stats)
					sn[own ? 5 : (rst == GEN_READY ? 3 : 4)]++;
				if (own) {
					cta_gen_rs(consts, wf, goal, W, parent, &s_rsKid);
					rsRec = &s_rsKid;
				} else {
					while (rst != GEN_READY) {
						__nanosleep(40);
						rst = ld_acquire_cta(C.rstate + n);
					}
					rsRec = C.rsKid + n;
				}
			}
			PPX_TICK(3); // Reeds-Shepp shot
			while (st != GEN_READY) {
				__nanosleep(40);
				st = ld_acquire_cta(C.gstate + n);
			}
			PPX_TICK(4); // waiting for a worker
			// ---- commit: AStar::Expand's loop over the children (a_star.h:385-402) with ProcessPossibleShortcut
			// (hybrid_a_star.h:199-205), child `lane` on lane `lane`
			if (!kidsEarly && (int)lane < nCs)
				loadRec(C.kids + ((size_t)n * nCs + lane));
			if ((int)lane == nCs) {
				if (want && !rsEarly)
					loadRec(rsRec);
				if (!want)
					ok = false;
			}
			const int kind = (int)lane == nCs ? 1 : 0;
			const unsigned okMask = __ballot_sync(0xffffffffu, ok);
			nChildren += __popc(okMask);
			if (__any_sync(0xffffffffu, ok && ki < 0)) {
				status = PPX_ERR_CAPACITY;
				break;
			}
			const double g = parentG + ccost;
			const double f = g + ch;
			// current holder of the child's discrete cell: 0 none, 1 explored, fn + 2 frontier node fn,
			// 0x80000000 | j = the node child j of this very expansion has just pushed
			uint32_t curE = ok ? T.get(ki) : 1u;
			double hx = 0, hy = 0, hth = 0, hf = 0;
			if (ok && curE >= 2u) {
				const int fn = (int)(curE - 2u);
				hx = W.pose[3 * fn];
				hy = W.pose[3 * fn + 1];
				hth = W.pose[3 * fn + 2];
				hf = W.f[fn];
			}
			PPX_TICK(5); // commit: loads (records, table, holders)
			const unsigned grp = __match_any_sync(0xffffffffu, ok ? ki : -1 - (int)lane);
			bool push;
			{
				const Pose hp = { hx, hy, hth }, cp = { cx, cy, cth };
				push = ok && (curE == 0u || (curE >= 2u && identical_poses(hp, cp) && hf > f));
			}
			// children that share a cell meet the frontier one after the other, in child order: re-decide those
			unsigned shared = __ballot_sync(0xffffffffu, ok && (grp & (grp - 1)) != 0u);
			while (shared) {
				const int j = __ffs(shared) - 1;
				shared &= shared - 1;
				if ((int)lane == j) {
					const Pose hp = { hx, hy, hth }, cp = { cx, cy, cth };
					push = curE == 0u || (curE >= 2u && identical_poses(hp, cp) && hf > f);
				}
				if (!__shfl_sync(0xffffffffu, push ? 1 : 0, j))
					continue;
				const double jx = __shfl_sync(0xffffffffu, cx, j), jy = __shfl_sync(0xffffffffu, cy, j);
				const double jth = __shfl_sync(0xffffffffu, cth, j), jf = __shfl_sync(0xffffffffu, f, j);
				if ((int)lane > j && ((grp >> j) & 1u)) { // later members of j's group now find j's node in the cell
					curE = 0x80000000u | (uint32_t)j;
					hx = jx; hy = jy; hth = jth; hf = jf;
				}
			}
			const unsigned pushMask = __ballot_sync(0xffffffffu, push);
			const int nPush = __popc(pushMask);
			PPX_TICK(7); // commit: decisions
			if (numNodes + nPush > W.maxNodes) {
				status = PPX_ERR_CAPACITY;
				break;
			}
			// a pushing child removes the node that held its cell: a frontier node, or an earlier sibling
			const unsigned killed = __reduce_or_sync(0xffffffffu, (push && (curE & 0x80000000u)) ? (1u << (curE & 31u)) : 0u);
			const bool dead = (killed >> lane) & 1u;
			const int rank = __popc(pushMask & ((1u << lane) - 1u));
			const int node = numNodes + rank;
			if (push) {
				W.pose[3 * node] = cx;
				W.pose[3 * node + 1] = cy;
				W.pose[3 * node + 2] = cth;
				W.g[node] = g;
				W.f[node] = f;
				W.h[node] = ch;
				W.parent[node] = n;
				W.keyIdx[node] = ki;
				W.meta[node] = (uint32_t)(dir & 3) | ((uint32_t)(kind & 1) << 2) | ((uint32_t)(lane & 31u) << 3) | (dead ? 0x100u : 0u);
				W.last[node] = clast;
				W.length[node] = clen;
				C.gstate[node] = GEN_PENDING;
				C.rstate[node] = ch < 10.0 ? GEN_PENDING : GEN_NA; // rs_shot_wanted: certain below 10
				if (!dead)
					T.set(ki, (uint32_t)node + 2u);
				if (curE >= 2u && !(curE & 0x80000000u))
					W.meta[curE - 2u] |= 0x100u; // m_frontier.Remove(frontierNode)
			}
			__syncwarp();
			__threadfence_block(); // node records before the heap entries the workers read
			unsigned todo = pushMask & ~killed;
			while (todo) {
				const int b = __ffs(todo) - 1;
				todo &= todo - 1;
				const double fb = __shfl_sync(0xffffffffu, f, b);
				const int nb = __shfl_sync(0xffffffffu, node, b);
				const int rb = __shfl_sync(0xffffffffu, rank, b);
				if (lane == 0)
					cta_heap_push(H, heapSize, fb, seqNext + (uint32_t)rb, (uint32_t)nb);
				heapSize = __shfl_sync(0xffffffffu, heapSize, 0);
			}
			numNodes += nPush;
			seqNext += (uint32_t)nPush;
			if (lane == 0) {
				s_numNodes = numNodes;
				s_heapSize = heapSize;
				s_epoch = ++epoch;
			}
			__syncwarp();
			PPX_TICK(6); // commit: node records + heap pushes
		}
#undef PPX_TICK
		if (stats && lane == 0) {
			for (int i = 0; i < 8; i++)
				atomicAdd(stats + i, sc[i]);
			atomicAdd(stats + 8, sn[0]);
			atomicAdd(stats + 9, sn[1]);
			atomicAdd(stats + 10, sn[2]);
			atomicAdd(stats + 11, (unsigned long long)nExpanded);
			atomicAdd(stats + 12, sn[3]);
			atomicAdd(stats + 13, sn[4]);
			atomicAdd(stats + 14, sn[5]);
		}

		if (lane == 0)
			s_run = 2; // the workers leave their loop after the pass they are in
		__syncwarp();
		if (lane == 0) {
			ppx_hybrid_result r;
			r.status = status == PPX_SEARCH_RUNNING ? -1 : status;
			r.n_expanded = nExpanded;
			r.n_nodes = numNodes;
			r.n_children = nChildren;
			r.n_rs_attempts = nRsAttempts;
			r.path_len = 0;
			r.cost = PPX_INF;
			r.expanded_hash = hash;
			if (status == PPX_OK && solution >= 0) {
				r.cost = W.g[solution]; // GetOptimalCost, a_star.h:300-305
				int depth = 0;
				for (int m = solution; m >= 0; m = W.parent[m])
					depth++;
				r.path_len = depth;
				if (paths && limits.max_path > 0) {
					ppx_hybrid_path_state* out = paths + (size_t)q * limits.max_path;
					int pos = depth - 1;
					for (int m = solution; m >= 0; m = W.parent[m], pos--) {
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
		__syncthreads(); // end of the query (pairs with the workers')
	}
}
