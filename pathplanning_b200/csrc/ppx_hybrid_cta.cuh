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

constexpr int kCtaHeapSmem = 1 + 32 + 1024; // heap entries kept in shared memory (levels 0-2 of the 32-ary heap)
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

enum { GEN_PENDING = 0, GEN_CLAIMED = 1, GEN_READY = 2 };

struct CtaWs { // per CTA (query slot), carved from the slab
	SearchWs W;
	KidRec* kids; // [maxNodes][nCs]
	int* gstate; // [maxNodes]
};

__host__ __device__ inline size_t cta_ws_bytes(int maxNodes, int tableSize, int nCs)
{
	return ws_bytes(maxNodes, tableSize) + align16((size_t)maxNodes * nCs * sizeof(KidRec)) + align16((size_t)maxNodes * 4);
}

__device__ inline CtaWs cta_ws_carve(char* base, const WsLayout& L, int nCs)
{
	CtaWs C;
	C.W = ws_carve(base, L);
	char* p = base + ws_bytes(L.maxNodes, L.tableSize);
	C.kids = (KidRec*)p;
	p += align16((size_t)L.maxNodes * nCs * sizeof(KidRec));
	C.gstate = (int*)p;
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

// ---- frontier: 32-ary min-heap on (totalCost asc, push sequence desc), first kCtaHeapSmem entries in shared memory ----
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
__device__ __forceinline__ void cta_heap_push(const CtaHeap& H, int& size, double f, uint32_t seq, uint32_t node)
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
__device__ __forceinline__ int cta_heap_pop(const CtaHeap& H, int& size, unsigned lane)
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

template <int kWarps>
__global__ void __launch_bounds__(kWarps * 32) k_hybrid_search_cta(const SearchConsts* __restrict__ consts, int64_t nq,
	const double* __restrict__ starts, const double* __restrict__ goals, const unsigned long long* __restrict__ seeds,
	const int* __restrict__ wfSlot, const float* __restrict__ wfCost, WsLayout L, size_t slabBytes, char* __restrict__ slabs,
	ppx_hybrid_limits limits, ppx_hybrid_result* __restrict__ results, ppx_hybrid_path_state* __restrict__ paths,
	ppx_hybrid_log_entry* __restrict__ logs, unsigned long long* __restrict__ queue, const int* __restrict__ order)
{
	__shared__ double s_hf[kCtaHeapSmem];
	__shared__ uint32_t s_hs[kCtaHeapSmem];
	__shared__ uint32_t s_hn[kCtaHeapSmem];
	__shared__ unsigned long long s_mt[312];
	__shared__ Child s_rs;
	__shared__ double s_rsh;
	__shared__ long long s_q;
	__shared__ volatile int s_run, s_heapSize;

	const HybridDev& P = consts->P;
	const MapDev& map = consts->map;
	const unsigned lane = threadIdx.x & 31;
	const int wib = threadIdx.x >> 5;
	const int nCs = 2 * P.numDeltas;
	const int perPass = min(3, 32 / nCs);
	const CtaWs C = cta_ws_carve(slabs + (size_t)blockIdx.x * slabBytes, L, nCs);
	const SearchWs& W = C.W;
	const size_t cells = (size_t)map.rows * map.cols;
	const CtaHeap H { s_hf, s_hs, s_hn, W.heapF, W.heapSeq, W.heapNode };

	while (true) {
		if (threadIdx.x == 0) {
			long long q = (long long)atomicAdd(queue, 1ull);
			if (q < nq && order)
				q = order[q]; // longest-expected-first schedule (k_query_order); results do not depend on it
			s_q = q;
			s_run = 0;
			s_heapSize = 0;
		}
		__syncthreads();
		const long long q = s_q;
		if (q >= nq)
			break;
		for (int i = threadIdx.x; i < L.tableSize; i += kWarps * 32)
			W.table[i] = 0u;
		const Pose goal = { goals[3 * q], goals[3 * q + 1], goals[3 * q + 2] };
		const float* wf = wfCost + (size_t)(wfSlot ? wfSlot[q] : q) * cells;
		__syncthreads();

		if (wib != 0) {
			// ---------------------------------------------------------------- worker
			while (s_run == 0)
				__nanosleep(100);
			while (s_run == 1) {
				const int hs = s_heapSize;
				int cand = -1;
				float cf = 0.0f;
				if ((int)lane < hs) {
					cand = (int)((volatile uint32_t*)s_hn)[lane];
					cf = (float)((volatile double*)s_hf)[lane];
				}
				__threadfence_block();
				if (cand >= 0 && (ld_acquire_cta(C.gstate + cand) != GEN_PENDING || (W.meta[cand] & 0x100u)))
					cand = -1; // has children already, or was removed from the frontier
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
				if (np == 0) {
					__nanosleep(200);
					continue;
				}
				const int t = (int)lane / nCs, cslot = (int)lane % nCs;
				if (t < np) {
					const int node = t == 0 ? picked[0] : (t == 1 ? picked[1] : picked[2]);
					cta_gen_kid(consts, wf, goal, C, nCs, node, cslot);
				}
				__syncwarp();
				__threadfence_block();
				if ((int)lane < np)
					st_release_cta(C.gstate + (lane == 0 ? picked[0] : (lane == 1 ? picked[1] : picked[2])), GEN_READY);
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
		int nExpanded = 0, nChildren = 0, nRsAttempts = 0;
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
					W.table[ki] = 2u;
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

		while (status == PPX_SEARCH_RUNNING) {
			if (limits.max_expansions > 0 && nExpanded >= limits.max_expansions) {
				status = PPX_ERR_LIMIT;
				break;
			}
			// ---- pop (AStar::SearchPath loop, a_star.h:336-344, up to m_explored.Insert)
			int n = -1;
			Pose parent = { 0, 0, 0 };
			while (heapSize > 0) {
				const int cand = cta_heap_pop(H, heapSize, lane);
				if (lane == 0)
					s_heapSize = heapSize;
				if (W.meta[cand] & 0x100u)
					continue; // removed from the frontier by ProcessPossibleShortcut
				parent = { W.pose[3 * cand], W.pose[3 * cand + 1], W.pose[3 * cand + 2] };
				n = cand;
				break;
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
			const int pki = W.keyIdx[n];
			const double hCost = W.h[n];
			const double parentG = W.g[n];
			nExpanded++;
			hash = (hash ^ (unsigned long long)(uint32_t)pki) * 1099511628211ULL;
			int want = 0;
			if (lane == 0) {
				W.table[pki] = 1u; // explored (and no longer in the frontier)
				if (log && nExpanded <= limits.max_log) {
					const Key3 pk = discretize_pose(P, parent);
					ppx_hybrid_log_entry& e = log[nExpanded - 1];
					e.key[0] = pk.x;
					e.key[1] = pk.y;
					e.key[2] = pk.k;
					e.direction = (int)(W.meta[n] & 3u);
					e.pose[0] = parent.x;
					e.pose[1] = parent.y;
					e.pose[2] = parent.th;
				}
				want = rs_shot_wanted(rng, hCost) ? 1 : 0;
			}
			want = __shfl_sync(0xffffffffu, want, 0);
			nRsAttempts += want;
			// ---- the children of n: from a worker, or generated here when no worker has taken the node yet
			int st = ld_acquire_cta(C.gstate + n);
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
				}
			}
			// ---- the Reeds-Shepp shot (hybrid_a_star.cpp:81-88) overlaps a worker that is still generating
			if (want) {
				RSSolution sol;
				hy_rs_solve_warp(P, parent, goal, sol);
				if (lane == 0) {
					Child c;
					hy_rs_child(P, map, parent, sol, c, &consts->fm);
					s_rsh = c.ok ? hy_heuristic(P, map, wf, goal, c.pose) : 0.0;
					s_rs = c;
				}
				__syncwarp();
			}
			while (st != GEN_READY) {
				__nanosleep(40);
				st = ld_acquire_cta(C.gstate + n);
			}
			// ---- commit: AStar::Expand's loop over the children (a_star.h:385-402) with ProcessPossibleShortcut
			// (hybrid_a_star.h:199-205), child `lane` on lane `lane`
			bool ok = false;
			int ki = -1, dir = 0, kind = 0;
			double cx = 0, cy = 0, cth = 0, ccost = 0, ch = 0, clen = 0;
			float clast = 1.0f;
			if ((int)lane < nCs) {
				const KidRec& r = C.kids[(size_t)n * nCs + lane];
				ok = r.ok != 0;
				if (ok) {
					ki = r.ki; dir = r.dir;
					cx = r.x; cy = r.y; cth = r.th; ccost = r.cost; ch = r.h; clen = r.length; clast = r.last;
				}
			} else if ((int)lane == nCs && want && s_rs.ok) {
				ok = true;
				kind = 1;
				ki = key_table_index(W, s_rs.key);
				dir = s_rs.dir;
				cx = s_rs.pose.x; cy = s_rs.pose.y; cth = s_rs.pose.th;
				ccost = s_rs.cost; ch = s_rsh; clen = s_rs.length; clast = s_rs.last;
			}
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
			uint32_t curE = ok ? W.table[ki] : 1u;
			double hx = 0, hy = 0, hth = 0, hf = 0;
			if (ok && curE >= 2u) {
				const int fn = (int)(curE - 2u);
				hx = W.pose[3 * fn];
				hy = W.pose[3 * fn + 1];
				hth = W.pose[3 * fn + 2];
				hf = W.f[fn];
			}
			const unsigned grp = __match_any_sync(0xffffffffu, ok ? ki : -1 - (int)lane);
			bool push;
			if (!__any_sync(0xffffffffu, (grp & (grp - 1)) != 0u)) {
				const Pose hp = { hx, hy, hth }, cp = { cx, cy, cth };
				push = ok && (curE == 0u || (curE >= 2u && identical_poses(hp, cp) && hf > f));
			} else {
				// children that share a cell meet the frontier one after the other, in child order
				push = false;
				const int last = 31 - __clz(okMask);
				for (int j = 0; j <= last; j++) {
					bool pj = false;
					if ((int)lane == j && ok) {
						const Pose hp = { hx, hy, hth }, cp = { cx, cy, cth };
						pj = curE == 0u || (curE >= 2u && identical_poses(hp, cp) && hf > f);
						push = pj;
					}
					pj = __shfl_sync(0xffffffffu, pj ? 1 : 0, j) != 0;
					if (!pj)
						continue;
					const int kj = __shfl_sync(0xffffffffu, ki, j);
					const double jx = __shfl_sync(0xffffffffu, cx, j), jy = __shfl_sync(0xffffffffu, cy, j);
					const double jth = __shfl_sync(0xffffffffu, cth, j), jf = __shfl_sync(0xffffffffu, f, j);
					if ((int)lane > j && ok && ki == kj) {
						curE = 0x80000000u | (uint32_t)j;
						hx = jx; hy = jy; hth = jth; hf = jf;
					}
				}
			}
			const unsigned pushMask = __ballot_sync(0xffffffffu, push);
			const int nPush = __popc(pushMask);
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
				if (!dead)
					W.table[ki] = (uint32_t)node + 2u;
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
			if (lane == 0)
				s_heapSize = heapSize;
			__syncwarp();
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
