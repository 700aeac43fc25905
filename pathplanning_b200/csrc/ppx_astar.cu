// libppx.so: grid A* (SURVEY N4 / BASELINE config C1) -- AStarN2 = AStar<GridCellPosition, NullAction, ..., GraphSearch>
// (algo/a_star_n2.h:29-34, algo/a_star.h:326-434) with AStarStatePropagatorFcnN2 (algo/a_star_n2.cpp:12-28) for a batch
// of (start, goal) cell pairs on the context's occupancy grid, one warp per query.
//
// The reference takes the step cost and the heuristic as std::function callbacks, which cannot cross a C ABI; this
// entry point fixes both to the Euclidean cell distance, the pair interfaces/python/scripts/example_a_star_grid.py
// passes (config C1).  Everything else is the reference's: neighbour order of GridCellPosition::GetNeighbors
// (utils/grid.cpp:16-50), occupied neighbours and corner-cutting diagonals dropped, the Frontier's pop order
// ("smallest totalCost, most recently pushed among equals", utils/frontier.h) as a binary heap on (totalCost, -push
// sequence), Find / Contains on the cell as a dense table (0 absent, 1 explored, n+2 in the frontier as node n),
// ProcessPossibleShortcut as a dead flag on the replaced node.  Lanes 0..7 evaluate the neighbours of the popped cell,
// lane 0 commits them in order.
#include <algorithm>
#include "ppx_ctx.h"

namespace {

constexpr int kAstarWarps = 4;

struct AstarWs {
	uint32_t* table; // [cells]
	int* cell; // [maxNodes] row * cols + col
	int* parent;
	double* g;
	double* f;
	uint8_t* dead;
	double* heapF;
	uint32_t* heapSeq;
	uint32_t* heapNode;
};

struct AstarLayout {
	size_t oTable, oCell, oParent, oG, oF, oDead, oHeapF, oHeapSeq, oHeapNode, bytes;
	int maxNodes;
	size_t cells;
};

AstarLayout astar_layout(size_t cells, int maxNodes)
{
	AstarLayout L;
	size_t off = 0;
	auto carve = [&](size_t b) {
		size_t o = off;
		off += (b + 255) & ~(size_t)255;
		return o;
	};
	L.oTable = carve(cells * 4);
	L.oCell = carve((size_t)maxNodes * 4);
	L.oParent = carve((size_t)maxNodes * 4);
	L.oG = carve((size_t)maxNodes * 8);
	L.oF = carve((size_t)maxNodes * 8);
	L.oDead = carve((size_t)maxNodes);
	L.oHeapF = carve((size_t)maxNodes * 8);
	L.oHeapSeq = carve((size_t)maxNodes * 4);
	L.oHeapNode = carve((size_t)maxNodes * 4);
	L.bytes = off;
	L.maxNodes = maxNodes;
	L.cells = cells;
	return L;
}

__device__ __forceinline__ bool before(double fa, uint32_t sa, double fb, uint32_t sb) { return fa < fb || (fa == fb && sa > sb); }

__device__ void heap_push(const AstarWs& W, int& heapSize, uint32_t& seqCounter, double f, int node)
{
	uint32_t seq = seqCounter++;
	int i = heapSize++;
	while (i > 0) {
		int p = (i - 1) >> 1;
		double pf = W.heapF[p];
		uint32_t ps = W.heapSeq[p];
		if (!before(f, seq, pf, ps))
			break;
		W.heapF[i] = pf;
		W.heapSeq[i] = ps;
		W.heapNode[i] = W.heapNode[p];
		i = p;
	}
	W.heapF[i] = f;
	W.heapSeq[i] = seq;
	W.heapNode[i] = (uint32_t)node;
}

__device__ int heap_pop(const AstarWs& W, int& heapSize)
{
	int top = (int)W.heapNode[0];
	int n = --heapSize;
	if (n > 0) {
		double f = W.heapF[n];
		uint32_t seq = W.heapSeq[n], node = W.heapNode[n];
		int i = 0;
		while (true) {
			int l = 2 * i + 1, r = l + 1;
			if (l >= n)
				break;
			int c = l;
			double cf = W.heapF[l];
			uint32_t cs = W.heapSeq[l];
			if (r < n) {
				double rf = W.heapF[r];
				uint32_t rs = W.heapSeq[r];
				if (before(rf, rs, cf, cs)) {
					c = r;
					cf = rf;
					cs = rs;
				}
			}
			if (!before(cf, cs, f, seq))
				break;
			W.heapF[i] = cf;
			W.heapSeq[i] = cs;
			W.heapNode[i] = W.heapNode[c];
			i = c;
		}
		W.heapF[i] = f;
		W.heapSeq[i] = seq;
		W.heapNode[i] = node;
	}
	return top;
}

__device__ __forceinline__ double euclid(int r0, int c0, int r1, int c1)
{
	double dr = (double)(r0 - r1), dc = (double)(c0 - c1);
	return sqrt(dr * dr + dc * dc);
}

// The reference takes the step cost and the heuristic as std::function callbacks (a_star_n2.h:14-27; Python lambdas in
// interfaces/python/src/pyplanning.cpp:140-154).  A callback cannot cross the C ABI into a kernel; the device offers the
// table of functors those callbacks are in practice: kind 0 Euclidean, 1 Manhattan, 2 octile, 3 (cost) Euclidean times
// the per-cell cost of the neighbour / (heuristic) zero.
struct AstarFn {
	int costKind, heurKind;
	const float* cellCost; // [rows * cols], cost kind 3
	int cols;
};

__device__ __forceinline__ double grid_metric(int kind, int r0, int c0, int r1, int c1)
{
	const double dr = fabs((double)(r0 - r1)), dc = fabs((double)(c0 - c1));
	if (kind == 1)
		return dr + dc;
	if (kind == 2)
		return fmax(dr, dc) + (sqrt(2.0) - 1.0) * fmin(dr, dc);
	return sqrt(dr * dr + dc * dc);
}
__device__ __forceinline__ double step_cost(const AstarFn& F, int r, int c, int nr, int nc)
{
	if (F.costKind == 3)
		return grid_metric(0, r, c, nr, nc) * (double)F.cellCost[(size_t)nr * F.cols + nc];
	return grid_metric(F.costKind, r, c, nr, nc);
}
__device__ __forceinline__ double heuristic(const AstarFn& F, int r, int c, int gr, int gc)
{
	return F.heurKind == 3 ? 0.0 : grid_metric(F.heurKind, r, c, gr, gc);
}

__global__ void __launch_bounds__(kAstarWarps * 32) k_astar_n2(AstarFn F, const uint8_t* __restrict__ occ8, int rows, int cols, int64_t nq,
	const int32_t* __restrict__ starts, const int32_t* __restrict__ goals, AstarLayout L, char* __restrict__ slabs, int maxPath,
	ppx_astar_result* __restrict__ results, int32_t* __restrict__ paths, unsigned long long* __restrict__ queue)
{
	const unsigned lane = threadIdx.x & 31;
	const size_t gwarp = (size_t)blockIdx.x * kAstarWarps + (threadIdx.x >> 5);
	char* slab = slabs + gwarp * L.bytes;
	AstarWs W;
	W.table = (uint32_t*)(slab + L.oTable);
	W.cell = (int*)(slab + L.oCell);
	W.parent = (int*)(slab + L.oParent);
	W.g = (double*)(slab + L.oG);
	W.f = (double*)(slab + L.oF);
	W.dead = (uint8_t*)(slab + L.oDead);
	W.heapF = (double*)(slab + L.oHeapF);
	W.heapSeq = (uint32_t*)(slab + L.oHeapSeq);
	W.heapNode = (uint32_t*)(slab + L.oHeapNode);

	while (true) {
		long long q = 0;
		if (lane == 0)
			q = (long long)atomicAdd(queue, 1ull);
		q = __shfl_sync(0xffffffffu, q, 0);
		if (q >= nq)
			break;
		for (size_t i = lane; i < L.cells; i += 32)
			W.table[i] = 0u;
		__syncwarp();
		const int sr = starts[2 * q], sc = starts[2 * q + 1], gr = goals[2 * q], gc = goals[2 * q + 1];
		int heapSize = 0, numNodes = 0, nExplored = 0, status = 1, solution = -1; // lane 0's copies are the state
		uint32_t seqCounter = 0;
		const bool okEnds = sr >= 0 && sr < rows && sc >= 0 && sc < cols && gr >= 0 && gr < rows && gc >= 0 && gc < cols;
		if (!okEnds)
			status = PPX_ERR_INVALID;
		if (lane == 0 && okEnds) { // InitializeSearch, a_star.h:350-364: root in the frontier AND in the explored set
			W.cell[0] = sr * cols + sc;
			W.parent[0] = -1;
			W.g[0] = 0.0;
			W.f[0] = 0.0;
			W.dead[0] = 0;
			numNodes = 1;
			heap_push(W, heapSize, seqCounter, 0.0, 0);
			W.table[sr * cols + sc] = 2u; // node 0 in the frontier
			nExplored = 1;
		}
		__syncwarp();
		while (status == 1) {
			int n = -1;
			if (lane == 0) {
				while (heapSize > 0) {
					int m = heap_pop(W, heapSize);
					if (W.dead[m])
						continue; // m_frontier.Remove
					n = m;
					break;
				}
				if (n < 0)
					status = -1; // Status::Failure: frontier exhausted
				else if (W.cell[n] == gr * cols + gc) { // IsSolution
					solution = n;
					status = 0;
					n = -1;
				} else {
					if (n != 0)
						nExplored++; // the root was inserted by InitializeSearch already
					W.table[W.cell[n]] = 1u; // explored, no longer in the frontier
				}
			}
			n = __shfl_sync(0xffffffffu, n, 0);
			status = __shfl_sync(0xffffffffu, status, 0);
			if (n < 0)
				break;
			const int cell = W.cell[n];
			const int r = cell / cols, c = cell % cols;
			// GetNeighbors order (utils/grid.cpp:16-50): left column (same, bottom, top), right column, bottom, top
			int nr = r, nc = c;
			bool exists = false;
			if (lane < 8) {
				const int dr[8] = { 0, -1, 1, 0, -1, 1, -1, 1 };
				const int dc[8] = { -1, -1, -1, 1, 1, 1, 0, 0 };
				nr = r + dr[lane];
				nc = c + dc[lane];
				exists = nr >= 0 && nr < rows && nc >= 0 && nc < cols;
				if (exists && occ8[(size_t)nr * cols + nc])
					exists = false; // IsOccupied(n)
				if (exists && nr != r && nc != c && occ8[(size_t)nr * cols + c] && occ8[(size_t)r * cols + nc])
					exists = false; // diagonal between two occupied cells
			}
			// the eight table entries -- and the f of the frontier nodes they name -- are read by the eight lanes at once: the
			// neighbours are eight different cells, so no commit below can change another neighbour's entry, and read one
			// after the other by lane 0 they were up to sixteen dependent L2 round trips per expansion
			double step = 0.0, h = 0.0, oldF = 0.0;
			uint32_t entry = 0;
			if (exists) {
				step = step_cost(F, r, c, nr, nc);
				h = heuristic(F, nr, nc, gr, gc);
				entry = W.table[(size_t)nr * cols + nc];
				if (entry >= 2u)
					oldF = W.f[entry - 2u];
			}
			const double parentG = W.g[n];
			for (int j = 0; j < 8; j++) { // commit in order (lane 0 holds the search state)
				const bool ex = __shfl_sync(0xffffffffu, (int)exists, j) != 0;
				const double st = __shfl_sync(0xffffffffu, step, j), hj = __shfl_sync(0xffffffffu, h, j);
				const int cr = __shfl_sync(0xffffffffu, nr, j), cc = __shfl_sync(0xffffffffu, nc, j);
				const uint32_t e = __shfl_sync(0xffffffffu, entry, j);
				const double ef = __shfl_sync(0xffffffffu, oldF, j);
				if (!ex || lane != 0 || status != 1)
					continue;
				const int ci = cr * cols + cc;
				const double g = parentG + st, f = g + hj;
				bool push = false;
				if (e == 0u)
					push = true;
				else if (e >= 2u) { // ProcessPossibleShortcut, a_star.h:412-421
					const int fn = (int)(e - 2u);
					if (ef > f) {
						W.dead[fn] = 1;
						push = true;
					}
				}
				if (push) {
					if (numNodes >= L.maxNodes) {
						status = PPX_ERR_CAPACITY;
						continue;
					}
					const int m = numNodes++;
					W.cell[m] = ci;
					W.parent[m] = n;
					W.g[m] = g;
					W.f[m] = f;
					W.dead[m] = 0;
					heap_push(W, heapSize, seqCounter, f, m);
					W.table[ci] = (uint32_t)m + 2u;
				}
			}
			__syncwarp();
			status = __shfl_sync(0xffffffffu, status, 0);
		}
		if (lane == 0) {
			ppx_astar_result res;
			res.status = status == 1 ? -1 : status;
			res.n_explored = nExplored;
			res.n_nodes = numNodes;
			res.path_len = 0;
			res.cost = PPX_INF;
			if (status == 0 && solution >= 0) {
				res.cost = W.g[solution]; // GetOptimalCost
				int depth = 0;
				for (int m = solution; m >= 0; m = W.parent[m])
					depth++;
				res.path_len = depth;
				if (paths && maxPath > 0) {
					int32_t* out = paths + (size_t)q * maxPath * 2;
					int pos = depth - 1;
					for (int m = solution; m >= 0; m = W.parent[m], pos--) {
						if (pos >= maxPath)
							continue;
						out[2 * pos] = W.cell[m] / cols;
						out[2 * pos + 1] = W.cell[m] % cols;
					}
				}
			}
			results[q] = res;
		}
		__syncwarp();
	}
}


// ---- BidirectionalAStar<GridCellPosition, ...> (algo/bidirectional_a_star.h:41-204): two A* searches (ExploredMap: state ->
// node) advancing in turn -- pop + Expand forward, FindIntersection, pop + Expand reverse, FindIntersection -- until the two
// frontier tops prove the best meeting point optimal (:161-172).  One warp per query; each direction has its own workspace.
// Table encoding per direction: 0 absent, 0x80000000 | n explored as node n, n + 2 in the frontier as node n.
struct BiDir {
	AstarWs W;
	int heapSize, live, numNodes, nExplored, solution;
	uint32_t seq;
	int gr, gc; // this direction's goal
};

__device__ int bi_clean_top(BiDir& D) // drops removed entries from the top; returns the top node or -1
{
	while (D.heapSize > 0) {
		int m = (int)D.W.heapNode[0];
		if (!D.W.dead[m])
			return m;
		heap_pop(D.W, D.heapSize);
	}
	return -1;
}

__global__ void __launch_bounds__(kAstarWarps * 32) k_bi_astar_n2(AstarFn F, const uint8_t* __restrict__ occ8, int rows, int cols, int64_t nq,
	const int32_t* __restrict__ starts, const int32_t* __restrict__ goals, AstarLayout L, char* __restrict__ slabs, int maxPath,
	ppx_astar_result* __restrict__ results, int32_t* __restrict__ paths, unsigned long long* __restrict__ queue)
{
	const unsigned lane = threadIdx.x & 31;
	const size_t gwarp = (size_t)blockIdx.x * kAstarWarps + (threadIdx.x >> 5);
	BiDir D[2];
	for (int d = 0; d < 2; d++) {
		char* slab = slabs + (2 * gwarp + d) * L.bytes;
		AstarWs& W = D[d].W;
		W.table = (uint32_t*)(slab + L.oTable);
		W.cell = (int*)(slab + L.oCell);
		W.parent = (int*)(slab + L.oParent);
		W.g = (double*)(slab + L.oG);
		W.f = (double*)(slab + L.oF);
		W.dead = (uint8_t*)(slab + L.oDead);
		W.heapF = (double*)(slab + L.oHeapF);
		W.heapSeq = (uint32_t*)(slab + L.oHeapSeq);
		W.heapNode = (uint32_t*)(slab + L.oHeapNode);
	}
	while (true) {
		long long q = 0;
		if (lane == 0)
			q = (long long)atomicAdd(queue, 1ull);
		q = __shfl_sync(0xffffffffu, q, 0);
		if (q >= nq)
			break;
		for (size_t i = lane; i < L.cells; i += 32) {
			D[0].W.table[i] = 0u;
			D[1].W.table[i] = 0u;
		}
		__syncwarp();
		const int sr = starts[2 * q], sc = starts[2 * q + 1], gr = goals[2 * q], gc = goals[2 * q + 1];
		const bool okEnds = sr >= 0 && sr < rows && sc >= 0 && sc < cols && gr >= 0 && gr < rows && gc >= 0 && gc < cols;
		int status = okEnds ? 1 : PPX_ERR_INVALID;
		double bestCost = PPX_INF, costOffset = 0.0;
		for (int d = 0; d < 2; d++) { // SetInitState / SetGoalState / InitializeSearch of both searches (:143-150)
			BiDir& B = D[d];
			B.heapSize = B.live = B.numNodes = B.nExplored = 0;
			B.solution = -1;
			B.seq = 0;
			B.gr = d == 0 ? gr : sr;
			B.gc = d == 0 ? gc : sc;
			if (lane == 0 && okEnds) {
				const int r0 = d == 0 ? sr : gr, c0 = d == 0 ? sc : gc;
				B.W.cell[0] = r0 * cols + c0;
				B.W.parent[0] = -1;
				B.W.g[0] = 0.0;
				B.W.f[0] = 0.0;
				B.W.dead[0] = 0;
				B.numNodes = 1;
				heap_push(B.W, B.heapSize, B.seq, 0.0, 0);
				B.live = 1;
				// the root is in the frontier AND in the explored map (a_star.h:359-360); a frontier entry whose state is also
				// explored cannot be told apart by one table word: the root is the only such node, flagged by index 0
				B.W.table[r0 * cols + c0] = 0x80000000u;
				B.nExplored = 1;
			}
		}
		if (okEnds) // :152: both heuristics evaluated at the planner's goal
			costOffset = heuristic(F, gr, gc, D[0].gr, D[0].gc) + heuristic(F, gr, gc, D[1].gr, D[1].gc);
		__syncwarp();
		while (status == 1) {
			int done = 0;
			if (lane == 0 && (D[0].live == 0 || D[1].live == 0))
				done = 1; // while (!fFrontier.Empty() && !rFrontier.Empty())
			done = __shfl_sync(0xffffffffu, done, 0);
			if (done) {
				status = -1;
				break;
			}
			for (int d = 0; d < 2; d++) {
				BiDir& A = D[d];
				BiDir& Bo = D[1 - d];
				// Pop
				int n = -1;
				if (lane == 0) {
					n = bi_clean_top(A);
					if (n >= 0) {
						heap_pop(A.W, A.heapSize);
						A.live--;
						// Expand: m_explored.Insert(node) -- unordered_map::insert keeps an existing entry (the root's own)
						const int cell = A.W.cell[n];
						if (!(A.W.table[cell] & 0x80000000u)) {
							A.W.table[cell] = 0x80000000u | (uint32_t)n;
							A.nExplored++;
						}
					}
				}
				n = __shfl_sync(0xffffffffu, n, 0);
				if (n < 0)
					continue; // cannot happen: both frontiers were non-empty at the top of the loop... the reverse one may have
					          // been emptied only by its own pops
				const int cell = A.W.cell[n];
				const int r = cell / cols, c = cell % cols;
				int nr = r, nc = c;
				bool exists = false;
				if (lane < 8) {
					const int dr[8] = { 0, -1, 1, 0, -1, 1, -1, 1 };
					const int dc[8] = { -1, -1, -1, 1, 1, 1, 0, 0 };
					nr = r + dr[lane];
					nc = c + dc[lane];
					exists = nr >= 0 && nr < rows && nc >= 0 && nc < cols;
					if (exists && occ8[(size_t)nr * cols + nc])
						exists = false;
					if (exists && nr != r && nc != c && occ8[(size_t)nr * cols + c] && occ8[(size_t)r * cols + nc])
						exists = false;
				}
				double step = 0.0, h = 0.0;
				if (exists) {
					step = step_cost(F, r, c, nr, nc);
					h = heuristic(F, nr, nc, A.gr, A.gc);
				}
				const double parentG = A.W.g[n];
				for (int j = 0; j < 8; j++) {
					const bool ex = __shfl_sync(0xffffffffu, (int)exists, j) != 0;
					const double st = __shfl_sync(0xffffffffu, step, j), hj = __shfl_sync(0xffffffffu, h, j);
					const int cr = __shfl_sync(0xffffffffu, nr, j), cc = __shfl_sync(0xffffffffu, nc, j);
					if (!ex || lane != 0 || status != 1)
						continue;
					const int ci = cr * cols + cc;
					const uint32_t e = A.W.table[ci];
					const double g = parentG + st, f = g + hj;
					bool push = false;
					if (e == 0u)
						push = true;
					else if (!(e & 0x80000000u)) { // in the frontier: ProcessPossibleShortcut
						const int fn = (int)(e - 2u);
						if (A.W.f[fn] > f) {
							A.W.dead[fn] = 1;
							A.live--;
							push = true;
						}
					}
					if (push) {
						if (A.numNodes >= L.maxNodes) {
							status = PPX_ERR_CAPACITY;
							continue;
						}
						const int m = A.numNodes++;
						A.W.cell[m] = ci;
						A.W.parent[m] = n;
						A.W.g[m] = g;
						A.W.f[m] = f;
						A.W.dead[m] = 0;
						heap_push(A.W, A.heapSize, A.seq, f, m);
						A.live++;
						A.W.table[ci] = (uint32_t)m + 2u;
					}
				}
				// FindIntersection(node, this search, the other search) (:184-198)
				if (lane == 0) {
					const uint32_t eo = Bo.W.table[cell];
					if (eo & 0x80000000u) {
						const int nb = (int)(eo & 0x7fffffffu);
						const double cand = A.W.g[n] + Bo.W.g[nb];
						if (cand < bestCost) {
							bestCost = cand;
							A.solution = n;
							Bo.solution = nb;
						}
					}
				}
				__syncwarp();
				status = __shfl_sync(0xffffffffu, status, 0);
				if (status != 1)
					break;
			}
			if (status != 1)
				break;
			// stopping criterion (:161-172)
			int stop = 0;
			if (lane == 0 && D[0].solution >= 0 && D[1].solution >= 0) {
				if (D[0].live == 0 || D[1].live == 0)
					stop = 1;
				else {
					const int ft = bi_clean_top(D[0]), rt = bi_clean_top(D[1]);
					if (D[0].W.g[ft] + D[1].W.g[rt] >= bestCost + costOffset)
						stop = 1;
				}
			}
			stop = __shfl_sync(0xffffffffu, stop, 0);
			if (stop)
				status = 0;
		}
		if (lane == 0) {
			ppx_astar_result res;
			res.status = status == 1 ? -1 : status;
			res.n_explored = D[0].nExplored + D[1].nExplored;
			res.n_nodes = D[0].numNodes + D[1].numNodes;
			res.path_len = 0;
			res.cost = PPX_INF;
			if (status == 0) {
				res.cost = D[0].W.g[D[0].solution] + D[1].W.g[D[1].solution]; // GetOptimalCost (:81-84)
				int df = 0, dr = 0;
				for (int m = D[0].solution; m >= 0; m = D[0].W.parent[m])
					df++;
				for (int m = D[1].solution; m >= 0; m = D[1].W.parent[m])
					dr++;
				res.path_len = df + dr; // GetPath (:66-72): forward path, then the reverse search's path reversed (the meeting cell twice)
				if (paths && maxPath > 0) {
					int32_t* out = paths + (size_t)q * maxPath * 2;
					int pos = df - 1;
					for (int m = D[0].solution; m >= 0; m = D[0].W.parent[m], pos--)
						if (pos < maxPath) {
							out[2 * pos] = D[0].W.cell[m] / cols;
							out[2 * pos + 1] = D[0].W.cell[m] % cols;
						}
					pos = df;
					for (int m = D[1].solution; m >= 0; m = D[1].W.parent[m], pos++)
						if (pos < maxPath) {
							out[2 * pos] = D[1].W.cell[m] / cols;
							out[2 * pos + 1] = D[1].W.cell[m] % cols;
						}
				}
			}
			results[q] = res;
		}
		__syncwarp();
	}
}

}

extern "C" int ppx_astar_n2_ex(ppx_ctx* ctx, int64_t nq, const int32_t* starts, const int32_t* goals, const ppx_astar_params* params,
	int32_t max_nodes, int32_t max_path, ppx_astar_result* results, int32_t* paths)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	PPX_REQUIRE(ctx, ctx->has_map && ctx->d_occ8, "ppx_astar_n2 needs a map with an occupancy grid");
	PPX_REQUIRE(ctx, nq >= 0 && max_path >= 0 && max_nodes >= 0 && params, "bad arguments");
	PPX_REQUIRE(ctx, nq == 0 || (starts && goals && results), "bad arguments");
	PPX_REQUIRE(ctx, max_path == 0 || paths, "max_path > 0 needs the paths array");
	PPX_REQUIRE(ctx, params->cost_kind >= 0 && params->cost_kind <= 3 && params->heuristic_kind >= 0 && params->heuristic_kind <= 3,
		"unknown cost / heuristic functor");
	PPX_REQUIRE(ctx, params->cost_kind != 3 || params->cell_cost, "cost functor 3 needs the per-cell cost grid");
	if (nq == 0)
		return PPX_OK;
	const size_t cells = ctx->cells;
	const int dirs = params->bidirectional ? 2 : 1;
	const int maxNodes = max_nodes > 0 ? max_nodes : (int)std::min<size_t>(4 * cells + 16, (size_t)1 << 28);
	AstarLayout L = astar_layout(cells, maxNodes);
	int warps = (int)std::min<int64_t>(nq, (int64_t)ctx->num_sms * kAstarWarps * 2);
	{ // one workspace per resident warp (two for a bidirectional search): keep the total under a quarter of the free device memory
		size_t freeB = 0, totalB = 0;
		PPX_CUDA(ctx, cudaMemGetInfo(&freeB, &totalB));
		const size_t budget = std::max<size_t>(freeB / 4, L.bytes * dirs);
		warps = (int)std::max<size_t>(1, std::min<size_t>((size_t)warps, budget / (L.bytes * dirs)));
	}
	const int ctas = (warps + kAstarWarps - 1) / kAstarWarps;
	char* slabs;
	PPX_CHECK(ppx_scratch(ctx, 7, (size_t)ctas * kAstarWarps * L.bytes * dirs, (void**)&slabs));
	size_t off = 0;
	auto carve = [&](size_t b) {
		size_t o = off;
		off += (b + 255) & ~(size_t)255;
		return o;
	};
	const size_t bEnds = (size_t)nq * 8, bRes = (size_t)nq * sizeof(ppx_astar_result), bPaths = (size_t)nq * max_path * 8;
	const size_t bCost = params->cost_kind == 3 ? cells * sizeof(float) : 0;
	const size_t oS = carve(bEnds), oG = carve(bEnds), oR = carve(bRes), oP = carve(bPaths), oC = carve(bCost);
	char* base;
	PPX_CHECK(ppx_scratch(ctx, 5, off, (void**)&base));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oS, starts, bEnds, cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(base + oG, goals, bEnds, cudaMemcpyHostToDevice, ctx->stream));
	if (bCost)
		PPX_CUDA(ctx, cudaMemcpyAsync(base + oC, params->cell_cost, bCost, cudaMemcpyHostToDevice, ctx->stream));
	if (bPaths)
		PPX_CUDA(ctx, cudaMemsetAsync(base + oP, 0xff, bPaths, ctx->stream)); // -1 past path_len
	PPX_CUDA(ctx, cudaMemsetAsync(ctx->d_queue, 0, sizeof(unsigned long long), ctx->stream));
	AstarFn F { params->cost_kind, params->heuristic_kind, bCost ? (const float*)(base + oC) : nullptr, ctx->desc.cols };
	if (params->bidirectional)
		k_bi_astar_n2<<<ctas, kAstarWarps * 32, 0, ctx->stream>>>(F, ctx->d_occ8, ctx->desc.rows, ctx->desc.cols, nq, (const int32_t*)(base + oS),
			(const int32_t*)(base + oG), L, slabs, max_path, (ppx_astar_result*)(base + oR), max_path ? (int32_t*)(base + oP) : nullptr,
			ctx->d_queue);
	else
		k_astar_n2<<<ctas, kAstarWarps * 32, 0, ctx->stream>>>(F, ctx->d_occ8, ctx->desc.rows, ctx->desc.cols, nq, (const int32_t*)(base + oS),
			(const int32_t*)(base + oG), L, slabs, max_path, (ppx_astar_result*)(base + oR), max_path ? (int32_t*)(base + oP) : nullptr,
			ctx->d_queue);
	PPX_LAUNCHED(ctx);
	PPX_CUDA(ctx, cudaMemcpyAsync(results, base + oR, bRes, cudaMemcpyDeviceToHost, ctx->stream));
	if (bPaths)
		PPX_CUDA(ctx, cudaMemcpyAsync(paths, base + oP, bPaths, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

extern "C" int ppx_astar_n2(ppx_ctx* ctx, int64_t nq, const int32_t* starts, const int32_t* goals, int32_t max_nodes,
	int32_t max_path, ppx_astar_result* results, int32_t* paths)
{
	ppx_astar_params p {}; // Euclidean step cost and heuristic, unidirectional
	return ppx_astar_n2_ex(ctx, nq, starts, goals, &p, max_nodes, max_path, results, paths);
}
