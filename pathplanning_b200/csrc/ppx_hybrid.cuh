// Device-side restatement of Hybrid A*'s successor generation and heuristics
// (algo/hybrid_a_star.cpp:59-173, algo/hybrid_a_star.h:104-111,193-211, algo/heuristics.cpp:78-95,155-165)
// plus the bookkeeping primitives of the on-device graph search (algo/a_star.h:326-434 over
// utils/frontier.h).  Operation order and float/double mixing follow the reference line by line;
// the parity traps of SURVEY.md section 8 (H2, H3, H4, H7, H10, H11) are reproduced, not fixed.
#pragma once
#include "ppx_device.cuh"
#include "ppx_fastlane.cuh"

namespace ppx {

constexpr int kMaxDeltas = 9;
constexpr int kMaxChildren = 2 * kMaxDeltas + 1;

// Everything a successor / heuristic evaluation reads besides the map (kernel parameter).
struct HybridDev {
	// HybridAStar::SearchParameters, hybrid_a_star.h:29-50
	double rmin, swCost, revMul, fwdMul, voroMul, spatialRes, angRes;
	float revF, fwdF, swF; // the same multipliers as ReedsShepp::Solver / PathSegment::ComputeCost take them (float)
	double primLength; // spatialResolution * 1.5, hybrid_a_star.cpp:115
	int numDeltas; // m_deltas.size(), hybrid_a_star.cpp:20-28
	// KinematicBicycleModel::ConstantSteer constants per steering angle (kinematic_bicycle_model.cpp:14-18),
	// evaluated by the host mirror with the host libm like the reference does
	double beta[kMaxDeltas], cosBeta[kMaxDeltas], kappa[kMaxDeltas];
	float voroDiag; // m_voroFieldDiagResolution, hybrid_a_star.cpp:38
	// NonHolonomicHeuristic, heuristics.cpp:7-23 (table [nhX][nhY][nhA], row-major)
	const double* nh;
	int nhX, nhY, nhA;
	double nhOffX, nhOffY, nhMinMul;
	// ObstaclesHeuristic, heuristics.cpp:97-104
	float obsDiag, obsMul;
	// StateValidatorOccupancyMap public fields
	float minSafeRadius, minStep;
};

// Pose2<int>::WrapTheta as every TU but 2dplane.cpp instantiates it (trap H2): the generic
// while-loops run in int arithmetic, `t -= 2*M_PI` truncates toward zero.
PPX_HD int wrap_theta_int(int t)
{
	while ((double)t > PPX_PI)
		t = trunc_i32((double)t - 2 * PPX_PI);
	while ((double)t < -PPX_PI)
		t = trunc_i32((double)t + 2 * PPX_PI);
	return t;
}

struct Key3 {
	int x, y, k;
};
PPX_HD bool key_equal(const Key3& a, const Key3& b) { return a.x == b.x && a.y == b.y && a.k == b.k; }

// StatePropagator::DiscretizePose, hybrid_a_star.h:104-111 (traps H2, H10)
PPX_HD Key3 discretize_pose(const HybridDev& P, const Pose& p)
{
	Key3 k;
	k.x = trunc_i32(p.x / P.spatialRes);
	k.y = trunc_i32(p.y / P.spatialRes);
	k.k = wrap_theta_int(trunc_i32(wrap_theta(p.th) / P.angRes));
	return k;
}

// GraphSearch::IdenticalPoses, hybrid_a_star.h:207-211 (tol = 1e-3)
PPX_HD bool identical_poses(const Pose& a, const Pose& b)
{
	const double tol = 1e-3;
	double dx = a.x - b.x, dy = a.y - b.y;
	return sqrt(dx * dx + dy * dy) < tol && fabs(a.th - b.th) < tol * PPX_PI / 180.0;
}

// GVD::GetPathCost at OccupancyMap::WorldPositionToGridCell(position, false), hybrid_a_star.cpp:102-104.
// The reference indexes unchecked (the assert is compiled out): Grid<T>::operator[] is
// m_data + row * columns (utils/grid.h:87), so a column just outside the map aliases the neighbouring
// row of the same allocation -- reproduced here, because the end of a primitive may overshoot the
// bounds by up to one validator step (trap H6) and its cost decides the expansion order.  A flat
// index outside the allocation (undefined behaviour in the reference) reads 0.
PPX_HD float pathcost_at(const MapDev& m, double x, double y)
{
	int row = trunc_i32((x - m.ox) / m.resd);
	int col = trunc_i32((y - m.oy) / m.resd);
	long long flat = (long long)row * m.cols + col;
	if (flat < 0 || flat >= (long long)m.rows * m.cols)
		return 0.0f;
	return PPX_LDG(m.pathcost + flat);
}

// The largest `length` the loop of GetVoronoiCost (hybrid_a_star.cpp:100) visits: only the last
// sample survives because the loop assigns instead of accumulating (trap H7).  Returns false when
// the loop body never runs.
PPX_HD bool voronoi_last_length(double pathLength, float interpLength, double* lastLength)
{
	double length = 0.0;
	if (!(length < pathLength))
		return false;
	double lastv = 0.0;
	const double step = (double)interpLength;
	while (length < pathLength) {
		lastv = length;
		length += step;
		if (!(step > 0.0))
			break;
	}
	*lastLength = lastv;
	return true;
}

struct Child {
	bool ok;
	Pose pose; // State::continuousPose
	Key3 key; // State::discretePose
	int dir; // State::direction
	int kind; // 0 constant steer, 1 Reeds-Shepp
	double cost; // transition cost
	float last; // lastValidRatio (1 when the primitive was fully valid)
	double length; // action.path->GetLength() after truncation
};

// StatePropagator::GetConstantSteerChild, hybrid_a_star.cpp:111-147
PPX_HD void csteer_child(const HybridDev& P, const MapDev& map, const Pose& parent, const Key3& parentKey, int deltaIdx,
	int direction, Child& out, const FastMap* fm = nullptr)
{
	CSteer p;
	p.start = parent;
	p.length = P.primLength;
	p.beta = P.beta[deltaIdx];
	p.cosBeta = P.cosBeta[deltaIdx];
	p.kappa = P.kappa[deltaIdx];
	p.dir = direction;
	p.curved = fabs(p.kappa) > 1e-9;
	if (p.curved)
		PPX_SINCOS(p.beta + p.start.th, &p.sB0, &p.cB0);
	else
		PPX_SINCOS(p.start.th, &p.sT0, &p.cT0);
	out.kind = 0;
	out.dir = direction; // PathConstantSteer::GetDirection
	float last;
	bool valid = fm ? csteer_is_path_valid_filtered(map, *fm, p, P.minSafeRadius, P.minStep, &last, nullptr)
	                : csteer_is_path_valid(map, p, P.minSafeRadius, P.minStep, &last, nullptr);
	out.last = last;
	if (valid) {
		out.pose = p.at(1.0); // m_final of the constructor
		out.key = discretize_pose(P, out.pose);
	} else {
		// Truncate(lastValidRatio): m_final = Interpolate(ratio); m_length *= ratio
		out.pose = p.at((double)last);
		p.length *= (double)last;
		out.key = discretize_pose(P, out.pose);
		if (key_equal(out.key, parentKey)) {
			out.ok = false;
			return;
		}
	}
	out.length = p.length;
	double pathCost = (direction == 0 ? P.fwdMul : P.revMul) * p.length;
	double switchingCost = 0.0; // GetDirection(1.0) == GetDirection(1.0), hybrid_a_star.cpp:142
	// GetVoronoiCost(*action.path), hybrid_a_star.cpp:93-109
	float voronoiCost = 0.0f;
	double lastLength;
	if (voronoi_last_length(p.length, P.voroDiag, &lastLength)) {
		Pose s = p.at(lastLength / p.length);
		voronoiCost = pathcost_at(map, s.x, s.y);
	}
	voronoiCost *= P.voroDiag;
	out.cost = pathCost + switchingCost + P.voroMul * (double)voronoiCost;
	out.ok = true;
}

// PathReedsShepp::GetDirection(ratio), path_reeds_shepp.cpp:153-166
PPX_HD int rs_direction_at(const RSSeg& seg, double r, double pathLength, double ratio)
{
	if (pathLength == 0)
		return 2;
	double length = 0;
	for (int i = 0; i < 5; i++) {
		if (!seg.valid(i))
			break;
		length += seg.len[i] * r;
		if (ratio * pathLength <= length)
			return seg.dir(i);
	}
	return 2;
}

// PathReedsShepp::Interpolate(ratio), path_reeds_shepp.cpp:12-47
PPX_HD Pose rs_interpolate(const Pose& start, const RSSeg& seg, double r, double pathLength, double ratio)
{
	if (pathLength == 0)
		return start;
	RSWalker w;
	w.init(start);
	return w.at(seg, r, ratio * pathLength);
}

// StatePropagator::GetReedsSheppChild, hybrid_a_star.cpp:149-173
PPX_HD void rs_child_from(const HybridDev& P, const MapDev& map, const Pose& parent, const RSSolution& sol, Child& out,
	const FastMap* fm = nullptr);
PPX_HD void rs_child(const HybridDev& P, const MapDev& map, const Pose& parent, const Pose& goal, Child& out)
{
	RSSolution sol;
	rs_solve(parent, goal, P.rmin, P.revF, P.fwdF, P.swF, 0, sol);
	rs_child_from(P, map, parent, sol, out);
}

// ... given the optimal segment (the search kernel solves it with the whole warp, rs_solve_warp)
PPX_HD void rs_child_from(const HybridDev& P, const MapDev& map, const Pose& parent, const RSSolution& sol, Child& out,
	const FastMap* fm)
{
	const double pathLength = sol.mLength * P.rmin; // PathSegment::GetLength(minTurningRadius)
	out.kind = 1;
	out.last = 1.0f;
	out.length = pathLength;
	float last;
	bool valid = fm ? rs_is_path_valid_filtered(map, *fm, parent, sol.seg, P.rmin, P.minSafeRadius, P.minStep, &last, nullptr)
	                : rs_is_path_valid(map, parent, sol.seg, P.rmin, P.minSafeRadius, P.minStep, &last, nullptr);
	if (!valid) {
		out.ok = false;
		return;
	}
	// path->ComputeCost(sw, rev, fwd): the double parameters narrow to PathSegment::ComputeCost's floats
	double pathAndSwitchingCosts = (double)rs_compute_cost(sol.seg, sol.mLength, P.rmin, P.revF, P.fwdF, P.swF);
	out.dir = rs_direction_at(sol.seg, P.rmin, pathLength, 1.0);
	out.pose = rs_interpolate(parent, sol.seg, P.rmin, pathLength, 1.0); // m_final
	out.key = discretize_pose(P, out.pose);
	float voronoiCost = 0.0f;
	double lastLength;
	if (voronoi_last_length(pathLength, P.voroDiag, &lastLength)) {
		Pose s = rs_interpolate(parent, sol.seg, P.rmin, pathLength, lastLength / pathLength);
		voronoiCost = pathcost_at(map, s.x, s.y);
	}
	voronoiCost *= P.voroDiag;
	out.cost = pathAndSwitchingCosts + P.voroMul * (double)voronoiCost;
	out.ok = true;
}

// NonHolonomicHeuristic::GetHeuristicValue, heuristics.cpp:78-95.  Reads with k < 0 (out of bounds in
// the reference, trap H3) are DEFINED here as k + numAngular -- the same definition the oracle
// harness applies to the reference.
PPX_HD double nonholo_heuristic(const HybridDev& P, const Pose& goal, const Pose& state)
{
	// delta = m_goal - state (2dplane.h:64-78)
	double dx = goal.x - state.x, dy = goal.y - state.y;
	double s, c;
	PPX_SINCOS(-state.th, &s, &c);
	double px = c * dx + (-s) * dy;
	double py = s * dx + c * dy;
	double dth = wrap_theta(wrap_theta(goal.th - state.th));
	int i = trunc_i32(round((px + P.nhOffX) / P.spatialRes));
	int j = trunc_i32(round((py + P.nhOffY) / P.spatialRes));
	int k = trunc_i32(round(dth / P.angRes));
	if (k == P.nhA)
		k = 0;
	if (i < 0 || i >= P.nhX || j < 0 || j >= P.nhY)
		return P.nhMinMul * sqrt(px * px + py * py);
	if (k < 0)
		k += P.nhA;
	if (k < 0 || k >= P.nhA)
		k = 0; // unreachable for |dth| <= pi; keeps the read inside the table
	return PPX_LDG(P.nh + ((size_t)i * P.nhY + j) * P.nhA + k);
}

// ObstaclesHeuristic::GetHeuristicValue, heuristics.cpp:155-165; wfCost is m_cost of the wavefront
// from this goal (explored <=> finite cost once Update has drained its frontier).
PPX_HD double obstacle_heuristic(const HybridDev& P, const MapDev& m, const float* __restrict__ wfCost, const Pose& goal,
	const Pose& state)
{
	double dx = goal.x - state.x, dy = goal.y - state.y;
	double euclidean = sqrt(dx * dx + dy * dy);
	int row = trunc_i32((state.x - m.ox) / m.resd);
	int col = trunc_i32((state.y - m.oy) / m.resd);
	if (row < 0 || row >= m.rows || col < 0 || col >= m.cols)
		return euclidean;
	float cost = wfCost[(size_t)row * m.cols + col];
	if (cost == __int_as_float(0x7f800000))
		return euclidean;
	double heuristic = (double)(cost * P.obsMul - P.obsDiag);
	return (heuristic < euclidean) ? euclidean : heuristic; // std::max(heuristic, euclidean)
}

// AStarCombinedHeuristic (a_star.h:102-109) over {non-holonomic, obstacles} through the HeuristicAdapter
PPX_HD double combined_heuristic(const HybridDev& P, const MapDev& m, const float* __restrict__ wfCost, const Pose& goal,
	const Pose& state)
{
	double value = -PPX_INF;
	double h0 = nonholo_heuristic(P, goal, state);
	value = (value < h0) ? h0 : value;
	double h1 = obstacle_heuristic(P, m, wfCost, goal, state);
	value = (value < h1) ? h1 : value;
	return value;
}

// ---------------------------------------------------------------------------------------------
// std::mt19937_64 + the reference's Random<double>::SampleUniform(0, 1) (utils/random.h:12-27, trap H4):
// libstdc++ generate_canonical<double,53> takes ONE 64-bit draw, (double)u / 2^64 (clamped below 1),
// and uniform_real_distribution(0, nextafter(1)) scales it by (1 + 2^-52).
struct Mt64 {
	unsigned long long* mt; // 312 words
	int* idx;
	PPX_HD void seed(unsigned long long s)
	{
		mt[0] = s;
		for (int i = 1; i < 312; i++)
			mt[i] = 6364136223846793005ULL * (mt[i - 1] ^ (mt[i - 1] >> 62)) + (unsigned long long)i;
		*idx = 312;
	}
	PPX_HD unsigned long long next()
	{
		if (*idx >= 312) {
			const unsigned long long UM = 0xFFFFFFFF80000000ULL, LM = 0x7FFFFFFFULL, A = 0xB5026F5AA96619E9ULL;
			for (int i = 0; i < 312; i++) {
				unsigned long long x = (mt[i] & UM) | (mt[(i + 1) % 312] & LM);
				mt[i] = mt[(i + 156) % 312] ^ (x >> 1) ^ ((x & 1ULL) ? A : 0ULL);
			}
			*idx = 0;
		}
		unsigned long long x = mt[(*idx)++];
		x ^= (x >> 29) & 0x5555555555555555ULL;
		x ^= (x << 17) & 0x71D67FFFEDA60000ULL;
		x ^= (x << 37) & 0xFFF7EEE000000000ULL;
		x ^= (x >> 43);
		return x;
	}
	PPX_HD double uniform01()
	{
		double canon = (double)next() / 18446744073709551616.0;
		if (canon >= 1.0)
			canon = 0.99999999999999988897769753748;
		const double range = 1.0000000000000002220446049250313; // nextafter(1.0, max) - 0.0
		return 0.0 + 1.0 * (canon * range + 0.0);
	}
};


// ---------------------------------------------------------------------------------------------
// Graph-search bookkeeping (algo/a_star.h:326-434 over utils/frontier.h), one instance per query.
//
//  * Frontier (utils/frontier.h): a vector kept sorted by descending totalCost, new elements
//    inserted after the run of elements that are >= (frontier.h:39-48,111-129), popped from the back.
//    Pop order is therefore "smallest totalCost first, most recently pushed first among equals":
//    a binary min-heap on (totalCost asc, push sequence desc) pops in exactly that order.
//    Frontier::Remove (frontier.h:52-77) becomes a dead flag on the node (skipped when popped).
//  * Frontier::Find / Explored::Contains hash the DISCRETE pose only (hybrid_a_star.h:72-86).  With
//    trap H2 the heading bin is always in [-3, 3], so the key space is a dense table over the map
//    bounds: 0 absent, 1 explored, n+2 in the frontier as node n.
struct SearchWs {
	double* pose; // [maxNodes][3] continuousPose
	double* g; // pathCost
	double* f; // totalCost
	double* h; // heuristic value of the node's state (what GetNeighborStates recomputes as hCost)
	int* parent;
	int* keyIdx; // table index of the node's discrete pose
	uint32_t* meta; // dir (bits 0-1) | kind (bit 2) | child slot (bits 3-7) | dead (bit 8)
	float* last; // lastValidRatio of the action that created the node
	double* length; // action path length
	double* heapF;
	uint32_t* heapSeq;
	uint32_t* heapNode;
	uint32_t* table;
	unsigned long long* mt; // 312 words
	int maxNodes;
	int tx, ty, x0, y0; // key table geometry: (kx - x0) in [0, tx), (ky - y0) in [0, ty)
};

struct SearchState {
	int numNodes, heapSize, mtIdx;
	uint32_t seq;
	int status; // 1 running, 0 success, -1 failure, -4 capacity
	int solution;
	int nExpanded, nChildren, nRsAttempts;
	unsigned long long hash;
};

enum { PPX_SEARCH_RUNNING = 1 };

PPX_HD int key_table_index(const SearchWs& W, const Key3& k)
{
	int ix = k.x - W.x0, iy = k.y - W.y0, ik = k.k + 3;
	if (ix < 0 || ix >= W.tx || iy < 0 || iy >= W.ty || ik < 0 || ik > 6)
		return -1;
	return (ix * W.ty + iy) * 7 + ik;
}

// a is popped before b
PPX_HD bool heap_before(double fa, uint32_t sa, double fb, uint32_t sb) { return fa < fb || (fa == fb && sa > sb); }

PPX_HD void heap_push(const SearchWs& W, SearchState& S, double f, int node)
{
	uint32_t seq = S.seq++;
	int i = S.heapSize++;
	while (i > 0) {
		int p = (i - 1) >> 1;
		double pf = W.heapF[p];
		uint32_t ps = W.heapSeq[p];
		if (!heap_before(f, seq, pf, ps))
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

PPX_HD int heap_pop(const SearchWs& W, SearchState& S)
{
	int top = (int)W.heapNode[0];
	int n = --S.heapSize;
	if (n > 0) {
		double f = W.heapF[n];
		uint32_t seq = W.heapSeq[n];
		uint32_t node = W.heapNode[n];
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
				if (heap_before(rf, rs, cf, cs)) {
					c = r;
					cf = rf;
					cs = rs;
				}
			}
			if (!heap_before(cf, cs, f, seq))
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

PPX_HD int node_alloc(const SearchWs& W, SearchState& S, const Pose& pose, int keyIdx, int parent, double g, double f,
	double h, uint32_t meta, float last, double length)
{
	if (S.numNodes >= W.maxNodes) {
		S.status = PPX_ERR_CAPACITY;
		return -1;
	}
	int n = S.numNodes++;
	W.pose[3 * n] = pose.x;
	W.pose[3 * n + 1] = pose.y;
	W.pose[3 * n + 2] = pose.th;
	W.g[n] = g;
	W.f[n] = f;
	W.h[n] = h;
	W.parent[n] = parent;
	W.keyIdx[n] = keyIdx;
	W.meta[n] = meta;
	W.last[n] = last;
	W.length[n] = length;
	return n;
}

// AStar::InitializeSearch, a_star.h:350-364 (the table must be zeroed by the caller)
PPX_HD void search_init(const HybridDev& P, const SearchWs& W, SearchState& S, const Pose& start, double hStart)
{
	S.numNodes = 0;
	S.heapSize = 0;
	S.seq = 0;
	S.status = PPX_SEARCH_RUNNING;
	S.solution = -1;
	S.nExpanded = S.nChildren = S.nRsAttempts = 0;
	S.hash = 1469598103934665603ULL;
	Key3 k = discretize_pose(P, start); // CreateStateFromPose, hybrid_a_star.cpp:50-57 (direction NoMotion)
	int ki = key_table_index(W, k);
	if (ki < 0) {
		S.status = PPX_ERR_CAPACITY;
		return;
	}
	int n = node_alloc(W, S, start, ki, -1, 0.0, 0.0, hStart, 2u, 1.0f, 0.0); // root: pathCost = totalCost = 0
	if (n < 0)
		return;
	heap_push(W, S, 0.0, n);
	W.table[ki] = (uint32_t)n + 2u;
}

// One turn of the loop of AStar::SearchPath (a_star.h:336-344) up to and including
// m_explored.Insert(node) (a_star.h:381).  Returns the node to expand, or -1 when the search ended
// (S.status says how).
PPX_HD int search_pop(const SearchWs& W, SearchState& S, const Pose& goal)
{
	while (S.heapSize > 0) {
		int n = heap_pop(W, S);
		if (W.meta[n] & 0x100u)
			continue; // removed from the frontier by ProcessPossibleShortcut
		Pose p = { W.pose[3 * n], W.pose[3 * n + 1], W.pose[3 * n + 2] };
		if (identical_poses(p, goal)) { // GraphSearch::IsSolution, hybrid_a_star.h:193-196
			S.solution = n;
			S.status = PPX_OK;
			return -1;
		}
		W.table[W.keyIdx[n]] = 1u; // explored (and no longer in the frontier)
		S.nExpanded++;
		S.hash = (S.hash ^ (unsigned long long)(uint32_t)W.keyIdx[n]) * 1099511628211ULL;
		return n;
	}
	S.status = -1; // Status::Failure: frontier exhausted
	return -1;
}

// The body of the loop of AStar::Expand for one child (a_star.h:385-402) with
// GraphSearch::ProcessPossibleShortcut (hybrid_a_star.h:199-205).
PPX_HD void consume_child(const SearchWs& W, SearchState& S, int parentNode, double parentG, const Child& c, double h, int slot)
{
	S.nChildren++;
	double g = parentG + c.cost;
	double f = g + h;
	int ki = key_table_index(W, c.key);
	if (ki < 0) {
		S.status = PPX_ERR_CAPACITY;
		return;
	}
	uint32_t e = W.table[ki];
	uint32_t meta = (uint32_t)(c.dir & 3) | ((uint32_t)(c.kind & 1) << 2) | ((uint32_t)(slot & 31) << 3);
	if (e == 0u) {
		int n = node_alloc(W, S, c.pose, ki, parentNode, g, f, h, meta, c.last, c.length);
		if (n < 0)
			return;
		heap_push(W, S, f, n);
		W.table[ki] = (uint32_t)n + 2u;
	} else if (e >= 2u) {
		int fn = (int)(e - 2u);
		Pose fp = { W.pose[3 * fn], W.pose[3 * fn + 1], W.pose[3 * fn + 2] };
		if (identical_poses(fp, c.pose) && W.f[fn] > f) {
			int n = node_alloc(W, S, c.pose, ki, parentNode, g, f, h, meta, c.last, c.length);
			if (n < 0)
				return;
			W.meta[fn] |= 0x100u; // m_frontier.Remove(frontierNode)
			heap_push(W, S, f, n);
			W.table[ki] = (uint32_t)n + 2u;
		}
	}
}

// Decision of hybrid_a_star.cpp:81-82: attempt a Reeds-Shepp shot?  Draws from the query's RNG only
// when hCost >= 10 (short-circuit ||).
PPX_HD bool rs_shot_wanted(Mt64& rng, double hCost)
{
	if (hCost < 10.0)
		return true;
	return rng.uniform01() < 10.0 / (hCost * hCost);
}

// Child slot s of GetNeighborStates (hybrid_a_star.cpp:59-91): slots 0 .. 2*numDeltas-1 are the
// constant-steer primitives in (delta, Forward/Backward) order, slot 2*numDeltas is the RS shot.
PPX_HD void make_child(const HybridDev& P, const MapDev& map, const Pose& parent, const Key3& parentKey, const Pose& goal,
	int slot, Child& out)
{
	if (slot < 2 * P.numDeltas)
		csteer_child(P, map, parent, parentKey, slot >> 1, slot & 1, out);
	else
		rs_child(P, map, parent, goal, out);
}

}
