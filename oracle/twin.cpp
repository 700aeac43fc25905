// TEST INFRASTRUCTURE ONLY -- never linked into, imported by, or executed from the product path.
//
// "Twin": the arithmetic / state-machine headers of libppx (pathplanning_b200/csrc/*.cuh) compiled
// by g++ for the host (PPX_HOST_TWIN), behind a few C entry points.  There is no GPU in the build
// container, so this is how the control logic of the kernels (micro-step state machines, frontier
// tie orders, the reference's parity traps) is debugged against the compiled reference
// (oracle/_ref/libpp_ref.so) and the golden fixtures before a kernel goes to the B200.
// It is NOT the parity oracle: parity is pinned by the unmodified reference (ref_harness.cpp) and
// by tests/golden; and it is NOT a CPU fallback: nothing under pathplanning_b200/ loads it.
// Note that host libm (glibc) and device libm (CUDA) differ in the last ulp, so twin == reference
// does not by itself prove kernel == reference; the -m gpu tests do that.
#define PPX_HOST_TWIN 1
#include <vector>
#include <cstdio>
#include "ppx_device.cuh"
#include "ppx_fastlane.cuh"
#include "ppx_hybrid_host.h"
#include "ppx_resample.cuh"
#include "ppx_smooth.cuh"

using namespace ppx;

namespace {
struct TwinWorld {
	ppx_map_desc desc;
	std::vector<float> clearance;
	std::vector<int> occupancy;
	std::vector<float> pathcost;
	MapDev dev() const
	{
		MapDev m;
		m.rows = desc.rows;
		m.cols = desc.cols;
		m.res = desc.resolution;
		m.resd = (double)desc.resolution;
		m.ox = desc.world_origin[0];
		m.oy = desc.world_origin[1];
		m.lox = desc.local_origin[0];
		m.loy = desc.local_origin[1];
		m.lbx = desc.lb[0]; m.lby = desc.lb[1]; m.lbt = desc.lb[2];
		m.ubx = desc.ub[0]; m.uby = desc.ub[1]; m.ubt = desc.ub[2];
		m.clearance = clearance.data();
		m.occupancy = occupancy.data();
		m.pathcost = pathcost.data();
		return m;
	}
};

struct LenArray {
	const double* p;
	double get(int i) const { return p[i]; }
};
}

extern "C" {

void* twin_world_create(const ppx_map_desc* desc, const int32_t* sqdist, const int32_t* occupancy, const float* pathcost)
{
	auto* w = new TwinWorld();
	w->desc = *desc;
	size_t n = (size_t)desc->rows * desc->cols;
	w->clearance.resize(n);
	w->occupancy.assign(n, -1);
	w->pathcost.assign(n, 0.0f);
	for (size_t i = 0; i < n; i++)
		w->clearance[i] = (float)(sqrt((double)sqdist[i]) * (double)desc->resolution);
	if (occupancy)
		w->occupancy.assign(occupancy, occupancy + n);
	if (pathcost)
		w->pathcost.assign(pathcost, pathcost + n);
	return w;
}

void twin_world_destroy(void* h) { delete (TwinWorld*)h; }

static double* g_audit = nullptr; // [max deviation (cells), certified, parked, parked: near cell border, parked: map-border rule, parked: slow-only edge]
void twin_set_audit(double* a) { g_audit = a; }

// impl 2: filtered FP32 lane machine with exact fallback (RSFastLane)
// impl 0: thread-per-edge walker (rs_is_path_valid); impl 1: micro-step lane machine (RSLane)
void twin_check_rs(void* h, long n, const ppx_rs_edge* edges, double r, float radius, float minStep, int impl,
	uint8_t* valid, float* last, uint32_t* samples, uint32_t* microsteps)
{
	auto* w = (TwinWorld*)h;
	MapDev map = w->dev();
	for (long i = 0; i < n; i++) {
		Pose start = { edges[i].start[0], edges[i].start[1], edges[i].start[2] };
		if (impl == 0) {
			RSSeg seg;
			for (int j = 0; j < 5; j++)
				seg.len[j] = edges[i].len[j];
			seg.code = (uint32_t)edges[i].code;
			float l;
			unsigned ns;
			valid[i] = rs_is_path_valid(map, start, seg, r, radius, minStep, &l, &ns);
			last[i] = l;
			if (samples) samples[i] = ns;
			if (microsteps) microsteps[i] = 0;
		} else if (impl == 2) {
			// filtered lane machine (ppx_fastlane.cuh) run sequentially: parked samples go straight to the exact path
			LenArray L { edges[i].len };
			FastMap fm = fast_map_derive(w->desc);
			RSFastLane<LenArray> lane;
			lane.begin(L, fm, map, start, (uint32_t)edges[i].code, r);
			RSSeg seg;
			for (int j = 0; j < 5; j++)
				seg.len[j] = edges[i].len[j];
			seg.code = (uint32_t)edges[i].code;
			const float inv_r = (float)(1.0 / r), rc = (float)(r * fm.inv_resd);
			unsigned steps = 0, parkedCount = 0;
			double ex, ey;
			while (true) {
				steps++;
				int rcode = 2;
				double lengthAtSample = lane.length;
				if (!lane.slowOnly) {
					if (!lane.advance(L, r, inv_r, rc, fm.inv_res))
						continue;
					rcode = lane.sample(map, fm, radius, minStep);
				}
				if (rcode == 2) {
					parkedCount++;
					if (g_audit && !lane.slowOnly) {
						float fx = floorf(lane.px), fy = floorf(lane.py), gx = lane.px - fx, gy = lane.py - fy;
						bool near = fminf(gx, gy) < kCellEps || fmaxf(gx, gy) > 1.0f - kCellEps;
						g_audit[near ? 3 : 4] += 1; // 3: near a cell border, 4: border / interior rule
					} else if (g_audit)
						g_audit[5] += 1; // slow-only edge
					if (rs_exact_sample(lane, map, start, seg, r, radius, minStep, &ex, &ey))
						break;
					continue;
				}
				if (g_audit) { // exact position of the certified sample
					RSWalker wk;
					wk.init(start);
					Pose p = wk.at(seg, r, (lengthAtSample / lane.pathLength) * lane.pathLength);
					double dx = (p.x - map.ox) / map.resd - lane.rowb - (double)lane.px;
					double dy = (p.y - map.oy) / map.resd - lane.colb - (double)lane.py;
					double dev = fmax(fabs(dx), fabs(dy));
					if (dev > g_audit[0]) g_audit[0] = dev;
					g_audit[1] += 1;
				}
				if (rcode == 1)
					break;
			}
			if (g_audit) g_audit[2] += parkedCount;
			valid[i] = lane.verdict;
			last[i] = lane.last;
			if (samples) samples[i] = lane.ns;
			if (microsteps) microsteps[i] = steps;
		} else {
			LenArray L { edges[i].len };
			RSLane<LenArray> lane;
			lane.begin(L, start, (uint32_t)edges[i].code, r);
			unsigned steps = 0;
			while (true) {
				double arg = lane.pre(L, r);
				double s, c;
				sincos(arg, &s, &c);
				steps++;
				if (lane.post(map, s, c, r, radius, minStep))
					break;
			}
			valid[i] = lane.verdict;
			last[i] = lane.last;
			if (samples) samples[i] = lane.ns;
			if (microsteps) microsteps[i] = steps;
		}
	}
}


// ---------------------------------------------------------------------------------------------
// Hybrid A*: the successor / heuristic / bookkeeping code of ppx_hybrid.cuh run sequentially.  The
// non-holonomic table and the wavefront grid are INPUTS here (taken from the reference by the test),
// so this checks exactly the logic the search kernel runs per warp.
namespace {
struct TwinHybrid {
	TwinWorld* world;
	ppx_hybrid_params params;
	HybridDev dev;
	std::vector<double> table;
	int x0, y0, tx, ty;
};
}

void* twin_hybrid_create(void* hw, const ppx_hybrid_params* params, const ppx_validator_params* vp, const double* table)
{
	auto* w = (TwinWorld*)hw;
	auto* h = new TwinHybrid();
	h->world = w;
	h->params = *params;
	if (hybrid_derive(*params, *vp, w->desc, h->dev) != PPX_OK) {
		delete h;
		return nullptr;
	}
	size_t total = (size_t)h->dev.nhX * h->dev.nhY * h->dev.nhA;
	if (table) {
		h->table.assign(table, table + total);
	} else { // NonHolonomicHeuristic::Build with the twin's own solver
		h->table.resize(total);
		const HybridDev& P = h->dev;
		for (size_t idx = 0; idx < total; idx++) {
			int k = (int)(idx % P.nhA), j = (int)((idx / P.nhA) % P.nhY), i = (int)(idx / ((size_t)P.nhA * P.nhY));
			Pose pose = { i * P.spatialRes - P.nhOffX, j * P.spatialRes - P.nhOffY, wrap_theta(k * P.angRes) };
			Pose goal = { 0.0, 0.0, 0.0 };
			RSSolution sol;
			rs_solve(pose, goal, P.rmin, P.revF, P.fwdF, P.swF, 0, sol);
			h->table[idx] = (double)sol.cost;
		}
	}
	h->dev.nh = h->table.data();
	hybrid_key_table(*params, w->desc, h->x0, h->y0, h->tx, h->ty);
	return h;
}

void twin_hybrid_destroy(void* hh) { delete (TwinHybrid*)hh; }

void twin_hybrid_info(void* hh, int* dims, double* offsets, double* deltasOut, double* table)
{
	auto* h = (TwinHybrid*)hh;
	dims[0] = h->dev.nhX; dims[1] = h->dev.nhY; dims[2] = h->dev.nhA; dims[3] = h->dev.numDeltas;
	offsets[0] = h->dev.nhOffX; offsets[1] = h->dev.nhOffY;
	if (deltasOut)
		for (int i = 0; i < h->dev.numDeltas; i++)
			deltasOut[i] = h->dev.kappa[i];
	if (table)
		memcpy(table, h->table.data(), h->table.size() * sizeof(double));
}

void twin_hybrid_heuristic(void* hh, const float* wfCost, const double* goal, long n, const double* poses, double* combined,
	double* nonholo, double* obstacles)
{
	auto* h = (TwinHybrid*)hh;
	MapDev map = h->world->dev();
	Pose g = { goal[0], goal[1], goal[2] };
	for (long i = 0; i < n; i++) {
		Pose s = { poses[3 * i], poses[3 * i + 1], poses[3 * i + 2] };
		if (combined) combined[i] = combined_heuristic(h->dev, map, wfCost, g, s);
		if (nonholo) nonholo[i] = nonholo_heuristic(h->dev, g, s);
		if (obstacles) obstacles[i] = obstacle_heuristic(h->dev, map, wfCost, g, s);
	}
}

void twin_hybrid_expand(void* hh, const float* wfCost, const double* goal, long n, const double* parents, const uint64_t* seeds,
	int32_t* nChildren, ppx_hybrid_child* children, double* parentHeuristic)
{
	auto* h = (TwinHybrid*)hh;
	const HybridDev& P = h->dev;
	MapDev map = h->world->dev();
	Pose g = { goal[0], goal[1], goal[2] };
	std::vector<unsigned long long> mt(312);
	for (long i = 0; i < n; i++) {
		Pose parent = { parents[3 * i], parents[3 * i + 1], parents[3 * i + 2] };
		Key3 pk = discretize_pose(P, parent);
		double hCost = combined_heuristic(P, map, wfCost, g, parent);
		if (parentHeuristic) parentHeuristic[i] = hCost;
		int idx = 0;
		Mt64 rng { mt.data(), &idx };
		rng.seed(seeds[i]);
		int nslots = 2 * P.numDeltas + (rs_shot_wanted(rng, hCost) ? 1 : 0);
		int count = 0;
		for (int s = 0; s < nslots; s++) {
			Child c;
			make_child(P, map, parent, pk, g, s, c);
			if (!c.ok)
				continue;
			ppx_hybrid_child& o = children[i * PPX_MAX_CHILDREN + count++];
			o.pose[0] = c.pose.x; o.pose[1] = c.pose.y; o.pose[2] = c.pose.th;
			o.cost = c.cost;
			o.heuristic = combined_heuristic(P, map, wfCost, g, c.pose);
			o.length = c.length;
			o.key[0] = c.key.x; o.key[1] = c.key.y; o.key[2] = c.key.k;
			o.direction = c.dir;
			o.kind = c.kind;
			o.last = c.last;
		}
		nChildren[i] = count;
	}
}

// The loop of k_hybrid_search for one query, sequentially.
void twin_hybrid_search(void* hh, const float* wfCost, const double* start, const double* goal, uint64_t seed,
	const ppx_hybrid_limits* limits, ppx_hybrid_result* result, ppx_hybrid_path_state* path, ppx_hybrid_log_entry* log, double* fLog)
{
	auto* h = (TwinHybrid*)hh;
	const HybridDev& P = h->dev;
	MapDev map = h->world->dev();
	const int tableSize = h->tx * h->ty * 7;
	const int maxNodes = limits->max_nodes > 0 ? limits->max_nodes : 2 * tableSize;
	std::vector<double> pose(3 * (size_t)maxNodes), g(maxNodes), f(maxNodes), hv(maxNodes), length(maxNodes), heapF(maxNodes);
	std::vector<int> parent(maxNodes), keyIdx(maxNodes);
	std::vector<uint32_t> meta(maxNodes), heapSeq(maxNodes), heapNode(maxNodes), table(tableSize, 0u);
	std::vector<float> lastv(maxNodes);
	std::vector<unsigned long long> mt(312);
	SearchWs W;
	W.pose = pose.data(); W.g = g.data(); W.f = f.data(); W.h = hv.data(); W.parent = parent.data(); W.keyIdx = keyIdx.data();
	W.meta = meta.data(); W.last = lastv.data(); W.length = length.data(); W.heapF = heapF.data();
	W.heapSeq = heapSeq.data(); W.heapNode = heapNode.data(); W.table = table.data(); W.mt = mt.data();
	W.maxNodes = maxNodes; W.tx = h->tx; W.ty = h->ty; W.x0 = h->x0; W.y0 = h->y0;
	SearchState S;
	Mt64 rng { W.mt, &S.mtIdx };
	rng.seed(seed);
	Pose st = { start[0], start[1], start[2] }, gl = { goal[0], goal[1], goal[2] };
	search_init(P, W, S, st, combined_heuristic(P, map, wfCost, gl, st));
	while (S.status == PPX_SEARCH_RUNNING) {
		if (limits->max_expansions > 0 && S.nExpanded >= limits->max_expansions) {
			S.status = PPX_ERR_LIMIT;
			break;
		}
		int n = search_pop(W, S, gl);
		if (n < 0)
			break;
		Pose par = { W.pose[3 * n], W.pose[3 * n + 1], W.pose[3 * n + 2] };
		Key3 pk = discretize_pose(P, par);
		double hCost = W.h[n];
		if (log && S.nExpanded <= limits->max_log) {
			ppx_hybrid_log_entry& e = log[S.nExpanded - 1];
			e.key[0] = pk.x; e.key[1] = pk.y; e.key[2] = pk.k;
			e.direction = (int)(W.meta[n] & 3u);
			e.pose[0] = par.x; e.pose[1] = par.y; e.pose[2] = par.th;
			if (fLog) { fLog[2 * (S.nExpanded - 1)] = W.g[n]; fLog[2 * (S.nExpanded - 1) + 1] = W.f[n]; }
		}
		int want = rs_shot_wanted(rng, hCost) ? 1 : 0;
		S.nRsAttempts += want;
		int nslots = 2 * P.numDeltas + want;
		double parentG = W.g[n];
		for (int s = 0; s < nslots && S.status == PPX_SEARCH_RUNNING; s++) {
			Child c;
			make_child(P, map, par, pk, gl, s, c);
			if (!c.ok)
				continue;
			double hc = combined_heuristic(P, map, wfCost, gl, c.pose);
			consume_child(W, S, n, parentG, c, hc, s);
		}
	}
	ppx_hybrid_result r;
	r.status = S.status == PPX_SEARCH_RUNNING ? -1 : S.status;
	r.n_expanded = S.nExpanded; r.n_nodes = S.numNodes; r.n_children = S.nChildren; r.n_rs_attempts = S.nRsAttempts;
	r.path_len = 0; r.cost = PPX_INF; r.expanded_hash = S.hash;
	if (S.status == PPX_OK && S.solution >= 0) {
		r.cost = W.g[S.solution];
		int depth = 0;
		for (int m = S.solution; m >= 0; m = W.parent[m]) depth++;
		r.path_len = depth;
		if (path && limits->max_path > 0) {
			int pos = depth - 1;
			for (int m = S.solution; m >= 0; m = W.parent[m], pos--) {
				if (pos >= limits->max_path) continue;
				ppx_hybrid_path_state& o = path[pos];
				o.pose[0] = W.pose[3 * m]; o.pose[1] = W.pose[3 * m + 1]; o.pose[2] = W.pose[3 * m + 2];
				o.length = W.length[m]; o.last = W.last[m];
				uint32_t mm = W.meta[m];
				o.kind = W.parent[m] < 0 ? (int8_t)-1 : (int8_t)((mm >> 2) & 1u);
				o.direction = (int8_t)(mm & 3u); o.slot = (int8_t)((mm >> 3) & 31u); o._pad = 0;
			}
		}
	}
	*result = r;
}

// k_hybrid_resample for one query, sequentially (hybrid_a_star.cpp:260-291).
long twin_hybrid_resample(void* hh, const double* goal, const ppx_hybrid_path_state* path, int pathLen, float pathInterpolation,
	long cap, double* poses, uint8_t* cusp, double* pathLength)
{
	auto* h = (TwinHybrid*)hh;
	const HybridDev& P = h->dev;
	std::vector<double> finalLength(pathLen > 0 ? pathLen : 1);
	ResamplePath rp;
	Pose gl = { goal[0], goal[1], goal[2] };
	int n = resample_plan(P, path, pathLen, gl, pathInterpolation, finalLength.data(), rp, poses, cusp, (int)cap);
	for (int i = 0; i < n; i++) {
		Pose p = resample_pose(P, rp, poses[3 * i]);
		poses[3 * i] = p.x; poses[3 * i + 1] = p.y; poses[3 * i + 2] = p.th;
	}
	if (pathLength)
		*pathLength = rp.length;
	return n;
}

// k_smooth for one path, sequentially: phase A / phase B of ppx_smooth.cuh (smoother.cpp:33-158).
int twin_smooth(void* hw, const ppx_smoother_params* p, const int32_t* nearObstacle, const int32_t* nearEdge, long n,
	const double* poses, const uint8_t* cusp, double* out, int* iterations)
{
	auto* w = (TwinWorld*)hw;
	MapDev map = w->dev();
	SmoothDev P;
	P.stepTolerance = p->step_tolerance; P.maxIterations = p->max_iterations; P.learningRate = p->learning_rate;
	P.pathWeight = p->path_weight; P.smoothWeight = p->smooth_weight; P.voronoiWeight = p->voronoi_weight;
	P.collisionWeight = p->collision_weight; P.curvatureWeight = p->curvature_weight; P.collisionRatio = p->collision_ratio;
	P.maxCurvature = p->max_curvature; P.alpha = p->alpha; P.dMax = p->d_max; P.minSafeRadius = p->min_safe_radius;
	P.nearObstacle = nearObstacle; P.nearEdge = nearEdge;
	const float unsafeRadius = P.minSafeRadius * (1 + P.collisionRatio);
	std::vector<V2> pos(n), nxt(n), cm1(n), cc(n), cp1(n);
	std::vector<uint8_t> flags(n, 0);
	for (long i = 0; i < n; i++) {
		pos[i] = v2(poses[3 * i], poses[3 * i + 1]);
		flags[i] = smooth_is_index(cusp, (int)n, (int)i) ? 1 : 0;
	}
	int st, it = -1;
	if (n < 5) {
		st = 2;
	} else {
		float step = P.stepTolerance;
		while (true) {
			it++;
			if (it >= P.maxIterations) { st = 0; break; }
			if (step < P.stepTolerance) { st = 1; break; }
			for (long i = 0; i < n; i++) {
				if (!(flags[i] & 1)) continue;
				bool on = smooth_curvature(P, pos[i - 1], pos[i], pos[i + 1], cm1[i], cc[i], cp1[i]);
				flags[i] = on ? 3 : 1;
			}
			float mx = 0.0f;
			for (long i = 0; i < n; i++) {
				V2 g = v2(0.0, 0.0);
				if (i > 0 && flags[i - 1] == 3) g = g + cp1[i - 1];
				if (flags[i] & 1) {
					g = smooth_own_terms(P, map, g, v2(poses[3 * i], poses[3 * i + 1]), pos[i - 2], pos[i - 1], pos[i], pos[i + 1],
						pos[i + 2], unsafeRadius);
					if (flags[i] == 3) g = g + cc[i];
				}
				if (i + 1 < n && flags[i + 1] == 3) g = g + cm1[i + 1];
				nxt[i] = pos[i] - (double)P.learningRate * g;
				mx = fmaxf(mx, (float)v2_norm(g));
			}
			step = mx;
			pos.swap(nxt);
		}
	}
	bool bad = false;
	for (long i = 0; i < n; i++) {
		float d;
		if (!is_state_valid(map, pos[i].x, pos[i].y, poses[3 * i + 2], P.minSafeRadius, &d)) bad = true;
		out[3 * i] = pos[i].x; out[3 * i + 1] = pos[i].y; out[3 * i + 2] = poses[3 * i + 2];
	}
	if (iterations) *iterations = it;
	return bad ? -1 : st;
}

void twin_mt64_uniform(uint64_t seed, long n, double* out)
{
	std::vector<unsigned long long> mt(312);
	int idx = 0;
	Mt64 rng { mt.data(), &idx };
	rng.seed(seed);
	for (long i = 0; i < n; i++)
		out[i] = rng.uniform01();
}

} // extern "C"
