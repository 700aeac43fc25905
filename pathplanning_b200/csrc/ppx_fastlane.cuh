// Filtered ("certified float") lane machine of the batched Reeds-Shepp edge checker.
//
// What IsPathValid (state_validator_occupancy_map.cpp:28-71) returns for an edge -- verdict, last
// valid ratio, number of samples -- is a function of the SEQUENCE OF GRID CELLS its samples fall
// in: the clearance is a per-cell float, the step `max(min(clearance - r, distToBorder), minStep)`
// is float arithmetic on it, and `length` is a double sum of those floats.  The exact (FP64,
// reference operation order) evaluation of a sample position is therefore only needed to decide
// which cell the sample is in.  This machine evaluates sample positions in FP32 (MUFU sin/cos, cell
// units relative to the edge's first cell) and CERTIFIES the cell: a sample whose FP32 position lies
// more than kCellEps from every cell boundary, whose cell lies in the safe interior of map and
// bounds, and whose step is certainly not limited by the distance to the border has the same cell
// and the same step as the exact evaluation, bit for bit.  Every other sample (about 0.5 %) is PARKED
// and re-evaluated by the exact FP64 path (RSWalker, the reference's Interpolate restated), batched
// per warp.  kCellEps is 3x the worst-case FP32 error bound and 10x the largest error the
// device audit has seen (k_check_rs<audit>: max |fast - exact| over all certified samples, 9.7e-5 cells).
#pragma once
#include "ppx_device.cuh"

#ifdef __CUDACC__
#define PPX_FSIN(x) __sinf(x)
#define PPX_FCOS(x) __cosf(x)
#else
#define PPX_FSIN(x) sinf(x)
#define PPX_FCOS(x) cosf(x)
#endif

namespace ppx {

constexpr float kCellEps = 1.0e-3f; // cells; certified samples keep this distance from cell borders
constexpr float kFastMaxCells = 1024.0f; // longest path (in cells) the FP32 frame is used for

// Per-map constants of the filter (host-derived, passed by value).
struct FastMap {
	int rb0, rb1, cb0, cb1; // rows/cols whose cells lie entirely inside map AND bounds, one cell of margin
	float inv_res; // 1 / resolution
	double inv_resd;
	int enabled; // 0: every sample takes the exact path (non-trivial heading bounds, odd maps)
};

inline FastMap fast_map_derive(const ppx_map_desc& d)
{
	FastMap f {};
	double res = (double)d.resolution;
	// local position lx = x - local_origin must satisfy lb <= lx <= ub; in cell units from world_origin:
	double uLx = (d.lb[0] + d.local_origin[0] - d.world_origin[0]) / res;
	double uUx = (d.ub[0] + d.local_origin[0] - d.world_origin[0]) / res;
	double uLy = (d.lb[1] + d.local_origin[1] - d.world_origin[1]) / res;
	double uUy = (d.ub[1] + d.local_origin[1] - d.world_origin[1]) / res;
	// ... while the step of IsPathValid measures the distance to the border on the WORLD position
	// (`state.x() - bounds[0].x()`, state_validator_occupancy_map.cpp:54-59): the margin `ib` the filter
	// certifies a step against must be a lower bound in that frame too.  The two frames differ by
	// local_origin (OccupancyMap::SetPosition); the certified rows/cols are those inside both.
	double wLx = (d.lb[0] - d.world_origin[0]) / res, wUx = (d.ub[0] - d.world_origin[0]) / res;
	double wLy = (d.lb[1] - d.world_origin[1]) / res, wUy = (d.ub[1] - d.world_origin[1]) / res;
	bool finite = std::isfinite(uLx) && std::isfinite(uUx) && std::isfinite(uLy) && std::isfinite(uUy)
		&& std::isfinite(wLx) && std::isfinite(wUx) && std::isfinite(wLy) && std::isfinite(wUy);
	uLx = uLx > wLx ? uLx : wLx;
	uUx = uUx < wUx ? uUx : wUx;
	uLy = uLy > wLy ? uLy : wLy;
	uUy = uUy < wUy ? uUy : wUy;
	auto lo = [](double u) { double c = ceil(u) + 1; return c < 0 ? 0.0 : c; };
	auto hi = [](double u, int n) { double c = floor(u) - 2; return c > n - 1 ? (double)(n - 1) : c; };
	f.enabled = finite && d.lb[2] <= -PPX_PI && d.ub[2] >= PPX_PI && d.rows < (1 << 20) && d.cols < (1 << 20);
	if (f.enabled) {
		double a = lo(uLx), b = hi(uUx, d.rows), c = lo(uLy), e = hi(uUy, d.cols);
		if (a > 2e6) a = 2e6;
		if (c > 2e6) c = 2e6;
		if (b < -2e6) b = -2e6;
		if (e < -2e6) e = -2e6;
		f.rb0 = (int)a; f.rb1 = (int)b; f.cb0 = (int)c; f.cb1 = (int)e;
	}
	f.inv_res = (float)(1.0 / res);
	f.inv_resd = 1.0 / res;
	return f;
}

// reduce an angle in [-3pi, 3pi] to [-pi, pi] (two-term 2*pi)
PPX_HD float fast_wrap(float a)
{
	float k = rintf(a * 0.15915494309189535f);
	a = a - k * 6.28125f; // exact product for small integer k (6.28125 has 9 significant bits)
	a = a - k * 1.9353071795864769e-3f;
	return a;
}

PPX_HD int imin(int a, int b) { return a < b ? a : b; }

enum LaneState : int { LANE_IDLE = 0, LANE_ACTIVE = 1, LANE_PARKED = 2 };

template <class LenStore>
struct RSFastLane {
	double pathLength, length, lastValidLength, lenAcc;
	float kx, ky, thK; // current knot, cell units relative to (rowb, colb); heading wrapped
	uint32_t code;
	int m;
	int rowb, colb;
	unsigned ns;
	bool slowOnly;
	bool verdict;
	float last;
	float px, py; // last evaluated fast position (audit)
	// cached constants of motion m (load_motion)
	double lm;
	float ml, sgnAdv, sgnTurn;
	bool have, straight;

	PPX_HD int steer(int i) const { return (code >> (4 * i)) & 3; }
	PPX_HD int dir(int i) const { return (code >> (4 * i + 2)) & 3; }
	PPX_HD bool valid(const LenStore& L, int i) const { return i < 5 && dir(i) != 2 && L.get(i) != PPX_INF; }

	// `L` must already hold the 5 normalised motion lengths of the edge.
	PPX_HD void begin(const LenStore& L, const FastMap& fm, const MapDev& map, const Pose& start, uint32_t code_, double r)
	{
		code = code_;
		double total = 0.0; // PathSegment::m_length as AddMotion accumulates it
		bool odd = false;
		for (int i = 0; i < 5; i++) {
			if (!valid(L, i))
				break;
			double l = L.get(i);
			odd |= !(l >= 0.0); // negative or NaN lengths: exact path only
			total += fabs(l);
		}
		pathLength = total * r;
		length = 0.0;
		lastValidLength = 0.0;
		lenAcc = 0.0;
		m = 0;
		ns = 0;
		double ux = (start.x - map.ox) * fm.inv_resd;
		double uy = (start.y - map.oy) * fm.inv_resd;
		double th = start.th;
		odd |= !(fabs(ux) < 1e6) || !(fabs(uy) < 1e6) || !(fabs(th) < 1e4);
		odd |= !(pathLength > 0.0) || !(pathLength * fm.inv_resd < (double)kFastMaxCells) || !(r > 1e-3) || !fm.enabled;
		slowOnly = odd;
		load_motion(L, r);
		if (odd) {
			rowb = colb = 0;
			kx = ky = thK = 0.0f;
			return;
		}
		double fx = floor(ux), fy = floor(uy);
		rowb = (int)fx;
		colb = (int)fy;
		kx = (float)(ux - fx);
		ky = (float)(uy - fy);
		th = th - PPX_2PI * rint(th * (1.0 / PPX_2PI));
		thK = (float)th;
	}

	// Motion m's constants, cached between knot commits (the motion changes ~once per 3 micro-steps).
	PPX_HD void load_motion(const LenStore& L, double r)
	{
		have = valid(L, m);
		lm = have ? L.get(m) * r : 0.0; // motionLength, path_reeds_shepp.cpp:25
		ml = (float)lm;
		int st = steer(m), dr = dir(m);
		straight = (st == 1) || !have;
		sgnAdv = dr == 1 ? -1.0f : 1.0f; // Backward
		sgnTurn = st == 2 ? -1.0f : 1.0f; // Right
	}

	// One micro-step: advance along motion m towards the current sample target.  Returns
	//   0: knot committed (or nothing to do), 1: a sample position (px, py) is ready.
	PPX_HD int advance(const LenStore& L, double r, float inv_r, float rc, float inv_res)
	{
		float rem = (float)(length - lenAcc);
		bool full = have && rem > ml;
		float adv = sgnAdv * fminf(rem, ml); // ml == 0 without a motion
		float a = adv * inv_r * sgnTurn; // signed heading change of the (partial) motion
		float phi = 0.5f * a;
		float ang = fast_wrap(straight ? thK : thK + phi);
		float chord = (2.0f * rc) * PPX_FSIN(sgnTurn * phi);
		float Lc = straight ? adv * inv_res : chord;
		px = kx + Lc * PPX_FCOS(ang);
		py = ky + Lc * PPX_FSIN(ang);
		if (!full)
			return 1;
		kx = px;
		ky = py;
		if (!straight)
			thK = fast_wrap(thK + a);
		lenAcc += lm;
		m++;
		load_motion(L, r);
		return 0;
	}

	// Certified sample at (px, py).  Returns 0: continue, 1: edge finished (verdict/last set), 2: park.
	PPX_HD int sample(const MapDev& map, const FastMap& fm, float minSafeRadius, float minStep)
	{
		float fx = floorf(px), fy = floorf(py);
		float gx = px - fx, gy = py - fy;
		bool near = fminf(gx, gy) < kCellEps || fmaxf(gx, gy) > 1.0f - kCellEps;
		int row = rowb + (int)fx, col = colb + (int)fy;
		int ibr = imin(row - fm.rb0, fm.rb1 - row);
		int ibc = imin(col - fm.cb0, fm.cb1 - col);
		int ib = imin(ibr, ibc);
		if (near || ib < 0)
			return 2;
		float d = PPX_LDG(map.clearance + (size_t)row * map.cols + col);
		if (!(d >= minSafeRadius)) { // IsStateValid false
			ns++;
			verdict = false;
			// 0 / pathLength == 0 exactly; spares the division's slow path for the common first-sample failure
			last = lastValidLength == 0.0 ? 0.0f : (float)(lastValidLength / pathLength);
			return 1;
		}
		float deltaLength = d - minSafeRadius;
		if (!(deltaLength <= (float)ib * map.res))
			return 2; // the distance to the border may be the smaller one: exact path decides
		ns++;
		lastValidLength = length;
		length += (double)std_maxf(deltaLength, minStep);
		if (length < pathLength)
			return 0;
		verdict = true;
		last = 1.0f;
		return 1;
	}
};

// Exact evaluation of ONE sample of IsPathValid at the lane's current `length`: the loop body of
// state_validator_occupancy_map.cpp:40-66 with PathReedsShepp::Interpolate restated by RSWalker
// (walks the motions from the start pose, like the reference does for every ratio).
// Returns true when the edge is finished.  *ex, *ey receive the exact sample position (audit).
template <class Lane>
PPX_HD bool rs_exact_sample(Lane& lane, const MapDev& map, const Pose& start, const RSSeg& seg, double r,
	float minSafeRadius, float minStep, double* ex, double* ey)
{
	float dist = 0.0f;
	if (lane.pathLength == 0.0) { // :34-38
		lane.ns = 1;
		lane.last = 1.0f;
		lane.verdict = is_state_valid(map, start.x, start.y, start.th, minSafeRadius, &dist);
		*ex = start.x;
		*ey = start.y;
		return true;
	}
	if (!(lane.length < lane.pathLength)) { // NaN path length: the reference's loop does not run
		lane.verdict = true;
		lane.last = 1.0f;
		return true;
	}
	RSWalker w;
	w.init(start);
	double ratio = lane.length / lane.pathLength;
	Pose p = w.at(seg, r, ratio * lane.pathLength);
	*ex = p.x;
	*ey = p.y;
	lane.ns++;
	if (!is_state_valid(map, p.x, p.y, p.th, minSafeRadius, &dist)) {
		lane.verdict = false;
		lane.last = (float)(lane.lastValidLength / lane.pathLength);
		return true;
	}
	lane.lastValidLength = lane.length;
	lane.length += next_step(map, p.x, p.y, dist, minSafeRadius, minStep);
	if (lane.length < lane.pathLength)
		return false;
	lane.verdict = true;
	lane.last = 1.0f;
	return true;
}


// ---------------------------------------------------------------------------------------------
// IsPathValid on a constant-steer primitive (csteer_is_path_valid) with the same cell filter: the
// sample at arc length `length` is  start + chord * (cos, sin)(beta + theta0 + h),  h = dist*kappa/2,
// chord = dist * sin(h)/h  (the chord form of kinematic_bicycle_model.cpp:5-32; unlike the difference
// of sines it does not cancel for small curvatures).  Uncertified samples are evaluated exactly in
// place: one primitive per thread, nothing to batch.
PPX_HD bool csteer_is_path_valid_filtered(const MapDev& map, const FastMap& fm, const CSteer& p, float minSafeRadius,
	float minStep, float* last, unsigned* samples)
{
	const double pathLength = p.length;
	double ux = (p.start.x - map.ox) * fm.inv_resd;
	double uy = (p.start.y - map.oy) * fm.inv_resd;
	bool usable = fm.enabled && fabs(ux) < 1e6 && fabs(uy) < 1e6 && fabs(p.start.th) < 1e4 && pathLength > 0.0
		&& pathLength * fm.inv_resd < (double)kFastMaxCells && fabs(p.cosBeta) > 0.05 && fabs(p.kappa) < 1e3 && fabs(p.beta) < 10.0;
	if (!usable)
		return csteer_is_path_valid(map, p, minSafeRadius, minStep, last, samples);
	const double fx0 = floor(ux), fy0 = floor(uy);
	const int rowb = (int)fx0, colb = (int)fy0;
	const float kx = (float)(ux - fx0), ky = (float)(uy - fy0);
	double a0 = (p.curved ? p.beta : 0.0) + p.start.th; // the straight branch moves along theta0
	a0 = a0 - PPX_2PI * rint(a0 * (1.0 / PPX_2PI));
	const float A = (float)a0;
	const float kap = p.curved ? (float)p.kappa : 0.0f, invCos = (float)(1.0 / p.cosBeta);
	const float sgn = p.dir == 1 ? -1.0f : 1.0f;
	float dist = 0.0f;
	unsigned ns = 0;
	double lastValidLength = 0.0, length = 0.0;
	bool ok = true;
	while (length < pathLength) {
		float d = sgn * (float)length * invCos;
		float h = 0.5f * d * kap;
		float sinc = fabsf(h) < 1e-3f ? 1.0f - h * h * (1.0f / 6.0f) : PPX_FSIN(h) / h;
		float chord = d * fm.inv_res * sinc;
		float ang = fast_wrap(A + h);
		float px = kx + chord * PPX_FCOS(ang), py = ky + chord * PPX_FSIN(ang);
		float fx = floorf(px), fy = floorf(py);
		float gx = px - fx, gy = py - fy;
		bool near = fminf(gx, gy) < kCellEps || fmaxf(gx, gy) > 1.0f - kCellEps;
		int row = rowb + (int)fx, col = colb + (int)fy;
		int ib = imin(imin(row - fm.rb0, fm.rb1 - row), imin(col - fm.cb0, fm.cb1 - col));
		bool certified = !near && ib >= 0 && fabsf(h) < 3.0f;
		bool valid = false;
		float deltaLength = 0.0f;
		if (certified) {
			dist = PPX_LDG(map.clearance + (size_t)row * map.cols + col);
			valid = dist >= minSafeRadius;
			deltaLength = dist - minSafeRadius;
			if (valid && !(deltaLength <= (float)ib * map.res))
				certified = false; // the distance to the border may limit the step
		}
		ns++;
		if (certified) {
			if (!valid) {
				*last = lastValidLength == 0.0 ? 0.0f : (float)(lastValidLength / pathLength);
				ok = false;
				break;
			}
			lastValidLength = length;
			length += (double)std_maxf(deltaLength, minStep);
		} else { // exact sample, csteer_is_path_valid's loop body
			Pose s = p.at(length / pathLength);
			if (!is_state_valid(map, s.x, s.y, s.th, minSafeRadius, &dist)) {
				*last = (float)(lastValidLength / pathLength);
				ok = false;
				break;
			}
			lastValidLength = length;
			length += next_step(map, s.x, s.y, dist, minSafeRadius, minStep);
		}
	}
	if (ok)
		*last = 1.0f;
	if (samples) *samples = ns;
	return ok;
}


// rs_is_path_valid (IsPathValid on one Reeds-Shepp path, sequentially) through the same filter: the
// lane machine run by a single thread, uncertified samples evaluated exactly in place.
struct SegLen {
	const double* p;
	PPX_HD double get(int i) const { return p[i]; }
};

PPX_HD bool rs_is_path_valid_filtered(const MapDev& map, const FastMap& fm, const Pose& start, const RSSeg& seg, double r,
	float minSafeRadius, float minStep, float* last, unsigned* samples)
{
	SegLen L { seg.len };
	RSFastLane<SegLen> lane;
	lane.begin(L, fm, map, start, seg.code, r);
	const float inv_r = (float)(1.0 / r), rc = (float)(r * fm.inv_resd);
	while (true) {
		int rcode = 2;
		if (!lane.slowOnly) {
			if (!lane.advance(L, r, inv_r, rc, fm.inv_res))
				continue;
			rcode = lane.sample(map, fm, minSafeRadius, minStep);
		}
		if (rcode == 2) {
			double ex, ey;
			if (rs_exact_sample(lane, map, start, seg, r, minSafeRadius, minStep, &ex, &ey))
				break;
		} else if (rcode == 1)
			break;
	}
	*last = lane.last;
	if (samples) *samples = lane.ns;
	return lane.verdict;
}

}
