// Device-side restatement of the reference's geometry / path / validator arithmetic.
// Every function cites the reference lines it follows and keeps the reference's operation ORDER
// and its float/double mixing, because collision verdicts and `last` ratios are compared
// bit-for-bit with the reference.  Compiled with -fmad=false (no FMA contraction).
#pragma once
#include "ppx_hd.h"
#include "ppx.h"

// A translation unit whose warps diverge widely (the per-query search kernel) defines PPX_SINCOS as
// a single out-of-line copy to keep its instruction footprint small; the edge kernels inline it.
#ifndef PPX_SINCOS
#define PPX_SINCOS sincos
#endif
// Same for the other libm calls of the Reeds-Shepp solver (fmod alone inlines to ~70 instructions,
// 25 times over the 12 formulas).
#ifndef PPX_FMOD
#define PPX_FMOD fmod
#define PPX_ATAN2 atan2
#define PPX_ACOS acos
#define PPX_ASIN asin
#define PPX_SIN sin
#endif

#define PPX_PI 3.14159265358979323846
#define PPX_PI_2 1.57079632679489661923
#define PPX_2PI (2 * PPX_PI)
#define PPX_INF (__longlong_as_double(0x7ff0000000000000LL))

namespace ppx {

// World model as the kernels see it (passed by value as a kernel parameter).
struct MapDev {
	int rows, cols;
	float res; // OccupancyMap::resolution
	double resd; // (double)res: the reference divides a double by the float field
	double ox, oy; // m_worldGridOrigin
	double lox, loy; // m_localOrigin
	double lbx, lby, lbt, ubx, uby, ubt; // StateSpaceSE2::bounds
	const float* __restrict__ clearance; // (float)(sqrt((double)sqdist) * res), gvd.h:38
	const int* __restrict__ occupancy;
	const float* __restrict__ pathcost;
};

struct Pose {
	double x, y, th;
};

// Pose2<double>::WrapTheta, geometry/2dplane.h:36-45.  Non-finite / astronomically large
// headings would spin forever in the reference; the kernel returns them unchanged instead.
PPX_HD double wrap_theta(double t)
{
	if (!(fabs(t) < 1e15))
		return t;
	while (t > PPX_PI)
		t -= 2 * PPX_PI;
	while (t < -PPX_PI)
		t += 2 * PPX_PI;
	return t;
}

// static_cast<int>(double) as x86-64 cvttsd2si does it (out of range / NaN -> INT_MIN),
// occupancy_map.h:108-109.
PPX_HD int trunc_i32(double v)
{
	if (!(v > -2147483649.0 && v < 2147483648.0))
		return INT_MIN;
	return __double2int_rz(v);
}

// std::min(a,b) / std::max(a,b) with the library's comparison direction (matters for NaN).
PPX_HD double std_min(double a, double b) { return (b < a) ? b : a; }
PPX_HD float std_minf(float a, float b) { return (b < a) ? b : a; }
PPX_HD float std_maxf(float a, float b) { return (a < b) ? b : a; }

// Maths::Modulo, utils/maths.h:9-16
PPX_HD double modulo(double in, double mod)
{
	double out = PPX_FMOD(in, mod);
	if (out < 0)
		out += mod;
	return out;
}

// StateValidatorOccupancyMap::IsStateValid, state_validator_occupancy_map.cpp:15-26.
// Returns validity; *dist receives the clearance of the pose's cell when it is inside the map.
PPX_HD bool is_state_valid(const MapDev& m, double x, double y, double th, float minSafeRadius, float* dist)
{
	// Pose2d localState(WorldPositionToLocalPosition(position), theta): the ctor wraps theta
	double lx = x - m.lox;
	double ly = y - m.loy;
	double lth = wrap_theta(th);
	// WorldPositionToGridCell(position) (bounded), occupancy_map.h:106-117,180-183
	int row = trunc_i32((x - m.ox) / m.resd);
	int col = trunc_i32((y - m.oy) / m.resd);
	bool inside = row >= 0 && row < m.rows && col >= 0 && col < m.cols;
	// StateSpaceSE2::ValidateBounds, state_space_se2.cpp:15-25
	bool inBounds = !(lx < m.lbx || lx > m.ubx) && !(ly < m.lby || ly > m.uby) && !(lth < m.lbt || lth > m.ubt);
	if (!inBounds || !inside)
		return false;
	float d = PPX_LDG(m.clearance + (size_t)row * m.cols + col);
	*dist = d;
	return d >= minSafeRadius;
}

// The step of IsPathValid after a valid sample, state_validator_occupancy_map.cpp:54-65.
PPX_HD double next_step(const MapDev& m, double x, double y, float distance, float minSafeRadius, float minStep)
{
	// std::min({a,b,c,d}) keeps the first element unless a later one compares less
	double b0 = x - m.lbx;
	double b1 = m.ubx - x;
	double b2 = y - m.lby;
	double b3 = m.uby - y;
	double mn = b0;
	if (b1 < mn) mn = b1;
	if (b2 < mn) mn = b2;
	if (b3 < mn) mn = b3;
	float distToMapBorder = (float)mn;
	float deltaLength = distance - minSafeRadius;
	deltaLength = std_minf(deltaLength, distToMapBorder);
	return (double)std_maxf(deltaLength, minStep);
}

// ---------------------------------------------------------------------------------------------
// Reeds-Shepp path (paths/path_reeds_shepp.cpp)

struct RSSeg {
	double len[5];
	uint32_t code; // 4 bits per motion: steer | dir<<2
	PPX_HD int steer(int i) const { return (code >> (4 * i)) & 3; }
	PPX_HD int dir(int i) const { return (code >> (4 * i + 2)) & 3; }
	// Motion::IsValid, reeds_shepp.cpp:305-308
	PPX_HD bool valid(int i) const { return i < 5 && len[i] != PPX_INF && dir(i) != 2; }
};

// PathSegment::m_length as AddMotion accumulates it (reeds_shepp.cpp:319-335)
PPX_HD double rs_total_length(const RSSeg& s)
{
	double total = 0.0;
	for (int i = 0; i < 5; i++) {
		if (!s.valid(i))
			break;
		total += fabs(s.len[i]);
	}
	return total;
}

// Pose2 operator+ (2dplane.h:49-62) with the rotation of the left pose cached (sL, cL).
PPX_HD Pose pose_compose(const Pose& l, double sL, double cL, double rx, double ry, double rth)
{
	Pose out;
	out.x = (cL * rx + (-sL) * ry) + l.x;
	out.y = (sL * rx + cL * ry) + l.y;
	out.th = wrap_theta(wrap_theta(l.th + rth));
	return out;
}

// PathReedsShepp::Straight / Turn (path_reeds_shepp.cpp:121-151) applied to a knot pose whose
// sin/cos are cached.  `amount` = motion.length * motionRatio.
PPX_HD Pose rs_advance(const Pose& k, double sK, double cK, int steer, int dir, double amount, double r)
{
	if (steer == 1) { // Straight
		double length = amount;
		if (dir == 1)
			length = -length;
		length *= r;
		Pose e;
		e.x = k.x + length * cK;
		e.y = k.y + length * sK;
		e.th = wrap_theta(k.th); // Pose2d ctor
		return e;
	}
	double turnAngle = amount;
	if (dir == 1)
		turnAngle = -turnAngle;
	double phi = turnAngle / 2;
	double sinPhi, cosPhi;
	PPX_SINCOS(phi, &sinPhi, &cosPhi);
	double L = 2 * sinPhi * r;
	double x = L * cosPhi;
	double y = L * sinPhi;
	if (steer == 2) {
		y *= -1;
		turnAngle *= -1;
	}
	// start + Pose2d(x, y, turnAngle): the temporary's ctor wraps turnAngle first
	return pose_compose(k, sK, cK, x, y, wrap_theta(turnAngle));
}

// Incremental walker over an RS path.  PathReedsShepp::Interpolate (path_reeds_shepp.cpp:12-47)
// re-walks all motions from the start for every ratio; motions that end before the target are
// always applied with motionRatio == 1.0 exactly, so their end poses ("knots") do not depend on
// the target and can be kept while targets are non-decreasing ("knot caching").
struct RSWalker {
	Pose k; // pose at the start of motion m
	double sK, cK; // sin/cos of k.th
	double lenAcc; // `length` accumulated over completed motions
	int m;

	PPX_HD void init(const Pose& start)
	{
		k = start;
		PPX_SINCOS(k.th, &sK, &cK);
		lenAcc = 0.0;
		m = 0;
	}

	// Pose at `target` = ratio * totalLength (targets must be non-decreasing across calls).
	PPX_HD Pose at(const RSSeg& s, double r, double target)
	{
		while (true) {
			if (!s.valid(m))
				return k; // ran off the end: interp is the end of the last applied motion
			double motionLength = s.len[m] * r;
			if (motionLength == 0) {
				m++;
				continue;
			}
			double motionRatio = (target - lenAcc) / motionLength;
			motionRatio = std_min(motionRatio, 1.0);
			Pose p = rs_advance(k, sK, cK, s.steer(m), s.dir(m), s.len[m] * motionRatio, r);
			double lenEnd = lenAcc + motionLength;
			if (lenEnd >= target)
				return p;
			k = p;
			PPX_SINCOS(k.th, &sK, &cK);
			lenAcc = lenEnd;
			m++;
		}
	}
};

// StateValidatorOccupancyMap::IsPathValid on a PathReedsShepp,
// state_validator_occupancy_map.cpp:28-71.
PPX_HD bool rs_is_path_valid(const MapDev& map, const Pose& start, const RSSeg& seg, double r,
	float minSafeRadius, float minStep, float* last, unsigned* samples)
{
	const double pathLength = rs_total_length(seg) * r; // PathSegment::GetLength(r)
	unsigned ns = 0;
	float dist = 0.0f;
	if (pathLength == 0.0) {
		*last = 1.0f;
		bool ok = is_state_valid(map, start.x, start.y, start.th, minSafeRadius, &dist);
		if (samples) *samples = 1;
		return ok;
	}
	RSWalker w;
	w.init(start);
	double lastValidLength = 0.0;
	double length = 0.0;
	bool ok = true;
	while (length < pathLength) {
		double ratio = length / pathLength;
		Pose p = w.at(seg, r, ratio * pathLength);
		ns++;
		if (!is_state_valid(map, p.x, p.y, p.th, minSafeRadius, &dist)) {
			*last = (float)(lastValidLength / pathLength);
			ok = false;
			break;
		}
		lastValidLength = length;
		length += next_step(map, p.x, p.y, dist, minSafeRadius, minStep);
	}
	if (ok)
		*last = 1.0f;
	if (samples) *samples = ns;
	return ok;
}

// ---------------------------------------------------------------------------------------------
// Rectangular car footprint (BASELINE config 3).  The reference checks ONE disc (two `// TODO add offset`,
// state_validator_occupancy_map.cpp:23,60); this variant has no reference semantics and is defined here (ppx.h,
// ppx_footprint_params): the length x width rectangle, centred `rear_to_center` ahead of the pose, is covered by n discs
// of one radius rho along the heading; a pose is valid iff every disc centre passes IsStateValid with minSafeRadius = rho
// (bounds, map, clearance of its cell).  Along a path the sampling is the reference's sphere tracing on the tightest disc,
// de-rated by gamma = the largest ratio between the travel of a disc centre and the travel of the pose.
struct RectDev {
	int n; // discs
	double first, pitch; // offset of disc 0 along the heading, distance between disc centres (m)
	float rho; // common radius
	float gamma; // >= 1
	float minStep;
};

// validity of the pose and the smallest slack (clearance - rho) over the discs, the smallest distance to the bounds
PPX_HD bool rect_state_valid(const MapDev& m, const RectDev& R, double x, double y, double th, float* slack, float* border)
{
	double s, c;
	PPX_SINCOS(th, &s, &c);
	float minSlack = __int_as_float(0x7f800000), minBorder = __int_as_float(0x7f800000);
	bool ok = true;
	for (int i = 0; i < R.n; i++) {
		const double o = R.first + i * R.pitch;
		const double dx = x + o * c, dy = y + o * s;
		float d = 0.0f;
		if (!is_state_valid(m, dx, dy, th, R.rho, &d))
			ok = false;
		minSlack = std_minf(minSlack, d - R.rho);
		double b = dx - m.lbx;
		if (m.ubx - dx < b) b = m.ubx - dx;
		if (dy - m.lby < b) b = dy - m.lby;
		if (m.uby - dy < b) b = m.uby - dy;
		minBorder = std_minf(minBorder, (float)b);
	}
	*slack = minSlack;
	*border = minBorder;
	return ok;
}

PPX_HD bool rs_is_path_valid_rect(const MapDev& map, const RectDev& R, const Pose& start, const RSSeg& seg, double r, float* last,
	unsigned* samples)
{
	const double pathLength = rs_total_length(seg) * r;
	unsigned ns = 0;
	float slack = 0.0f, border = 0.0f;
	if (pathLength == 0.0) {
		*last = 1.0f;
		if (samples) *samples = 1;
		return rect_state_valid(map, R, start.x, start.y, start.th, &slack, &border);
	}
	RSWalker w;
	w.init(start);
	double lastValidLength = 0.0, length = 0.0;
	bool ok = true;
	while (length < pathLength) {
		Pose p = w.at(seg, r, (length / pathLength) * pathLength);
		ns++;
		if (!rect_state_valid(map, R, p.x, p.y, p.th, &slack, &border)) {
			*last = (float)(lastValidLength / pathLength);
			ok = false;
			break;
		}
		lastValidLength = length;
		float delta = std_minf(slack, border) / R.gamma;
		length += (double)std_maxf(delta, R.minStep);
	}
	if (ok)
		*last = 1.0f;
	if (samples) *samples = ns;
	return ok;
}

// ---------------------------------------------------------------------------------------------
// Lane state machine of the re-converging Reeds-Shepp edge checker ("lane refill").
//
// IsPathValid on an RS path is a chain of data-dependent samples, each of which walks a
// data-dependent number of motions.  Run as one thread per edge that costs 26 of 32 lanes
// (ncu r1a: 6.1 active threads per instruction).  Here the same arithmetic is cut into uniform
// MICRO-STEPS -- each is "apply (part of) one motion to the current knot", needs exactly one
// sincos, and ends either in a knot commit (the motion ended before the target) or in a sample
// (IsStateValid + next step).  A warp keeps 32 such lanes busy and refills finished lanes from
// its edge queue, so every lane executes the same instruction stream all the time.
//   pre()  -> argument of the one sincos of this micro-step
//   post() -> consumes sin/cos, returns true when the edge is finished (verdict/last/ns set)
// The sin/cos of a committed knot's heading follow from the angle-addition identities applied to
// the cached sin/cos of the previous knot and the turn's half-angle sin/cos that the chord formula
// already needs, instead of a second sincos of the new heading.
template <class LenStore>
struct RSLane {
	double pathLength;
	uint32_t code;
	Pose k;
	double sK, cK, lenAcc;
	int m; // -1: the micro-step that evaluates sincos(start.theta)
	double length, lastValidLength;
	unsigned ns;
	bool verdict;
	float last;
	// scratch pre() -> post()
	double amount, motionLength, target;
	int kind; // 0 init, 1 motion, 2 sample at the knot (ran past the last valid motion)

	PPX_HD int steer(int i) const { return (code >> (4 * i)) & 3; }
	PPX_HD int dir(int i) const { return (code >> (4 * i + 2)) & 3; }
	PPX_HD bool valid(const LenStore& L, int i) const { return i < 5 && dir(i) != 2 && L.get(i) != PPX_INF; }

	// `L` must already hold the 5 normalised motion lengths of the edge.
	PPX_HD void begin(const LenStore& L, const Pose& start, uint32_t code_, double r)
	{
		code = code_;
		k = start;
		double total = 0.0; // PathSegment::m_length as AddMotion accumulates it
		for (int i = 0; i < 5; i++) {
			if (!valid(L, i))
				break;
			total += fabs(L.get(i));
		}
		pathLength = total * r;
		lenAcc = 0.0;
		m = -1;
		length = 0.0;
		lastValidLength = 0.0;
		ns = 0;
	}

	PPX_HD double pre(const LenStore& L, double r)
	{
		if (m < 0) {
			kind = 0;
			return k.th;
		}
		target = (length / pathLength) * pathLength; // Interpolate(length / pathLength): ratio * totalLength
		while (valid(L, m) && L.get(m) * r == 0)
			m++; // `continue` on zero-length motions, path_reeds_shepp.cpp:27-28
		if (!valid(L, m)) {
			kind = 2;
			return 0.0;
		}
		kind = 1;
		double len = L.get(m);
		motionLength = len * r;
		double motionRatio = (target - lenAcc) / motionLength;
		motionRatio = std_min(motionRatio, 1.0);
		amount = len * motionRatio;
		if (steer(m) == 1)
			return 0.0;
		double turnAngle = amount;
		if (dir(m) == 1)
			turnAngle = -turnAngle;
		return turnAngle / 2;
	}

	PPX_HD bool sample(const MapDev& map, const Pose& p, float minSafeRadius, float minStep)
	{
		float dist = 0.0f;
		ns++;
		if (!is_state_valid(map, p.x, p.y, p.th, minSafeRadius, &dist)) {
			verdict = false;
			last = (float)(lastValidLength / pathLength);
			return true;
		}
		lastValidLength = length;
		length += next_step(map, p.x, p.y, dist, minSafeRadius, minStep);
		if (length < pathLength)
			return false;
		verdict = true;
		last = 1.0f;
		return true;
	}

	PPX_HD bool post(const MapDev& map, double s, double c, double r, float minSafeRadius, float minStep)
	{
		if (kind == 0) {
			sK = s;
			cK = c;
			m = 0;
			if (pathLength == 0.0) { // state_validator_occupancy_map.cpp:34-38
				float dist;
				ns = 1;
				last = 1.0f;
				verdict = is_state_valid(map, k.x, k.y, k.th, minSafeRadius, &dist);
				return true;
			}
			return false;
		}
		if (kind == 2)
			return sample(map, k, minSafeRadius, minStep);
		const int st = steer(m), dr = dir(m);
		Pose p;
		double sT = 0.0, cT = 1.0;
		if (st == 1) { // PathReedsShepp::Straight, path_reeds_shepp.cpp:121-133
			double len = amount;
			if (dr == 1)
				len = -len;
			len *= r;
			p.x = k.x + len * cK;
			p.y = k.y + len * sK;
			p.th = wrap_theta(k.th);
		} else { // PathReedsShepp::Turn, :135-151
			double turnAngle = amount;
			if (dr == 1)
				turnAngle = -turnAngle;
			double chord = 2 * s * r;
			double x = chord * c;
			double y = chord * s;
			sT = 2 * s * c;
			cT = c * c - s * s;
			if (st == 2) {
				y *= -1;
				turnAngle *= -1;
				sT = -sT;
			}
			p = pose_compose(k, sK, cK, x, y, wrap_theta(turnAngle));
		}
		double lenEnd = lenAcc + motionLength;
		if (lenEnd >= target)
			return sample(map, p, minSafeRadius, minStep);
		// the motion ends before the target: commit its end pose as the next knot
		k = p;
		double sN = sK * cT + cK * sT;
		double cN = cK * cT - sK * sT;
		sK = sN;
		cK = cN;
		lenAcc = lenEnd;
		m++;
		return false;
	}
};

// ---------------------------------------------------------------------------------------------
// Constant-steer primitive (paths/path_constant_steer.cpp, models/kinematic_bicycle_model.cpp:5-32)

struct CSteer {
	Pose start;
	double length, beta, cosBeta, kappa;
	int dir;
	double sB0, cB0; // sin/cos(beta + start.th), cached
	bool curved; // abs(DthetaDdist) > 1e-9
	double sT0, cT0; // sin/cos(start.th) for the straight branch

	PPX_HD void init(const ppx_csteer_edge& e)
	{
		start = { e.start[0], e.start[1], e.start[2] };
		length = e.length;
		beta = e.beta;
		cosBeta = e.cos_beta;
		kappa = e.dtheta_ddist;
		dir = e.direction;
		curved = fabs(kappa) > 1e-9;
		if (curved)
			PPX_SINCOS(beta + start.th, &sB0, &cB0);
		else
			PPX_SINCOS(start.th, &sT0, &cT0);
	}

	// PathConstantSteer::Interpolate -> KinematicBicycleModel::ConstantSteer(m_init, steering, m_length*ratio, dir)
	PPX_HD Pose at_dist(double dist) const
	{
		if (dir == 1)
			dist = -dist;
		dist = dist / cosBeta;
		Pose to = start;
		if (curved) {
			to.th += dist * kappa;
			double s1, c1;
			PPX_SINCOS(beta + to.th, &s1, &c1);
			to.x += 1 / kappa * (s1 - sB0);
			to.y += 1 / kappa * (-c1 + cB0);
		} else {
			to.x += dist * cT0;
			to.y += dist * sT0;
		}
		return to;
	}
	PPX_HD Pose at(double ratio) const { return at_dist(length * ratio); }
};

PPX_HD bool csteer_is_path_valid(const MapDev& map, const CSteer& p, float minSafeRadius, float minStep,
	float* last, unsigned* samples)
{
	const double pathLength = p.length;
	float dist = 0.0f;
	unsigned ns = 0;
	if (pathLength == 0.0) {
		*last = 1.0f;
		if (samples) *samples = 1;
		return is_state_valid(map, p.start.x, p.start.y, p.start.th, minSafeRadius, &dist);
	}
	double lastValidLength = 0.0, length = 0.0;
	bool ok = true;
	while (length < pathLength) {
		Pose s = p.at(length / pathLength);
		ns++;
		if (!is_state_valid(map, s.x, s.y, s.th, minSafeRadius, &dist)) {
			*last = (float)(lastValidLength / pathLength);
			ok = false;
			break;
		}
		lastValidLength = length;
		length += next_step(map, s.x, s.y, dist, minSafeRadius, minStep);
	}
	if (ok)
		*last = 1.0f;
	if (samples) *samples = ns;
	return ok;
}

// ---------------------------------------------------------------------------------------------
// Straight SE(2) segment (paths/path_se2.cpp:5-22)
PPX_HD bool line_is_path_valid(const MapDev& map, const Pose& a, const Pose& b, float minSafeRadius,
	float minStep, float* last)
{
	double dx = b.x - a.x, dy = b.y - a.y;
	const double pathLength = sqrt(dx * dx + dy * dy);
	float dist = 0.0f;
	if (pathLength == 0.0) {
		*last = 1.0f;
		return is_state_valid(map, a.x, a.y, a.th, minSafeRadius, &dist);
	}
	double lastValidLength = 0.0, length = 0.0;
	while (length < pathLength) {
		double ratio = length / pathLength;
		double x = (1 - ratio) * a.x + ratio * b.x;
		double y = (1 - ratio) * a.y + ratio * b.y;
		double th = (1 - ratio) * a.th + ratio * b.th;
		if (!is_state_valid(map, x, y, th, minSafeRadius, &dist)) {
			*last = (float)(lastValidLength / pathLength);
			return false;
		}
		lastValidLength = length;
		length += next_step(map, x, y, dist, minSafeRadius, minStep);
	}
	*last = 1.0f;
	return true;
}

// ---------------------------------------------------------------------------------------------
// Reeds-Shepp solver (geometry/reeds_shepp.cpp)

PPX_HD bool angle_invalid(double t) { return t < 0 || t > PPX_PI; } // :13-16
PPX_HD double wrap_angle(double t) { return modulo(t + PPX_PI, 2 * PPX_PI) - PPX_PI; } // :18-21

// Per base word: 5 x (steer|dir<<2) nibbles and 5 x length selectors (0 t, 1 u, 2 v, 3 pi/2, 7 none).
// Built from reeds_shepp.cpp:299-... "Path" builders (:162-297).
//                         L=0 S=1 R=2 ; F=0 B=1
#define PPX_M(steer, dir) ((steer) | ((dir) << 2))
#define PPX_NONE 0xA /* steer=2, dir=2(NoMotion) */
PPX_CONST unsigned char c_rs_base_motion[12][5] = {
	{ PPX_M(0, 0), PPX_M(1, 0), PPX_M(0, 0), PPX_NONE, PPX_NONE }, // LfSfLf
	{ PPX_M(0, 0), PPX_M(1, 0), PPX_M(2, 0), PPX_NONE, PPX_NONE }, // LfSfRf
	{ PPX_M(0, 0), PPX_M(2, 1), PPX_M(0, 0), PPX_NONE, PPX_NONE }, // LfRbLf
	{ PPX_M(0, 0), PPX_M(2, 1), PPX_M(0, 1), PPX_NONE, PPX_NONE }, // LfRbLb
	{ PPX_M(0, 0), PPX_M(2, 0), PPX_M(0, 1), PPX_NONE, PPX_NONE }, // LfRfLb
	{ PPX_M(0, 0), PPX_M(2, 0), PPX_M(0, 1), PPX_M(2, 1), PPX_NONE }, // LfRufLubRb
	{ PPX_M(0, 0), PPX_M(2, 1), PPX_M(0, 1), PPX_M(2, 0), PPX_NONE }, // LfRubLubRf
	{ PPX_M(0, 0), PPX_M(2, 1), PPX_M(1, 1), PPX_M(0, 1), PPX_NONE }, // LfRbpi2SbLb
	{ PPX_M(0, 0), PPX_M(2, 1), PPX_M(1, 1), PPX_M(2, 1), PPX_NONE }, // LfRbpi2SbRb
	{ PPX_M(0, 0), PPX_M(1, 0), PPX_M(2, 0), PPX_M(0, 1), PPX_NONE }, // LfSfRfpi2Lb
	{ PPX_M(0, 0), PPX_M(1, 0), PPX_M(0, 0), PPX_M(2, 1), PPX_NONE }, // LfSfLfpi2Rb
	{ PPX_M(0, 0), PPX_M(2, 1), PPX_M(1, 1), PPX_M(0, 1), PPX_M(2, 0) }, // LfRbpi2SbLbpi2Rf
};
PPX_CONST unsigned char c_rs_base_len[12][5] = {
	{ 0, 1, 2, 7, 7 }, { 0, 1, 2, 7, 7 }, { 0, 1, 2, 7, 7 }, { 0, 1, 2, 7, 7 }, { 0, 1, 2, 7, 7 },
	{ 0, 1, 1, 2, 7 }, { 0, 1, 1, 2, 7 },
	{ 0, 3, 1, 2, 7 }, { 0, 3, 1, 2, 7 },
	{ 0, 1, 3, 2, 7 }, { 0, 1, 3, 2, 7 },
	{ 0, 3, 1, 3, 2 },
};

// GetPath(word, t, u, v) (:571-606): motions of base word w/4, then Timeflip (w%4 & 1) and
// Reflect (w%4 & 2) transforms (:495-516).  Returns m_length.
PPX_HD double rs_build_segment(int word, double t, double u, double v, RSSeg& seg)
{
	const int base = word >> 2, q = word & 3;
	double total = 0.0;
	uint32_t code = 0;
	#pragma unroll
	for (int i = 0; i < 5; i++) {
		unsigned mcode = c_rs_base_motion[base][i];
		unsigned sel = c_rs_base_len[base][i];
		double len = sel == 0 ? t : sel == 1 ? u : sel == 2 ? v : sel == 3 ? PPX_PI_2 : PPX_INF;
		if (sel != 7) {
			unsigned steer = mcode & 3, dir = mcode >> 2;
			if (q & 1)
				dir ^= 1; // TimeflipTransform
			if ((q & 2) && steer != 1)
				steer ^= 2; // ReflectTransform (Left 0 <-> Right 2)
			mcode = steer | (dir << 2);
			total += fabs(len);
		}
		seg.len[i] = len;
		code |= mcode << (4 * i);
	}
	seg.code = code;
	return total;
}

// PathSegment::ComputeCost (:347-376)
PPX_HD float rs_compute_cost(const RSSeg& seg, double mLength, double r, float rev, float fwd, float sw)
{
	if (!seg.valid(0))
		return __int_as_float(0x7f800000);
	if (rev == 1.0f && sw == 0.0f)
		return (float)(mLength * r);
	float cost = 0;
	int prevDirection = seg.dir(0);
	for (int i = 0; i < 5; i++) {
		if (!seg.valid(i))
			break;
		float motionCost = (float)(seg.len[i] * r);
		int d = seg.dir(i);
		if (d == 0)
			motionCost *= fwd;
		else if (d == 1)
			motionCost *= rev;
		if (d != prevDirection)
			motionCost += sw;
		prevDirection = d;
		cost += motionCost;
	}
	return cost;
}

// Intermediate quantities shared by the 6 formulas of one family on one mirrored goal.
struct RSGoalTerms {
	double gth; // goal.theta of the mirrored goal
	double xi, eta; // x -/+ sin(theta), y - 1 +/- cos(theta)
	double u1sq, u1, phi; // xi^2+eta^2, sqrt, PPX_ATAN2(eta, xi)
};

// GetMotionLengths (:537-569) for base word `base` on the mirrored goal.  `a` holds the terms of
// the "minus" family (x - sin, y - 1 + cos), `b` those of the "plus" family (x + sin, y - 1 - cos).
PPX_HD double rs_motion_lengths(int base, const RSGoalTerms& a, const RSGoalTerms& b, double& t, double& u, double& v)
{
	const double INF = PPX_INF;
	switch (base) {
	case 0: { // LfSfLf :23-37
		u = a.u1;
		t = a.phi;
		v = wrap_angle(a.gth - t);
		if (angle_invalid(t) || angle_invalid(v))
			return INF;
		return t + u + v;
	}
	case 1: { // LfSfRf :39-59 (the `u1squared < 4` branch has no return: NaN flows on)
		u = sqrt(b.u1sq - 4);
		double phi = PPX_ATAN2(2.0, u);
		t = wrap_angle(b.phi + phi);
		v = wrap_angle(t - b.gth);
		if (angle_invalid(t) || angle_invalid(v))
			return INF;
		return t + u + v;
	}
	case 2: { // LfRbLf :61-84
		if (a.u1 > 4)
			return INF;
		double alpha = PPX_ACOS(a.u1 / 4.0);
		t = modulo(PPX_PI_2 + alpha + a.phi, 2 * PPX_PI);
		u = modulo(PPX_PI - 2 * alpha, 2 * PPX_PI);
		v = modulo(a.gth - t - u, 2 * PPX_PI);
		if (angle_invalid(t) || angle_invalid(u) || angle_invalid(v))
			return INF;
		return t + u + v;
	}
	case 3: { // LfRbLb :86-106
		if (a.u1 > 4)
			return INF;
		double alpha = PPX_ACOS(a.u1 / 4.0);
		t = modulo(PPX_PI_2 + alpha + a.phi, 2 * PPX_PI);
		u = modulo(PPX_PI - 2 * alpha, 2 * PPX_PI);
		v = modulo(t + u - a.gth, 2 * PPX_PI);
		return t + u + v;
	}
	case 4: { // LfRfLb :108-129
		if (a.u1 > 4)
			return INF;
		u = PPX_ACOS((8 - a.u1 * a.u1) / 8.0);
		double va = PPX_SIN(u);
		double alpha = PPX_ASIN(2 * va / a.u1);
		t = modulo(PPX_PI_2 - alpha + a.phi, 2 * PPX_PI);
		v = modulo(t - u - a.gth, 2 * PPX_PI);
		return t + u + v;
	}
	case 5: { // LfRufLubRb :131-160
		if (b.u1 > 4)
			return INF;
		if (b.u1 > 2) {
			double alpha = PPX_ACOS(b.u1 / 4 - 0.5);
			t = modulo(PPX_PI_2 + b.phi - alpha, 2 * PPX_PI);
			u = modulo(PPX_PI - alpha, 2 * PPX_PI);
			v = modulo(b.gth - t + 2 * u, 2 * PPX_PI);
		} else {
			double alpha = PPX_ACOS(b.u1 / 4 + 0.5);
			t = modulo(PPX_PI_2 + b.phi + alpha, 2 * PPX_PI);
			u = modulo(alpha, 2 * PPX_PI);
			v = modulo(b.gth - t + 2 * u, 2 * PPX_PI);
		}
		return t + u + u + v;
	}
	case 6: { // LfRubLubRf :162-186
		if (b.u1 > 6)
			return INF;
		double va1 = 1.25 - b.u1 * b.u1 / 16; // 1.25f literal is exact
		if (va1 < 0 || va1 > 1)
			return INF;
		u = PPX_ACOS(va1);
		double va2 = PPX_SIN(u);
		double alpha = PPX_ASIN(2 * va2 / b.u1);
		t = modulo(PPX_PI_2 + b.phi + alpha, 2 * PPX_PI);
		v = modulo(t - b.gth, 2 * PPX_PI);
		return t + u + u + v;
	}
	case 7: { // LfRbpi2SbLb :188-211
		if (a.u1sq < 4)
			return INF;
		u = sqrt(a.u1sq - 4) - 2;
		if (u < 0)
			return INF;
		double alpha = PPX_ATAN2(2.0, u + 2);
		t = modulo(PPX_PI_2 + a.phi + alpha, 2 * PPX_PI);
		v = modulo(t + PPX_PI_2 - a.gth, 2 * PPX_PI);
		return t + PPX_PI_2 + u + v;
	}
	case 8: { // LfRbpi2SbRb :213-232
		if (b.u1 < 2)
			return INF;
		t = modulo(PPX_PI_2 + b.phi, 2 * PPX_PI);
		u = b.u1 - 2;
		v = modulo(b.gth - t - PPX_PI_2, 2 * PPX_PI);
		return t + PPX_PI_2 + u + v;
	}
	case 9: { // LfSfRfpi2Lb :234-257
		if (a.u1sq < 4)
			return INF;
		u = sqrt(a.u1sq - 4) - 2;
		if (u < 0)
			return INF;
		double alpha = PPX_ATAN2(u + 2, 2.0);
		t = modulo(PPX_PI_2 + a.phi - alpha, 2 * PPX_PI);
		v = modulo(t - PPX_PI_2 - a.gth, 2 * PPX_PI);
		return t + u + PPX_PI_2 + v;
	}
	case 10: { // LfSfLfpi2Rb :259-278
		if (b.u1 < 2)
			return INF;
		t = modulo(b.phi, 2 * PPX_PI);
		u = b.u1 - 2;
		v = modulo(-t - PPX_PI_2 + b.gth, 2 * PPX_PI);
		return t + u + PPX_PI_2 + v;
	}
	default: { // LfRbpi2SbLbpi2Rf :280-303
		if (b.u1sq < 16)
			return INF;
		u = sqrt(b.u1sq - 4) - 4;
		if (u < 0)
			return INF;
		double alpha = PPX_ATAN2(2.0, u + 4);
		t = modulo(PPX_PI_2 + b.phi + alpha, 2 * PPX_PI);
		v = modulo(t - b.gth, 2 * PPX_PI);
		return t + u + v + PPX_PI;
	}
	}
}

struct RSSolution {
	int word; // PathWords, -1 = NoPath
	RSSeg seg;
	double mLength; // PathSegment::m_length (normalised)
	float cost;
};

// Solver::GetOptimalPath (:654-683) when shortest == 0, Solver::GetShortestPath (:608-652) else.
PPX_HD void rs_solve(const Pose& start, const Pose& goal, double r, float rev, float fwd, float sw,
	int shortest, RSSolution& out)
{
	// GetGoalArray (:518-535): newGoal = goal - start (2dplane.h:64-78), scaled by 1/r
	double dx = goal.x - start.x, dy = goal.y - start.y;
	double sN, cN;
	PPX_SINCOS(-start.th, &sN, &cN);
	double gx = cN * dx + (-sN) * dy;
	double gy = sN * dx + cN * dy;
	double gth = wrap_theta(wrap_theta(goal.th - start.th));
	gx = gx / r;
	gy = gy / r;
	double sG, cG;
	PPX_SINCOS(gth, &sG, &cG);

	float bestCost = __int_as_float(0x7f800000);
	double bestLen = PPX_INF;
	int bestWord = -1;
	double bt = 0, bu = 0, bv = 0;

	#pragma unroll 1
	for (int q = 0; q < 4; q++) {
		// goals[q] = Pose2d(+-x, +-y, +-theta); sin(-theta) = -sin(theta), cos(-theta) = cos(theta)
		double x = (q & 1) ? -gx : gx;
		double y = (q & 2) ? -gy : gy;
		bool flipTh = (q == 1 || q == 2);
		double th = wrap_theta(flipTh ? -gth : gth);
		double s = flipTh ? -sG : sG;
		RSGoalTerms a, b;
		a.gth = b.gth = th;
		a.xi = x - s;
		a.eta = y - 1 + cG;
		b.xi = x + s;
		b.eta = y - 1 - cG;
		a.u1sq = a.xi * a.xi + a.eta * a.eta;
		b.u1sq = b.xi * b.xi + b.eta * b.eta;
		a.u1 = sqrt(a.u1sq);
		b.u1 = sqrt(b.u1sq);
		a.phi = PPX_ATAN2(a.eta, a.xi);
		b.phi = PPX_ATAN2(b.eta, b.xi);
		#pragma unroll 1
		for (int base = 0; base < 12; base++) {
			double t = 0, u = 0, v = 0;
			double length = rs_motion_lengths(base, a, b, t, u, v);
			int word = base * 4 + q;
			if (shortest) {
				// strict <, ties resolved by enum order: emulate by comparing (length, word)
				if (length < bestLen || (length == bestLen && word < bestWord)) {
					bestLen = length;
					bestWord = word;
					bt = t; bu = u; bv = v;
				}
			} else {
				if (length == PPX_INF)
					continue;
				RSSeg seg;
				double mLength = rs_build_segment(word, t, u, v, seg);
				float cost = rs_compute_cost(seg, mLength, r, rev, fwd, sw);
				if (cost < bestCost || (cost == bestCost && word < bestWord)) {
					bestCost = cost;
					bestWord = word;
					bt = t; bu = u; bv = v;
				}
			}
		}
	}
	out.word = bestWord;
	if (bestWord < 0) {
		// PathSegment(): all motions invalid, m_length 0
		for (int i = 0; i < 5; i++)
			out.seg.len[i] = PPX_INF;
		out.seg.code = 0xAAAAA;
		out.mLength = 0.0;
		out.cost = __int_as_float(0x7f800000);
		return;
	}
	out.mLength = rs_build_segment(bestWord, bt, bu, bv, out.seg);
	out.cost = rs_compute_cost(out.seg, out.mLength, r, rev, fwd, sw);
}


#ifdef __CUDACC__
// Solver::GetOptimalPath / GetShortestPath for EIGHT pairs per warp: the four lanes of a quad take the four mirrored goals
// of one pair (goal, timeflip, reflect, both: reeds_shepp.cpp:518-535), every lane walks the 12 base formulas in enum
// order -- the same formula on all 32 lanes at the same time, so the warp does not diverge on the formula switch -- and the
// quad takes the lexicographic minimum of (float cost, word), which is the reference's "first strict minimum in enum
// order" (word = base * 4 + mirror).  All four lanes of a quad return the pair's solution.  Must be called converged.
// rs_quad_part: the same over the base formulas baseFirst, baseFirst + baseStep, ... < 12 only, returning the quad's partial
// minimum (bestWord = 64: none) -- several warps can share the 12 formulas of a pair and combine their parts by the same
// lexicographic rule (k_rrt: one base formula per warp).
__device__ __forceinline__ void rs_quad_part(const Pose& start, const Pose& goal, double r, float rev, float fwd, float sw,
	int shortest, int baseFirst, int baseStep, float& bestCost, double& bestLen, int& bestWord, double& bt, double& bu, double& bv)
{
	const int q = (int)(threadIdx.x & 3);
	double dx = goal.x - start.x, dy = goal.y - start.y;
	double sN, cN;
	PPX_SINCOS(-start.th, &sN, &cN);
	double gx = cN * dx + (-sN) * dy;
	double gy = sN * dx + cN * dy;
	double gth = wrap_theta(wrap_theta(goal.th - start.th));
	gx = gx / r;
	gy = gy / r;
	double sG, cG;
	PPX_SINCOS(gth, &sG, &cG);
	const double x = (q & 1) ? -gx : gx;
	const double y = (q & 2) ? -gy : gy;
	const bool flipTh = (q == 1 || q == 2);
	const double th = wrap_theta(flipTh ? -gth : gth);
	const double sq = flipTh ? -sG : sG;
	RSGoalTerms a, b;
	a.gth = b.gth = th;
	a.xi = x - sq;
	a.eta = y - 1 + cG;
	b.xi = x + sq;
	b.eta = y - 1 - cG;
	a.u1sq = a.xi * a.xi + a.eta * a.eta;
	b.u1sq = b.xi * b.xi + b.eta * b.eta;
	a.u1 = sqrt(a.u1sq);
	b.u1 = sqrt(b.u1sq);
	a.phi = PPX_ATAN2(a.eta, a.xi);
	b.phi = PPX_ATAN2(b.eta, b.xi);
	const float INF = __int_as_float(0x7f800000);
	bestCost = INF; // GetOptimalPath: float cost; GetShortestPath: the double length (kept in bestLen)
	bestLen = PPX_INF;
	bestWord = 64;
	bt = 0, bu = 0, bv = 0;
	#pragma unroll 1
	for (int base = baseFirst; base < 12; base += baseStep) {
		double t = 0, u = 0, v = 0;
		const double length = rs_motion_lengths(base, a, b, t, u, v);
		const int word = base * 4 + q;
		if (shortest) {
			if (length < bestLen) { // words of one lane come in increasing order: strict < keeps the first
				bestLen = length;
				bestWord = word;
				bt = t; bu = u; bv = v;
			}
		} else if (length != PPX_INF) {
			RSSeg seg;
			const double mLength = rs_build_segment(word, t, u, v, seg);
			const float cost = rs_compute_cost(seg, mLength, r, rev, fwd, sw);
			if (cost < bestCost) {
				bestCost = cost;
				bestWord = word;
				bt = t; bu = u; bv = v;
			}
		}
	}
	#pragma unroll
	for (int off = 1; off <= 2; off <<= 1) {
		const float oc = __shfl_xor_sync(0xffffffffu, bestCost, off);
		const double ol = __shfl_xor_sync(0xffffffffu, bestLen, off);
		const int ow = __shfl_xor_sync(0xffffffffu, bestWord, off);
		const double t2 = __shfl_xor_sync(0xffffffffu, bt, off);
		const double u2 = __shfl_xor_sync(0xffffffffu, bu, off);
		const double v2 = __shfl_xor_sync(0xffffffffu, bv, off);
		const bool take = shortest ? (ol < bestLen || (ol == bestLen && ow < bestWord)) : (oc < bestCost || (oc == bestCost && ow < bestWord));
		if (take) {
			bestCost = oc;
			bestLen = ol;
			bestWord = ow;
			bt = t2; bu = u2; bv = v2;
		}
	}
}

__device__ __forceinline__ void rs_solve_quad(const Pose& start, const Pose& goal, double r, float rev, float fwd, float sw,
	int shortest, int& outWord, double& ot, double& ou, double& ov)
{
	float bestCost;
	double bestLen, bt, bu, bv;
	int bestWord;
	rs_quad_part(start, goal, r, rev, fwd, sw, shortest, 0, 1, bestCost, bestLen, bestWord, bt, bu, bv);
	outWord = bestWord < 48 ? bestWord : -1;
	ot = bt; ou = bu; ov = bv;
}

// Solver::GetOptimalPath evaluated by a whole warp: the 48 words are independent, lane L takes word
// L and (lanes 0..15) word 32 + L.  The reference's scan in enum order with a strict `<` keeps the
// first word of minimal float cost, i.e. the minimum of (cost, word) in lexicographic order over
// the words whose cost compares below +inf (NaN and +inf never win, trap H8); the warp takes the
// same minimum by shuffles.  All lanes return the same solution.  Must be called converged.
__device__ __forceinline__ void rs_solve_warp(const Pose& start, const Pose& goal, double r, float rev, float fwd, float sw,
	RSSolution& out)
{
	const unsigned lane = threadIdx.x & 31;
	double dx = goal.x - start.x, dy = goal.y - start.y;
	double sN, cN;
	PPX_SINCOS(-start.th, &sN, &cN);
	double gx = cN * dx + (-sN) * dy;
	double gy = sN * dx + cN * dy;
	double gth = wrap_theta(wrap_theta(goal.th - start.th));
	gx = gx / r;
	gy = gy / r;
	double sG, cG;
	PPX_SINCOS(gth, &sG, &cG);
	const int q = lane & 3;
	double x = (q & 1) ? -gx : gx;
	double y = (q & 2) ? -gy : gy;
	bool flipTh = (q == 1 || q == 2);
	double th = wrap_theta(flipTh ? -gth : gth);
	double sq = flipTh ? -sG : sG;
	RSGoalTerms a, b;
	a.gth = b.gth = th;
	a.xi = x - sq;
	a.eta = y - 1 + cG;
	b.xi = x + sq;
	b.eta = y - 1 - cG;
	a.u1sq = a.xi * a.xi + a.eta * a.eta;
	b.u1sq = b.xi * b.xi + b.eta * b.eta;
	a.u1 = sqrt(a.u1sq);
	b.u1 = sqrt(b.u1sq);
	a.phi = PPX_ATAN2(a.eta, a.xi);
	b.phi = PPX_ATAN2(b.eta, b.xi);
	const float INF = __int_as_float(0x7f800000);
	float bestCost = INF;
	int bestWord = 64;
	double bt = 0, bu = 0, bv = 0;
	#pragma unroll 1
	for (int round = 0; round < 2; round++) {
		int word = round * 32 + (int)lane;
		if (word < 48) {
			double t = 0, u = 0, v = 0;
			double length = rs_motion_lengths(word >> 2, a, b, t, u, v);
			if (length != PPX_INF) {
				RSSeg seg;
				double mLength = rs_build_segment(word, t, u, v, seg);
				float cost = rs_compute_cost(seg, mLength, r, rev, fwd, sw);
				if (cost < bestCost) { // the lane's round-1 word is the larger one: strict < keeps enum order
					bestCost = cost;
					bestWord = word;
					bt = t; bu = u; bv = v;
				}
			}
		}
	}
	#pragma unroll
	for (int off = 16; off > 0; off >>= 1) {
		float oc = __shfl_xor_sync(0xffffffffu, bestCost, off);
		int ow = __shfl_xor_sync(0xffffffffu, bestWord, off);
		double ot = __shfl_xor_sync(0xffffffffu, bt, off);
		double ou = __shfl_xor_sync(0xffffffffu, bu, off);
		double ov = __shfl_xor_sync(0xffffffffu, bv, off);
		if (oc < bestCost || (oc == bestCost && ow < bestWord)) {
			bestCost = oc;
			bestWord = ow;
			bt = ot; bu = ou; bv = ov;
		}
	}
	if (!(bestCost < INF))
		bestWord = -1;
	out.word = bestWord;
	if (bestWord < 0) {
		for (int i = 0; i < 5; i++)
			out.seg.len[i] = PPX_INF;
		out.seg.code = 0xAAAAA;
		out.mLength = 0.0;
		out.cost = INF;
		return;
	}
	out.mLength = rs_build_segment(bestWord, bt, bu, bv, out.seg);
	out.cost = rs_compute_cost(out.seg, out.mLength, r, rev, fwd, sw);
}
#endif

}
