// Resampling of a Hybrid A* solution into the pose array the smoother takes:
// the "Process graph search path" block of HybridAStar::SearchPath, algo/hybrid_a_star.cpp:260-291, over
// PathSE2CompositeNonHolonomic (paths/path_composite.h:43-100, path_composite.cpp:4-33).
//
// The solution chain of ppx_hybrid_search (root .. goal, one ppx_hybrid_path_state per node) describes the
// composite path: segment j is the action that reaches chain[j+1] from chain[j].pose -- a constant-steer
// primitive (steering slot, direction, length after truncation) or the Reeds-Shepp connection to the goal.
//   resample_plan   (one thread)  composite length, cusp ratios, the sampled ratios and the cusp flags
//   resample_pose   (any thread)  PathComposite::Interpolate(ratio) -> FindSegment + the segment's Interpolate
#pragma once
#include "ppx_hybrid.cuh"

namespace ppx {

constexpr int kMaxCusps = 64;

struct ResamplePath {
	const ppx_hybrid_path_state* chain; // len states
	int len;
	const double* finalLength; // [len - 1] PathInfo::finalLength of each segment (initLength = previous finalLength)
	double length; // composite m_length
	RSSeg rs; // segment of the Reeds-Shepp action (kind 1), if any
	double rsLength;
};

// std::set<double>::insert on a small sorted array; false when full
PPX_HD bool cusp_insert(double* set, int& n, double v)
{
	int i = 0;
	while (i < n && set[i] < v)
		i++;
	if (i < n && set[i] == v)
		return true;
	if (n >= kMaxCusps)
		return false;
	for (int k = n; k > i; k--)
		set[k] = set[k - 1];
	set[i] = v;
	n++;
	return true;
}

// Motion count of a segment, PathSegment::GetNumMotions (reeds_shepp.cpp:455-464)
PPX_HD int rs_num_motions(const RSSeg& s)
{
	int index = 0;
	while (s.valid(index)) {
		index++;
		if (index == 5)
			break;
	}
	return index;
}

// Sequential part.  finalLength must have room for len - 1 doubles; ratios are written to out[3 * i]
// (the pose array, overwritten by resample_pose afterwards), cusp[i] = 1 where i is in cuspIndices.
// Returns the number of samples, 0 without a path, -1 when maxPoints / kMaxCusps do not suffice.
PPX_HD int resample_plan(const HybridDev& P, const ppx_hybrid_path_state* chain, int len, const Pose& goal, float pathInterpolation,
	double* finalLength, ResamplePath& rp, double* out, uint8_t* cusp, int maxPoints)
{
	rp.chain = chain;
	rp.len = len;
	rp.finalLength = finalLength;
	rp.length = 0.0;
	rp.rsLength = 0.0;
	const int ns = len - 1;
	if (ns < 1)
		return 0;
	// PathComposite::PushBack (path_composite.h:43-50)
	for (int j = 0; j < ns; j++) {
		const ppx_hybrid_path_state& st = chain[j + 1];
		if (st.kind == 1) {
			RSSolution sol;
			Pose parent = { chain[j].pose[0], chain[j].pose[1], chain[j].pose[2] };
			rs_solve(parent, goal, P.rmin, P.revF, P.fwdF, P.swF, 0, sol);
			rp.rs = sol.seg;
			rp.rsLength = sol.mLength * P.rmin;
		}
		rp.length += st.length;
		finalLength[j] = rp.length;
	}
	// PathSE2CompositeNonHolonomic::GetCuspPointRatios (path_composite.cpp:4-27)
	double set[kMaxCusps];
	int nset = 0;
	{
		double length = 0.0;
		int prevDirection = 0;
		for (int j = 0; j < ns; j++) {
			const ppx_hybrid_path_state& st = chain[j + 1];
			const bool isRs = st.kind == 1;
			int d0 = isRs ? rs_direction_at(rp.rs, P.rmin, rp.rsLength, 0.0) : (int)st.direction;
			if (j == 0)
				prevDirection = d0;
			if (d0 != prevDirection)
				if (!cusp_insert(set, nset, length / rp.length))
					return -1;
			if (isRs && rp.rsLength != 0.0) { // PathReedsShepp::GetCuspPointRatios (path_reeds_shepp.cpp:96-119):
				// ratios relative to the Reeds-Shepp path itself go into the composite's set unchanged
				double l = rp.rs.len[0] * P.rmin;
				const int nm = rs_num_motions(rp.rs);
				for (int i = 1; i < nm; i++) {
					double ratio = l / rp.rsLength;
					if (ratio > 1.0)
						break;
					if (rp.rs.dir(i) != rp.rs.dir(i - 1))
						if (!cusp_insert(set, nset, ratio))
							return -1;
					l += rp.rs.len[i] * P.rmin;
				}
			}
			prevDirection = isRs ? rs_direction_at(rp.rs, P.rmin, rp.rsLength, 1.0) : (int)st.direction;
			length += st.length;
		}
	}
	if (!cusp_insert(set, nset, 0.0) || !cusp_insert(set, nset, 1.0))
		return -1;
	// hybrid_a_star.cpp:276-287
	const double pathLength = rp.length;
	int n = 0, it = 0;
	for (double length = 0.0; length <= pathLength; length += pathInterpolation) {
		if (n >= maxPoints)
			return -1;
		double cuspLength = (it < nset ? set[it] : 0.0) * pathLength;
		if (it < nset && length >= cuspLength - pathInterpolation / 2 && length < cuspLength + pathInterpolation / 2) {
			cusp[n] = 1;
			out[3 * n] = set[it];
			it++;
		} else {
			cusp[n] = 0;
			out[3 * n] = length / pathLength;
		}
		n++;
	}
	return n;
}

// PathComposite::Interpolate(ratio) (path_composite.h:52-62) with FindSegment (:83-98)
PPX_HD Pose resample_pose(const HybridDev& P, const ResamplePath& rp, double ratio)
{
	const int ns = rp.len - 1;
	const double length = ratio * rp.length;
	int j = 0; // std::upper_bound: first segment with length < finalLength
	while (j < ns && !(length < rp.finalLength[j]))
		j++;
	double pathRatio;
	if (j == ns) { // it == end(): FindSegment returns ratio 1.0, Interpolate steps back to the last segment
		j = ns - 1;
		pathRatio = 1.0;
	} else {
		const double segLength = rp.chain[j + 1].length;
		const double initLength = j > 0 ? rp.finalLength[j - 1] : 0.0;
		pathRatio = segLength == 0.0 ? 0.0 : (length - initLength) / segLength;
	}
	const ppx_hybrid_path_state& st = rp.chain[j + 1];
	const Pose start = { rp.chain[j].pose[0], rp.chain[j].pose[1], rp.chain[j].pose[2] };
	if (st.kind == 1)
		return rs_interpolate(start, rp.rs, P.rmin, rp.rsLength, pathRatio);
	// PathConstantSteer::Interpolate: ConstantSteer(m_init, steering, m_length * ratio, direction)
	CSteer p;
	const int deltaIdx = (int)st.slot >> 1;
	p.start = start;
	p.length = st.length;
	p.beta = P.beta[deltaIdx];
	p.cosBeta = P.cosBeta[deltaIdx];
	p.kappa = P.kappa[deltaIdx];
	p.dir = (int)st.direction;
	p.curved = fabs(p.kappa) > 1e-9;
	if (p.curved)
		PPX_SINCOS(p.beta + p.start.th, &p.sB0, &p.cB0);
	else
		PPX_SINCOS(p.start.th, &p.sT0, &p.cT0);
	return p.at(pathRatio);
}

}
