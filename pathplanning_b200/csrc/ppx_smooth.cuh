// Gradient-descent path smoother: Smoother::Smooth, algo/smoother.cpp:33-158, and
// CalculateCurvatureTerm, smoother.cpp:160-213, restated per point so that the points of a path can be
// evaluated in parallel while every gradient is summed in the reference's order.
//
// The reference walks `indices` in ascending order; loop index i adds to gradients[i] its path, collision,
// Voronoi and smoothing terms and then the curvature term adds to gradients[i-1], [i], [i+1].  gradients[j]
// therefore receives, in this order:  curvature(+1 part) of index j-1, the own terms of j, curvature(centre)
// of j, curvature(-1 part) of index j+1.  Phase A computes the three curvature parts of every index, phase B
// assembles each gradient in that order -- same operands, same order, so +,-,*,/ and sqrt give the reference's
// bits; acosf / cosf are libm (device and glibc differ in the last ulp, bar: 1e-6 on the poses).
//
// The reference mixes float parameters into double vectors; every conversion below follows the C++ usual
// arithmetic conversions of the original expressions (float op float -> float, float op double -> double,
// scalar * Point2d -> component * (double)scalar).
#pragma once
#include "ppx_device.cuh"

namespace ppx {

struct SmoothDev {
	float stepTolerance; // Smoother::Parameters, smoother.h:29-62
	int maxIterations;
	float learningRate, pathWeight, smoothWeight, voronoiWeight, collisionWeight, curvatureWeight, collisionRatio, maxCurvature;
	float alpha, dMax; // GVD::alpha, GVD::dMax (gvd.h:181)
	float minSafeRadius; // StateValidatorOccupancyMap::minSafeRadius
	const int* __restrict__ nearObstacle; // [rows * cols * 2] GVD::GetNearestObstacleCell (row, col)
	const int* __restrict__ nearEdge; // [rows * cols * 2] GVD::GetNearestVoronoiEdgeCell
};

struct V2 {
	double x, y;
};
PPX_HD V2 v2(double x, double y) { V2 r; r.x = x; r.y = y; return r; }
PPX_HD V2 operator+(V2 a, V2 b) { return v2(a.x + b.x, a.y + b.y); }
PPX_HD V2 operator-(V2 a, V2 b) { return v2(a.x - b.x, a.y - b.y); }
PPX_HD V2 operator-(V2 a) { return v2(-a.x, -a.y); }
PPX_HD V2 operator*(double s, V2 a) { return v2(a.x * s, a.y * s); }
PPX_HD V2 operator/(V2 a, double s) { return v2(a.x / s, a.y / s); }
PPX_HD double v2_dot(V2 a, V2 b) { return a.x * b.x + a.y * b.y; }
PPX_HD double v2_sqnorm(V2 a) { return a.x * a.x + a.y * a.y; }
PPX_HD double v2_norm(V2 a) { return sqrt(v2_sqnorm(a)); }
PPX_HD V2 v2_normalized(V2 a)
{
	double n = v2_norm(a);
	return n > 0.0 ? v2(a.x / n, a.y / n) : a;
}
// OrthogonalComplement, smoother.cpp:10-13: a - a.dot(b) * b / b.squaredNorm()
PPX_HD V2 v2_orth(V2 a, V2 b) { return a - (v2_dot(a, b) * b) / v2_sqnorm(b); }

#ifdef PPX_HOST_TWIN
#define PPX_ACOSF acosf
#define PPX_COSF cosf
#else
#define PPX_ACOSF acosf
#define PPX_COSF cosf
#endif

// CalculateCurvatureTerm (smoother.cpp:160-213): the three additions it would make, false when kappa <= maxCurvature
PPX_HD bool smooth_curvature(const SmoothDev& P, V2 xim1, V2 xi, V2 xip1, V2& gm1, V2& gc, V2& gp1)
{
	V2 dXi = xi - xim1;
	V2 dXip1 = xip1 - xi;
	float c = (float)v2_dot(v2_normalized(dXi), v2_normalized(dXip1));
	c = c < -1.0f ? -1.0f : (1.0f < c ? 1.0f : c); // std::clamp<float>
	float deltaPhi = PPX_ACOSF(c);
	double nXi = v2_norm(dXi);
	float kappa = (float)((double)deltaPhi / nXi);
	if (kappa <= P.maxCurvature)
		return false;
	float denominator = (float)(nXi * v2_norm(dXip1));
	V2 dCos_dXi = v2_orth(dXip1, dXi) / (double)denominator;
	V2 dCos_dXip1 = v2_orth(dXi, dXip1) / (double)denominator;
	double cosPhi = (double)PPX_COSF(deltaPhi);
	float dPhi_dCos = (float)(-1.0 / sqrt(1.0 - cosPhi * cosPhi)); // std::pow(float, 2) is evaluated in double
	float coef1 = (float)(1.0 / nXi * (double)dPhi_dCos);
	V2 coef2 = ((double)deltaPhi / v2_sqnorm(dXi)) * v2_normalized(dXi);
	V2 dKm1 = (double)coef1 * (-dCos_dXi) + coef2;
	V2 dKc = (double)coef1 * (dCos_dXi - dCos_dXip1) - coef2;
	V2 dKp1 = (double)coef1 * dCos_dXip1;
	float w = P.curvatureWeight * (kappa - P.maxCurvature);
	gm1 = (double)w * dKm1;
	gc = (double)w * dKc;
	gp1 = (double)w * dKp1;
	return true;
}

// OccupancyMap::GridCellToWorldPosition (occupancy_map.h:91-94,146-149): int * float resolution is a float product
PPX_HD V2 cell_to_world(const MapDev& m, int row, int col)
{
	return v2(m.ox + (double)((float)row * m.res), m.oy + (double)((float)col * m.res));
}

// The own terms of loop index i added to g in the reference's order (smoother.cpp:105-143)
PPX_HD V2 smooth_own_terms(const SmoothDev& P, const MapDev& m, V2 g, V2 orig, V2 pm2, V2 pm1, V2 curr, V2 pp1, V2 pp2, float unsafeRadius)
{
	g = g + (double)P.pathWeight * (orig - curr);
	int row = trunc_i32((curr.x - m.ox) / m.resd);
	int col = trunc_i32((curr.y - m.oy) / m.resd);
	if (row >= 0 && row < m.rows && col >= 0 && col < m.cols) { // IsInsideMap(curr)
		const size_t cell = ((size_t)row * m.cols + col) * 2;
		V2 dirObs = curr - cell_to_world(m, PPX_LDG(P.nearObstacle + cell), PPX_LDG(P.nearObstacle + cell + 1));
		float obstDist = (float)v2_norm(dirObs);
		if (obstDist < unsafeRadius)
			g = g + ((double)(P.collisionWeight * (obstDist - unsafeRadius)) * dirObs) / (double)obstDist;
		if (obstDist < P.dMax && P.voronoiWeight > 0.0f) {
			V2 dirEdge = curr - cell_to_world(m, PPX_LDG(P.nearEdge + cell), PPX_LDG(P.nearEdge + cell + 1));
			float voroDist = (float)v2_norm(dirEdge);
			if (voroDist > 0.0f) {
				float aPO = P.alpha + obstDist;
				float oMD = obstDist - P.dMax;
				float oPV = obstDist + voroDist;
				float dMaxSq = P.dMax * P.dMax;
				float pvdv = (P.alpha / aPO) * (oMD * oMD / dMaxSq) * (obstDist / (oPV * oPV));
				float pvdo = (P.alpha / aPO) * (voroDist / oPV) * (oMD / dMaxSq) * (-oMD / aPO - oMD / oPV + 2);
				g = g + (double)P.voronoiWeight * (((double)pvdo * dirObs) / (double)obstDist + ((double)pvdv * dirEdge) / (double)voroDist);
			}
		}
	}
	g = g + (double)P.smoothWeight * (pm2 - 4.0 * pm1 + 6.0 * curr - 4.0 * pp1 + pp2);
	return g;
}

// j is optimised iff 2 <= j <= n - 3 and no cusp index lies in [j - 2, j + 2] (the skip-ahead loop of
// smoother.cpp:55-78 only avoids re-testing windows that contain the same cusp)
PPX_HD bool smooth_is_index(const uint8_t* cusp, int n, int j)
{
	if (j < 2 || j > n - 3)
		return false;
	return !(cusp[j - 2] | cusp[j - 1] | cusp[j] | cusp[j + 1] | cusp[j + 2]);
}

}
