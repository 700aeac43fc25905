// Host half of HybridAStar::Initialize: the constants the reference evaluates once with the host
// libm (steering set, bicycle-model constants, heuristic table geometry).  Pure host code.
#pragma once
#include <algorithm>
#include <cmath>
#include "ppx.h"
#include "ppx_hybrid.cuh"

namespace ppx {

inline int hybrid_derive(const ppx_hybrid_params& p, const ppx_validator_params& vp, const ppx_map_desc& map, HybridDev& P)
{
	P = HybridDev {};
	P.rmin = p.min_turning_radius;
	P.swCost = p.direction_switching_cost;
	P.revMul = p.reverse_cost_multiplier;
	P.fwdMul = p.forward_cost_multiplier;
	P.voroMul = p.voronoi_cost_multiplier;
	P.spatialRes = p.spatial_resolution;
	P.angRes = p.angular_resolution;
	P.revF = (float)p.reverse_cost_multiplier;
	P.fwdF = (float)p.forward_cost_multiplier;
	P.swF = (float)p.direction_switching_cost;
	P.primLength = p.spatial_resolution * 1.5;
	// StatePropagator ctor, hybrid_a_star.cpp:13-29; KinematicBicycleModel(wheelbase, 0.0)
	const double wheelbase = p.wheelbase, rearToCenter = 0.0;
	double deltaMax; // GetSteeringAngleFromTurningRadius, kinematic_bicycle_model.cpp:34-41
	if (p.min_turning_radius < rearToCenter)
		deltaMax = M_PI_2;
	else
		deltaMax = atan(wheelbase / sqrt(std::pow(p.min_turning_radius, 2) - std::pow(rearToCenter, 2)));
	double deltas[kMaxDeltas];
	int nd = 0;
	deltas[nd++] = 0.0;
	for (unsigned int i = 0; i < p.num_generated_motion / 2; i++) {
		if (nd + 2 > kMaxDeltas)
			return PPX_ERR_INVALID;
		double delta = (i + 1) / 2.0 * deltaMax;
		deltas[nd++] = delta;
		deltas[nd++] = -delta;
	}
	P.numDeltas = nd;
	for (int i = 0; i < nd; i++) { // kinematic_bicycle_model.cpp:14-18
		double tanSteering = tan(deltas[i]);
		double beta = atan(rearToCenter * tanSteering / wheelbase);
		double cosBeta = cos(beta);
		P.beta[i] = beta;
		P.cosBeta[i] = cosBeta;
		P.kappa[i] = cosBeta * tanSteering / wheelbase;
	}
	P.voroDiag = (float)((double)map.resolution * std::sqrt(2.0)); // hybrid_a_star.cpp:38
	// NonHolonomicHeuristic::Build geometry, heuristics.cpp:40-49 and ctor :13-14
	const double sizeX = map.ub[0] - map.lb[0], sizeY = map.ub[1] - map.lb[1];
	unsigned int nx = (unsigned int)ceil(sizeX / p.spatial_resolution);
	if (nx % 2 == 0)
		nx++;
	unsigned int ny = (unsigned int)ceil(sizeY / p.spatial_resolution);
	if (ny % 2 == 0)
		ny++;
	P.nhX = (int)nx;
	P.nhY = (int)ny;
	P.nhA = (int)(unsigned int)ceil(2 * M_PI / p.angular_resolution);
	P.nhOffX = floor(nx / 2.0) * p.spatial_resolution;
	P.nhOffY = floor(ny / 2.0) * p.spatial_resolution;
	P.nhMinMul = std::min(p.reverse_cost_multiplier, p.forward_cost_multiplier);
	// ObstaclesHeuristic ctor, heuristics.cpp:97-104 (both members are const float)
	P.obsDiag = (float)(std::sqrt(2) * map.resolution);
	P.obsMul = (float)(std::min(p.reverse_cost_multiplier, p.forward_cost_multiplier) * map.resolution);
	P.minSafeRadius = vp.min_safe_radius;
	P.minStep = vp.min_path_interpolation_distance;
	return PPX_OK;
}

// Dense key table over the discrete poses reachable inside the bounds (+2 cells of slack: the end
// pose of a primitive is never validated).
inline void hybrid_key_table(const ppx_hybrid_params& p, const ppx_map_desc& map, int& x0, int& y0, int& tx, int& ty)
{
	auto cell = [&](double v) { return (int)(v / p.spatial_resolution); };
	x0 = cell(map.lb[0] + map.local_origin[0]) - 2;
	y0 = cell(map.lb[1] + map.local_origin[1]) - 2;
	tx = cell(map.ub[0] + map.local_origin[0]) + 2 - x0 + 1;
	ty = cell(map.ub[1] + map.local_origin[1]) + 2 - y0 + 1;
}

}
