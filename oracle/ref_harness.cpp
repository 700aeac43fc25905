// TEST INFRASTRUCTURE ONLY -- never linked into, imported by, or executed from the product path.
//
// C-callable harness over the UNMODIFIED reference sources (compiled where they lie under
// /root/reference by oracle/Makefile into oracle/_ref/libpp_ref.so).  It exists so that tests,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference arm can run the
// reference's own C++ path on the same inputs as the CUDA path.
//
// Harness-side definitions that the reference leaves undefined (traps):
//  * H4  the reference seeds one process-global mt19937_64 from std::random_device
//        (utils/random.h:14-16); the harness re-seeds it explicitly before each search.
//  * H3  NonHolonomicHeuristic::GetHeuristicValue reads m_values[i][j][k] with k<0
//        (algo/heuristics.cpp:84-94, out of bounds).  The harness re-points every m_values[i][j]
//        into a padded copy where index k<0 aliases k+numAngular, which makes the read defined
//        without touching reference source.  The CUDA path implements the same wrap.
//  * H15 StatePropagator::GetVoronoiCost indexes GVD::m_pathCostMap unchecked (hybrid_a_star.cpp:102-104,
//        assert compiled out) and the end of a primitive may overshoot the map by one validator step
//        (trap H6): rows beyond the grid read past the heap block (ASan: heap-buffer-overflow; the value
//        depends on the heap layout and decides expansion order).  The harness re-points the grid's
//        storage into the middle of a zero-filled block (kPadRows rows on either side), so those reads
//        return 0 without touching reference source.  The CUDA path reads 0 there as well; reads that
//        stay inside the allocation (a column overflow aliasing the next row) are reproduced as is.
//  * H1/H2 are properties of how the reference compiles (abs overload, int WrapTheta); the
//        harness asserts both at load time (ref_selfcheck).
//
// Private members are reached with the usual test-only access macro trick; all standard headers
// are included first so the macros only affect reference headers.
#include <stdlib.h>
#include <math.h>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <cstdio>
#include <memory>
#include <vector>
#include <array>
#include <tuple>
#include <set>
#include <map>
#include <queue>
#include <random>
#include <string>
#include <chrono>
#include <limits>
#include <numeric>
#include <fstream>
#include <sstream>
#include <iostream>
#include <algorithm>
#include <functional>
#include <filesystem>
#include <unordered_map>
#include <unordered_set>
#include <mutex>
#include <thread>
#include <signal.h>
#include <eigen3/Eigen/Dense>
#include <flann/flann.hpp>

#define private public
#define protected public
#include "core/base.h"
#include "algo/hybrid_a_star.h"
#include "algo/heuristics.h"
#include "algo/a_star_n2.h"
#include "algo/rrt.h"
#include "algo/rrt_star.h"
#include "geometry/2dplane.h"
#include "geometry/reeds_shepp.h"
#include "models/kinematic_bicycle_model.h"
#include "paths/path_constant_steer.h"
#include "paths/path_reeds_shepp.h"
#include "paths/path_se2.h"
#include "paths/path_r2.h"
#include "paths/path_composite.h"
#include "state_space/state_space_se2.h"
#include "state_space/state_space_r2.h"
#include "state_validator/gvd.h"
#include "state_validator/obstacle.h"
#include "state_validator/obstacle_list_occupancy_map.h"
#include "state_validator/occupancy_map.h"
#include "state_validator/state_validator_occupancy_map.h"
#include "state_validator/state_validator_free.h"
#include "utils/tree.h"
#undef private
#undef protected

using namespace Planner;

namespace {

// Counts IsStateValid calls (= samples visited by IsPathValid, which calls the virtual).
class CountingValidator : public StateValidatorOccupancyMap {
public:
	using StateValidatorOccupancyMap::StateValidatorOccupancyMap;
	bool IsStateValid(const Pose2d& state) override
	{
		samples++;
		return StateValidatorOccupancyMap::IsStateValid(state);
	}
	long samples = 0;
};

struct RefWorld {
	Ref<StateSpaceSE2> space;
	Ref<ObstacleListOccupancyMap> map;
	Ref<CountingValidator> validator;
	Ref<GVD> gvd;
	std::vector<Ref<Obstacle>> obstacles;
};

void SeedRandom(uint64_t seed)
{
	Random<double>::Init();
	*Random<double>::s_engine = std::mt19937_64(seed);
	Random<double>::s_uniformDistribution.reset();
}

ReedsShepp::PathSegment MakeSegment(const int* steer, const int* dir, const double* len)
{
	ReedsShepp::PathSegment seg;
	for (int i = 0; i < 5; i++) {
		if (dir[i] == (int)Direction::NoMotion)
			break;
		seg.AddMotion((Steer)steer[i], (Direction)dir[i], len[i]);
	}
	return seg;
}

void DumpSegment(const ReedsShepp::PathSegment& seg, int* steer, int* dir, double* len)
{
	for (int i = 0; i < 5; i++) {
		const auto& m = seg.GetMotion(i);
		steer[i] = (int)m.steer;
		dir[i] = (int)m.direction;
		len[i] = m.length;
	}
}

}

extern "C" {

// ---------------------------------------------------------------------------------------------
// Self check of the compile-dependent traps H1 / H2.  Returns 0 when the build behaves like an
// upstream build (curved primitives, int-arithmetic heading wrap), else a bit mask.
int ref_selfcheck(int* wrapped_out /* 75 ints: k=-37..37 -> wrapped k */)
{
	int bad = 0;
	KinematicBicycleModel model(2.6, 0.0);
	Pose2d to = model.ConstantSteer(Pose2d(0, 0, 0), 0.3, 1.5, Direction::Forward);
	if (to.theta == 0.0)
		bad |= 1; // H1: abs(int) picked
	HybridAStar::SearchParameters p;
	HybridAStar::StatePropagator prop(p);
	for (int k = -37; k <= 37; k++) {
		// pose whose heading falls in bin k
		double th = (k + (k >= 0 ? 0.5 : -0.5)) * p.angularResolution;
		Pose2d pose;
		pose.position = Point2d(0.5, 0.5);
		pose.theta = th;
		Pose2i d = prop.DiscretizePose(pose);
		if (wrapped_out)
			wrapped_out[k + 37] = d.theta;
	}
	return bad;
}

// ---------------------------------------------------------------------------------------------
// Reeds-Shepp
int ref_rs_optimal(const double* start, const double* goal, double rmin, float rev, float fwd, float sw,
	int* steer, int* dir, double* len, double* total_len, float* cost)
{
	ReedsShepp::PathWords word = ReedsShepp::PathWords::NoPath;
	auto seg = ReedsShepp::Solver::GetOptimalPath(Pose2d(start[0], start[1], start[2]), Pose2d(goal[0], goal[1], goal[2]),
		rmin, rev, fwd, sw, &word);
	DumpSegment(seg, steer, dir, len);
	if (total_len)
		*total_len = seg.GetLength(rmin);
	if (cost)
		*cost = seg.ComputeCost(rmin, rev, fwd, sw);
	return (int)word;
}

// start/goal are passed verbatim (no Pose2d ctor wrap) when raw != 0
int ref_rs_shortest(const double* start, const double* goal, double rmin,
	int* steer, int* dir, double* len, double* total_len, double* final_pose)
{
	ReedsShepp::PathWords word = ReedsShepp::PathWords::NoPath;
	Pose2d s(start[0], start[1], start[2]), g(goal[0], goal[1], goal[2]);
	auto seg = ReedsShepp::Solver::GetShortestPath(s, g, rmin, &word);
	DumpSegment(seg, steer, dir, len);
	if (total_len)
		*total_len = seg.GetLength(rmin);
	if (final_pose) {
		PathReedsShepp path(s, seg, rmin);
		final_pose[0] = path.GetFinalState().x();
		final_pose[1] = path.GetFinalState().y();
		final_pose[2] = path.GetFinalState().theta;
	}
	return (int)word;
}

void ref_rs_optimal_batch(long n, const double* starts, const double* goals, double rmin, float rev, float fwd, float sw,
	int* words, int* steer, int* dir, double* len, double* total_len, float* cost)
{
	for (long i = 0; i < n; i++)
		words[i] = ref_rs_optimal(starts + 3 * i, goals + 3 * i, rmin, rev, fwd, sw,
			steer + 5 * i, dir + 5 * i, len + 5 * i, total_len ? total_len + i : nullptr, cost ? cost + i : nullptr);
}

// Interpolate an RS path (optionally truncated first; truncate_ratio<0 => no truncation)
void ref_rs_interpolate(const double* start, const int* steer, const int* dir, const double* len, double rmin,
	double truncate_ratio, int nratios, const double* ratios, double* poses, double* out_length, double* final_pose)
{
	Pose2d s;
	s.position = Point2d(start[0], start[1]);
	s.theta = start[2];
	PathReedsShepp path(s, MakeSegment(steer, dir, len), rmin);
	if (truncate_ratio >= 0)
		path.Truncate(truncate_ratio);
	for (int i = 0; i < nratios; i++) {
		Pose2d p = path.Interpolate(ratios[i]);
		poses[3 * i + 0] = p.x();
		poses[3 * i + 1] = p.y();
		poses[3 * i + 2] = p.theta;
	}
	if (out_length)
		*out_length = path.GetLength();
	if (final_pose) {
		final_pose[0] = path.GetFinalState().x();
		final_pose[1] = path.GetFinalState().y();
		final_pose[2] = path.GetFinalState().theta;
	}
}

void ref_csteer_interpolate(const double* start, double steering, double length, int direction, double wheelbase,
	int nratios, const double* ratios, double* poses)
{
	auto model = makeRef<KinematicBicycleModel>(wheelbase, 0.0);
	Pose2d s;
	s.position = Point2d(start[0], start[1]);
	s.theta = start[2];
	PathConstantSteer path(model, s, steering, length, (Direction)direction);
	for (int i = 0; i < nratios; i++) {
		Pose2d p = path.Interpolate(ratios[i]);
		poses[3 * i + 0] = p.x();
		poses[3 * i + 1] = p.y();
		poses[3 * i + 2] = p.theta;
	}
}

// ---------------------------------------------------------------------------------------------
// World: state space + obstacle-list occupancy map + validator
void* ref_world_create(const double* lb, const double* ub, float resolution)
{
	auto* w = new RefWorld();
	std::array<Pose2d, 2> bounds = { Pose2d(lb[0], lb[1], lb[2]), Pose2d(ub[0], ub[1], ub[2]) };
	w->space = makeRef<StateSpaceSE2>(bounds);
	w->map = makeRef<ObstacleListOccupancyMap>(resolution);
	w->validator = makeRef<CountingValidator>(w->space, w->map);
	return w;
}

void ref_world_destroy(void* h) { delete (RefWorld*)h; }

void ref_world_set_position(void* h, double x, double y)
{
	// OccupancyMap::SetPosition only stores m_localOrigin; the grid origin is fixed at
	// InitializeSize time (occupancy_map.cpp:8-9), so re-run it like a user would have to.
	auto* w = (RefWorld*)h;
	w->map->SetPosition(Point2d(x, y));
}

int ref_world_add_rectangle(void* h, double dx, double dy, double x, double y, double theta)
{
	auto* w = (RefWorld*)h;
	auto obstacle = makeRef<Obstacle>();
	obstacle->SetShape(makeRef<RectangleShape>(dx, dy));
	obstacle->SetPose(Pose2d(x, y, theta));
	w->obstacles.push_back(obstacle);
	return w->map->AddObstacle(obstacle) ? 1 : 0;
}

// Mark single cells occupied (id>=0) the way AddObstacle does for each boundary cell
// (obstacle_list_occupancy_map.cpp:39-42).
void ref_world_set_cells(void* h, long n, const int* rows, const int* cols, int id)
{
	auto* w = (RefWorld*)h;
	for (long i = 0; i < n; i++) {
		(*w->map->m_occupancyMatrix)[rows[i]][cols[i]] = id;
		w->map->m_obstacleDistanceMap->SetObstacle(GridCellPosition(rows[i], cols[i]));
	}
}

void ref_world_update(void* h) { ((RefWorld*)h)->map->Update(); }

void ref_world_info(void* h, int* rows, int* cols, double* origin_xy, float* resolution)
{
	auto* w = (RefWorld*)h;
	*rows = w->map->Rows();
	*cols = w->map->Columns();
	origin_xy[0] = w->map->m_worldGridOrigin.x();
	origin_xy[1] = w->map->m_worldGridOrigin.y();
	origin_xy[2] = w->map->m_localOrigin.x();
	origin_xy[3] = w->map->m_localOrigin.y();
	*resolution = w->map->resolution;
}

void ref_world_copy_grids(void* h, int* occupancy, int* sqdist)
{
	auto* w = (RefWorld*)h;
	long n = (long)w->map->Rows() * w->map->Columns();
	if (occupancy)
		memcpy(occupancy, w->map->m_occupancyMatrix->m_data, n * sizeof(int));
	if (sqdist)
		memcpy(sqdist, w->map->m_obstacleDistanceMap->m_distance.m_data, n * sizeof(int));
}

void ref_world_set_validator_params(void* h, float minSafeRadius, float minPathInterpolationDistance)
{
	auto* w = (RefWorld*)h;
	w->validator->minSafeRadius = minSafeRadius;
	w->validator->minPathInterpolationDistance = minPathInterpolationDistance;
}

// GVD::Update (obstacle brushfire + Voronoi brushfire + potential), copies out the float
// potential (GVD::PathCostMap) and the squared Voronoi-edge distance.
void ref_world_gvd_update(void* h, float* pathcost, int* voro_sqdist)
{
	auto* w = (RefWorld*)h;
	if (!w->gvd)
		w->gvd = makeRef<GVD>(w->map);
	w->gvd->Update();
	long n = (long)w->map->Rows() * w->map->Columns();
	if (pathcost)
		memcpy(pathcost, w->gvd->m_pathCostMap.m_data, n * sizeof(float));
	if (voro_sqdist)
		memcpy(voro_sqdist, w->gvd->m_voronoiMap->m_distance.m_data, n * sizeof(int));
}

// bench.py, C4: the reference searches the SAME potential the device built.  Overwrites the two distance grids of the
// world's GVD (GVD::PathCostMap::Update recomputes the potential from them at every GVD::Update, gvd.cpp:266-283,
// hybrid_a_star.cpp:250) with the caller's; returns the number of cells that differed (obstacle, Voronoi).
void ref_world_set_gvd_distances(void* h, const int* sqdist, const int* voro_sqdist, long* diff)
{
	auto* w = (RefWorld*)h;
	if (!w->gvd)
		w->gvd = makeRef<GVD>(w->map);
	long n = (long)w->map->Rows() * w->map->Columns();
	int* a = w->map->m_obstacleDistanceMap->m_distance.m_data;
	int* b = w->gvd->m_voronoiMap->m_distance.m_data;
	long da = 0, db = 0;
	for (long i = 0; i < n; i++) {
		da += a[i] != sqdist[i];
		db += b[i] != voro_sqdist[i];
		a[i] = sqdist[i];
		b[i] = voro_sqdist[i];
	}
	if (diff) {
		diff[0] = da;
		diff[1] = db;
	}
}

// The nearest-obstacle / nearest-Voronoi-edge cells of the world's own GVD (after ref_world_gvd_update), (row, col) pairs.
void ref_world_gvd_cells(void* h, int32_t* nearestObstacle, int32_t* nearestEdge)
{
	auto* w = (RefWorld*)h;
	if (!w->gvd)
		return;
	const int rows = w->map->Rows(), cols = w->map->Columns();
	for (int r = 0; r < rows; r++)
		for (int c = 0; c < cols; c++) {
			const auto& o = w->gvd->GetNearestObstacleCell(r, c);
			const auto& e = w->gvd->GetNearestVoronoiEdgeCell(r, c);
			size_t i = ((size_t)r * cols + c) * 2;
			nearestObstacle[i] = o.row; nearestObstacle[i + 1] = o.col;
			nearestEdge[i] = e.row; nearestEdge[i + 1] = e.col;
		}
}

int ref_is_state_valid(void* h, const double* pose)
{
	auto* w = (RefWorld*)h;
	Pose2d p;
	p.position = Point2d(pose[0], pose[1]);
	p.theta = pose[2];
	return w->validator->IsStateValid(p) ? 1 : 0;
}

void ref_is_state_valid_batch(void* h, long n, const double* poses, uint8_t* valid)
{
	for (long i = 0; i < n; i++)
		valid[i] = (uint8_t)ref_is_state_valid(h, poses + 3 * i);
}

// Solve (GetOptimalPath) + build PathReedsShepp + IsPathValid for n start/goal pairs.
// Any output pointer may be null.  Returns wall seconds of the loop.
double ref_check_rs_edges(void* h, long n, const double* starts, const double* goals, double rmin,
	float rev, float fwd, float sw,
	uint8_t* valid, float* last, int* samples, int* words, int* steer, int* dir, double* len,
	double* total_len, float* cost)
{
	auto* w = (RefWorld*)h;
	auto t0 = std::chrono::steady_clock::now();
	for (long i = 0; i < n; i++) {
		Pose2d s(starts[3 * i], starts[3 * i + 1], starts[3 * i + 2]);
		Pose2d g(goals[3 * i], goals[3 * i + 1], goals[3 * i + 2]);
		ReedsShepp::PathWords word = ReedsShepp::PathWords::NoPath;
		auto seg = ReedsShepp::Solver::GetOptimalPath(s, g, rmin, rev, fwd, sw, &word);
		PathReedsShepp path(s, seg, rmin);
		float l = 0.0f;
		long before = w->validator->samples;
		bool ok = w->validator->IsPathValid(path, &l);
		if (valid) valid[i] = ok;
		if (last) last[i] = l;
		if (samples) samples[i] = (int)(w->validator->samples - before);
		if (words) words[i] = (int)word;
		if (steer) DumpSegment(seg, steer + 5 * i, dir + 5 * i, len + 5 * i);
		if (total_len) total_len[i] = seg.GetLength(rmin);
		if (cost) cost[i] = seg.ComputeCost(rmin, rev, fwd, sw);
	}
	return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// Check-only: pre-solved segments.  starts are taken verbatim (no wrap).
double ref_check_rs_segments(void* h, long n, const double* starts, const int* steer, const int* dir, const double* len,
	double rmin, uint8_t* valid, float* last, int* samples)
{
	auto* w = (RefWorld*)h;
	std::vector<PathReedsShepp> paths;
	paths.reserve(n);
	for (long i = 0; i < n; i++) {
		Pose2d s;
		s.position = Point2d(starts[3 * i], starts[3 * i + 1]);
		s.theta = starts[3 * i + 2];
		paths.emplace_back(s, MakeSegment(steer + 5 * i, dir + 5 * i, len + 5 * i), rmin);
	}
	auto t0 = std::chrono::steady_clock::now();
	for (long i = 0; i < n; i++) {
		float l = 0.0f;
		long before = w->validator->samples;
		bool ok = w->validator->IsPathValid(paths[i], &l);
		if (valid) valid[i] = ok;
		if (last) last[i] = l;
		if (samples) samples[i] = (int)(w->validator->samples - before);
	}
	return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// Rectangular footprint variant (BASELINE config 3; NOT a reference behaviour -- "port", parity unpinned by definition): the
// definition of include/ppx.h (ppx_footprint_params) restated over the reference's own pieces -- PathReedsShepp::Interpolate for
// the poses, StateValidatorOccupancyMap::IsStateValid with minSafeRadius = rho for every covering disc, the map's distance grid.
// fp = { length, width, rear_to_center, n_discs, min_step }.
double ref_check_rs_segments_rect(void* h, long n, const double* starts, const int* steer, const int* dir, const double* len,
	double rmin, const double* fp, uint8_t* valid, float* last, int* samples)
{
	auto* w = (RefWorld*)h;
	const int nd = (int)fp[3];
	const double seg = fp[0] / nd, first = fp[2] - fp[0] / 2.0 + seg / 2.0;
	const float rho = (float)std::sqrt(seg * seg / 4.0 + fp[1] * fp[1] / 4.0);
	double omax = 0.0;
	for (int i = 0; i < nd; i++)
		omax = std::fmax(omax, std::fabs(first + i * seg));
	const float gamma = (float)std::sqrt(1.0 + (omax / rmin) * (omax / rmin));
	const float minStep = (float)fp[4];
	const auto& bounds = w->space->bounds;
	const float savedRadius = w->validator->minSafeRadius;
	w->validator->minSafeRadius = rho;
	auto poseState = [&](const Pose2d& p, float& slack, float& border) {
		const double s = std::sin(p.theta), c = std::cos(p.theta);
		bool ok = true;
		slack = border = std::numeric_limits<float>::infinity();
		for (int i = 0; i < nd; i++) {
			const double o = first + i * seg;
			Pose2d d;
			d.position = Point2d(p.x() + o * c, p.y() + o * s);
			d.theta = p.theta;
			float dist = 0.0f;
			if (!w->validator->IsStateValid(d))
				ok = false;
			auto cell = w->map->WorldPositionToGridCell(d.position);
			if (w->map->IsInsideMap(cell)) {
				bool inb = !(d.x() - w->map->GetPosition().x() < bounds[0].x() || d.x() - w->map->GetPosition().x() > bounds[1].x()
					|| d.y() - w->map->GetPosition().y() < bounds[0].y() || d.y() - w->map->GetPosition().y() > bounds[1].y());
				double th = d.theta;
				while (th > M_PI) th -= 2 * M_PI;
				while (th < -M_PI) th += 2 * M_PI;
				inb = inb && !(th < bounds[0].theta || th > bounds[1].theta);
				if (inb)
					dist = w->map->GetObstacleDistanceMap()->GetDistanceToNearestObstacle(cell);
			}
			slack = std::min(slack, dist - rho);
			double b = std::min({ d.x() - bounds[0].x(), bounds[1].x() - d.x(), d.y() - bounds[0].y(), bounds[1].y() - d.y() });
			border = std::min(border, (float)b);
		}
		return ok;
	};
	auto t0 = std::chrono::steady_clock::now();
	for (long i = 0; i < n; i++) {
		Pose2d s;
		s.position = Point2d(starts[3 * i], starts[3 * i + 1]);
		s.theta = starts[3 * i + 2];
		PathReedsShepp path(s, MakeSegment(steer + 5 * i, dir + 5 * i, len + 5 * i), rmin);
		const double pathLength = path.GetLength();
		float slack, border, l = 1.0f;
		int ns = 0;
		bool ok = true;
		if (pathLength == 0.0) {
			ns = 1;
			ok = poseState(path.GetInitialState(), slack, border);
		} else {
			double lastValidLength = 0.0, length = 0.0;
			while (length < pathLength) {
				Pose2d state = path.Interpolate(length / pathLength);
				ns++;
				if (!poseState(state, slack, border)) {
					l = lastValidLength / pathLength;
					ok = false;
					break;
				}
				lastValidLength = length;
				float delta = std::min(slack, border) / gamma;
				length += std::max(delta, minStep);
			}
		}
		valid[i] = ok;
		last[i] = l;
		if (samples) samples[i] = ns;
	}
	w->validator->minSafeRadius = savedRadius;
	return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// Constant-steer primitives (A2): validity, last ratio, final pose (after truncation when invalid,
// exactly as GetConstantSteerChild does, hybrid_a_star.cpp:117-123).
void ref_check_csteer(void* h, long n, const double* starts, const double* steering, const double* length,
	const int* direction, double wheelbase, uint8_t* valid, float* last, double* final_pose, double* final_len)
{
	auto* w = (RefWorld*)h;
	auto model = makeRef<KinematicBicycleModel>(wheelbase, 0.0);
	for (long i = 0; i < n; i++) {
		Pose2d s;
		s.position = Point2d(starts[3 * i], starts[3 * i + 1]);
		s.theta = starts[3 * i + 2];
		PathConstantSteer path(model, s, steering[i], length[i], (Direction)direction[i]);
		float l = 0.0f;
		bool ok = w->validator->IsPathValid(path, &l);
		if (!ok)
			path.Truncate(l);
		if (valid) valid[i] = ok;
		if (last) last[i] = l;
		if (final_pose) {
			final_pose[3 * i + 0] = path.GetFinalState().x();
			final_pose[3 * i + 1] = path.GetFinalState().y();
			final_pose[3 * i + 2] = path.GetFinalState().theta;
		}
		if (final_len) final_len[i] = path.GetLength();
	}
}

// Straight SE(2) segments (paths/path_se2.cpp), used by RRT with PathConnectionSE2.
void ref_check_line_edges(void* h, long n, const double* from, const double* to, uint8_t* valid, float* last)
{
	auto* w = (RefWorld*)h;
	for (long i = 0; i < n; i++) {
		Pose2d a, b;
		a.position = Point2d(from[3 * i], from[3 * i + 1]);
		a.theta = from[3 * i + 2];
		b.position = Point2d(to[3 * i], to[3 * i + 1]);
		b.theta = to[3 * i + 2];
		PathSE2 path(a, b);
		float l = 0.0f;
		bool ok = w->validator->IsPathValid(path, &l);
		if (valid) valid[i] = ok;
		if (last) last[i] = l;
	}
}

// ---------------------------------------------------------------------------------------------
// Holonomic-with-obstacles wavefront (A8).  Returns wall seconds of Update().
double ref_wavefront(void* h, const double* goal, double rev, double fwd, float* cost, uint8_t* explored)
{
	auto* w = (RefWorld*)h;
	ObstaclesHeuristic heur(w->map, rev, fwd);
	Pose2d g(goal[0], goal[1], goal[2]);
	auto t0 = std::chrono::steady_clock::now();
	heur.Update(g);
	double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	long n = (long)w->map->Rows() * w->map->Columns();
	if (cost)
		memcpy(cost, heur.m_cost.m_data, n * sizeof(float));
	if (explored)
		for (long i = 0; i < n; i++)
			explored[i] = heur.m_explored.m_data[i] ? 1 : 0;
	return dt;
}

// Heuristic lookups for n poses after a wavefront from `goal`
void ref_obstacle_heuristic_values(void* h, const double* goal, double rev, double fwd, long n, const double* poses, double* values)
{
	auto* w = (RefWorld*)h;
	ObstaclesHeuristic heur(w->map, rev, fwd);
	Pose2d g(goal[0], goal[1], goal[2]);
	heur.Update(g);
	heur.SetGoal(g);
	for (long i = 0; i < n; i++) {
		Pose2d p;
		p.position = Point2d(poses[3 * i], poses[3 * i + 1]);
		p.theta = poses[3 * i + 2];
		values[i] = heur.GetHeuristicValue(p);
	}
}

// ---------------------------------------------------------------------------------------------
// Non-holonomic-without-obstacles table (A9).  values may be null to query the dimensions.
double ref_nonholo_build(const double* lb, const double* ub, double spatialRes, double angRes, double rmin,
	double rev, double fwd, double sw, int* dims /*nx,ny,na*/, double* offsets /*ox,oy*/, double* values)
{
	std::array<Pose2d, 2> bounds = { Pose2d(lb[0], lb[1], lb[2]), Pose2d(ub[0], ub[1], ub[2]) };
	auto t0 = std::chrono::steady_clock::now();
	auto heur = NonHolonomicHeuristic::Build(bounds, spatialRes, angRes, rmin, rev, fwd, sw);
	double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	dims[0] = heur->numSpatialX;
	dims[1] = heur->numSpatialY;
	dims[2] = heur->numAngular;
	if (offsets) {
		offsets[0] = heur->offsetX;
		offsets[1] = heur->offsetY;
	}
	if (values)
		for (unsigned i = 0; i < heur->numSpatialX; i++)
			for (unsigned j = 0; j < heur->numSpatialY; j++)
				for (unsigned k = 0; k < heur->numAngular; k++)
					values[((size_t)i * heur->numSpatialY + j) * heur->numAngular + k] = heur->m_values[i][j][k];
	return dt;
}

// ---------------------------------------------------------------------------------------------
// Hybrid A*
namespace {
struct ExpansionRecord {
	int ix, iy, ik, direction;
	double x, y, theta;
};

class RecordingPropagator : public HybridAStar::StatePropagator {
public:
	RecordingPropagator(const HybridAStar::SearchParameters& p) :
		HybridAStar::StatePropagator(p) { }
	std::vector<std::tuple<HybridAStar::State, HybridAStar::Action, double>> GetNeighborStates(const HybridAStar::State& state) override
	{
		if (record)
			expansions.push_back({ state.discretePose.x(), state.discretePose.y(), state.discretePose.theta, (int)state.direction,
				state.continuousPose.x(), state.continuousPose.y(), state.continuousPose.theta });
		auto out = HybridAStar::StatePropagator::GetNeighborStates(state);
		numChildren += out.size();
		return out;
	}
	bool record = true;
	std::vector<ExpansionRecord> expansions;
	size_t numChildren = 0;
};

struct RefHybrid {
	RefWorld* world;
	std::unique_ptr<HybridAStar> planner;
	Ref<RecordingPropagator> propagator;
	std::vector<double*> paddedBlocks; // H3 padding
	std::vector<double*> originalBlocks;
	float* paddedPathCost = nullptr; // H15 padding
	float* originalPathCost = nullptr;
};
constexpr long kPadRows = 64;
}

void* ref_hybrid_create(void* hw, double minTurningRadius, double directionSwitchingCost, double reverseCostMultiplier,
	double forwardCostMultiplier, double voronoiCostMultiplier, unsigned numGeneratedMotion, double spatialResolution,
	double angularResolution)
{
	auto* h = new RefHybrid();
	h->world = (RefWorld*)hw;
	HybridAStar::SearchParameters p(minTurningRadius, directionSwitchingCost, reverseCostMultiplier, forwardCostMultiplier,
		voronoiCostMultiplier, numGeneratedMotion, spatialResolution, angularResolution);
	h->planner = std::make_unique<HybridAStar>(p);
	// validator handed to the planner is the counting subclass (same virtual behaviour)
	Ref<StateValidatorOccupancyMap> validator = h->world->validator;
	if (!h->planner->Initialize(validator)) {
		delete h;
		return nullptr;
	}
	h->world->gvd = h->planner->m_gvd;
	// Swap in the recording propagator (same parameters / same base behaviour).
	h->propagator = makeRef<RecordingPropagator>(p);
	h->propagator->Initialize(validator, h->planner->m_graphSearch.m_heuristic, h->planner->m_gvd);
	h->planner->m_propagator = h->propagator;
	h->planner->m_graphSearch.Initialize(h->propagator, h->planner->m_graphSearch.m_heuristic);
	// H3: make negative-k reads defined (alias k+numAngular)
	auto& nh = *h->planner->m_nonHoloHeuristic;
	const unsigned na = nh.numAngular;
	for (unsigned i = 0; i < nh.numSpatialX; i++)
		for (unsigned j = 0; j < nh.numSpatialY; j++) {
			double* orig = nh.m_values[i][j];
			double* padded = new double[2 * na];
			for (unsigned k = 0; k < na; k++) {
				padded[k] = orig[k];
				padded[na + k] = orig[k];
			}
			h->originalBlocks.push_back(orig);
			h->paddedBlocks.push_back(padded);
			nh.m_values[i][j] = padded + na;
		}
	// H15: zero padding around GVD::m_pathCostMap
	{
		auto& grid = h->planner->m_gvd->m_pathCostMap;
		const long n = (long)grid.rows * grid.columns, pad = kPadRows * grid.columns;
		h->paddedPathCost = new float[n + 2 * pad]();
		memcpy(h->paddedPathCost + pad, grid.m_data, n * sizeof(float));
		h->originalPathCost = grid.m_data;
		grid.m_data = h->paddedPathCost + pad;
	}
	return h;
}

void ref_hybrid_destroy(void* hh)
{
	auto* h = (RefHybrid*)hh;
	if (h->paddedPathCost) {
		h->planner->m_gvd->m_pathCostMap.m_data = h->originalPathCost;
		delete[] h->paddedPathCost;
	}
	auto& nh = *h->planner->m_nonHoloHeuristic;
	size_t idx = 0;
	for (unsigned i = 0; i < nh.numSpatialX; i++)
		for (unsigned j = 0; j < nh.numSpatialY; j++)
			nh.m_values[i][j] = h->originalBlocks[idx++];
	for (double* p : h->paddedBlocks)
		delete[] p;
	delete h;
}

void ref_hybrid_table(void* hh, int* dims, double* offsets, double* values)
{
	auto* h = (RefHybrid*)hh;
	auto& nh = *h->planner->m_nonHoloHeuristic;
	dims[0] = nh.numSpatialX;
	dims[1] = nh.numSpatialY;
	dims[2] = nh.numAngular;
	if (offsets) {
		offsets[0] = nh.offsetX;
		offsets[1] = nh.offsetY;
	}
	if (values)
		for (unsigned i = 0; i < nh.numSpatialX; i++)
			for (unsigned j = 0; j < nh.numSpatialY; j++)
				for (unsigned k = 0; k < nh.numAngular; k++)
					values[((size_t)i * nh.numSpatialY + j) * nh.numAngular + k] = nh.m_values[i][j][k];
}

// Graph search only (wavefront + GVD update + A*), no smoothing: the part of
// HybridAStar::SearchPath at hybrid_a_star.cpp:248-257.
// times: [0] wavefront s, [1] gvd s, [2] graph search s
// Returns status (0 success, -1 failure).
int ref_hybrid_search(void* hh, const double* start, const double* goal, uint64_t seed, int do_wavefront,
	double* cost, long* num_expanded, long* num_children, long* num_samples, double* times)
{
	auto* h = (RefHybrid*)hh;
	auto& pl = *h->planner;
	Pose2d s(start[0], start[1], start[2]), g(goal[0], goal[1], goal[2]);
	pl.SetInitState(s);
	pl.SetGoalState(g);
	SeedRandom(seed);
	h->propagator->expansions.clear();
	h->propagator->numChildren = 0;
	long samplesBefore = h->world->validator->samples;
	auto t0 = std::chrono::steady_clock::now();
	if (do_wavefront)
		pl.m_obstacleHeuristic->Update(pl.m_goal);
	auto t1 = std::chrono::steady_clock::now();
	pl.m_gvd->Update();
	auto t2 = std::chrono::steady_clock::now();
	pl.m_graphSearch.SetInitState(pl.m_propagator->CreateStateFromPose(pl.m_init));
	pl.m_graphSearch.SetGoalState(pl.m_propagator->SetGoalState(pl.m_goal));
	Status st = pl.m_graphSearch.SearchPath();
	auto t3 = std::chrono::steady_clock::now();
	pl.m_stats.graphSearchStatus = st;
	if (cost) *cost = pl.m_graphSearch.GetOptimalCost();
	if (num_expanded) *num_expanded = (long)h->propagator->expansions.size();
	if (num_children) *num_children = (long)h->propagator->numChildren;
	if (num_samples) *num_samples = h->world->validator->samples - samplesBefore;
	if (times) {
		times[0] = std::chrono::duration<double>(t1 - t0).count();
		times[1] = std::chrono::duration<double>(t2 - t1).count();
		times[2] = std::chrono::duration<double>(t3 - t2).count();
	}
	return (int)st;
}

// Expansion-ordered node records of the last search: 4 ints + 3 doubles each
long ref_hybrid_expansions(void* hh, long cap, int* keys /*4*cap: ix,iy,ik,dir*/, double* poses /*3*cap*/)
{
	auto* h = (RefHybrid*)hh;
	long n = (long)h->propagator->expansions.size();
	for (long i = 0; i < n && i < cap; i++) {
		const auto& e = h->propagator->expansions[i];
		if (keys) {
			keys[4 * i + 0] = e.ix;
			keys[4 * i + 1] = e.iy;
			keys[4 * i + 2] = e.ik;
			keys[4 * i + 3] = e.direction;
		}
		if (poses) {
			poses[3 * i + 0] = e.x;
			poses[3 * i + 1] = e.y;
			poses[3 * i + 2] = e.theta;
		}
	}
	return n;
}

// Solution of the last search as the node chain root..goal: continuous poses + per-action kind
// (0 constant-steer, 1 Reeds-Shepp) and length.  Returns number of states (0 when no solution).
long ref_hybrid_solution(void* hh, long cap, double* poses, int* kinds, double* lengths)
{
	auto* h = (RefHybrid*)hh;
	auto states = h->planner->m_graphSearch.GetPath();
	auto actions = h->planner->m_graphSearch.GetActions();
	long n = (long)states.size();
	for (long i = 0; i < n && i < cap; i++) {
		poses[3 * i + 0] = states[i].continuousPose.x();
		poses[3 * i + 1] = states[i].continuousPose.y();
		poses[3 * i + 2] = states[i].continuousPose.theta;
		if (i > 0) {
			auto& path = actions[i - 1].path;
			if (kinds) kinds[i] = dynamic_cast<PathReedsShepp*>(path.get()) ? 1 : 0;
			if (lengths) lengths[i] = path->GetLength();
		} else {
			if (kinds) kinds[0] = -1;
			if (lengths) lengths[0] = 0.0;
		}
	}
	return n;
}

// The "Process graph search path" block of HybridAStar::SearchPath (hybrid_a_star.cpp:260-291) applied to the
// last search: the composite path of the solution resampled every `pathInterpolation` with the cusp points
// snapped in, BEFORE smoothing.  The loop is restated here (it is local to SearchPath, which goes on to smooth);
// every object it uses -- GetCompositePath, GetCuspPointRatios, Interpolate -- is the reference's own.
// Returns the number of samples (0 without solution); cusp[i] = 1 where i is in cuspIndices.
long ref_hybrid_resampled(void* hh, float pathInterpolation, long cap, double* poses, uint8_t* cusp, double* path_length)
{
	auto* h = (RefHybrid*)hh;
	auto& pl = *h->planner;
	if (pl.m_stats.graphSearchStatus < 0)
		return 0;
	auto path = pl.m_graphSearch.GetCompositePath();
	const double pathLength = path->GetLength();
	if (path_length) *path_length = pathLength;
	std::vector<double> ratios;
	std::unordered_set<int> cuspIndices;
	auto cuspRatios = path->GetCuspPointRatios();
	cuspRatios.insert(0.0);
	cuspRatios.insert(1.0);
	auto cuspRatioIt = cuspRatios.begin();
	for (double length = 0.0; length <= pathLength; length += pathInterpolation) {
		double cuspLength = *cuspRatioIt * pathLength;
		if (length >= cuspLength - pathInterpolation / 2 && length < cuspLength + pathInterpolation / 2) {
			cuspIndices.insert(ratios.size());
			ratios.push_back(*cuspRatioIt);
			cuspRatioIt++;
		} else {
			ratios.push_back(length / pathLength);
		}
	}
	std::vector<Pose2d> pts;
	try {
		pts = path->Interpolate(ratios);
	} catch (const std::out_of_range&) { // start already satisfies the goal: the composite path has no segment and
		return 0; // the reference's SearchPath throws here (path_composite.h:45-46); reported as "no samples"
	}
	long n = (long)pts.size();
	for (long i = 0; i < n && i < cap; i++) {
		poses[3 * i] = pts[i].x();
		poses[3 * i + 1] = pts[i].y();
		poses[3 * i + 2] = pts[i].theta;
		cusp[i] = cuspIndices.count((int)i) ? 1 : 0;
	}
	return n;
}

// N4: Obstacle::GetBoundaryWorldPosition of a RectangleShape / RegularPolygonShape / CircleShape at a pose
// (obstacle.cpp:87-127): kind 0 rectangle (a = dx, b = dy), 1 regular polygon (a = radius, b = count), 2 circle
// (a = radius, b = count).  Returns the number of vertices written to out (x, y pairs).
int ref_shape_vertices(int kind, double a, double b, double x, double y, double theta, double* out, int cap)
{
	Ref<Shape> shape;
	if (kind == 0)
		shape = makeRef<RectangleShape>(a, b);
	else if (kind == 1)
		shape = makeRef<RegularPolygonShape>(a, (int)b);
	else
		shape = makeRef<CircleShape>(a, (int)b);
	Obstacle o;
	o.SetShape(shape);
	o.SetPose(Pose2d(x, y, theta));
	auto pts = o.GetBoundaryWorldPosition();
	int n = 0;
	for (auto& p : pts) {
		if (n >= cap)
			break;
		out[2 * n] = p.x();
		out[2 * n + 1] = p.y();
		n++;
	}
	return n;
}

// N4: AddObstacle of such a shape into the world's map (obstacle_list_occupancy_map.cpp:29-45)
int ref_world_add_shape(void* h, int kind, double a, double b, double x, double y, double theta)
{
	auto* w = (RefWorld*)h;
	Ref<Shape> shape;
	if (kind == 0)
		shape = makeRef<RectangleShape>(a, b);
	else if (kind == 1)
		shape = makeRef<RegularPolygonShape>(a, (int)b);
	else
		shape = makeRef<CircleShape>(a, (int)b);
	auto o = makeRef<Obstacle>();
	o->SetShape(shape);
	o->SetPose(Pose2d(x, y, theta));
	w->obstacles.push_back(o);
	return w->map->AddObstacle(o) ? 1 : 0;
}

// N4: RemoveObstacle of the index-th obstacle added to the world (obstacle_list_occupancy_map.cpp:47-61)
int ref_world_remove_obstacle(void* h, int index)
{
	auto* w = (RefWorld*)h;
	if (index < 0 || index >= (int)w->obstacles.size())
		return 0;
	return w->map->RemoveObstacle(w->obstacles[index]) ? 1 : 0;
}

// N2: the nearest-obstacle / nearest-Voronoi-edge cell grids of the planner's GVD (after its last Update), (row, col)
// pairs, rows*cols*2 ints each -- the lookups Smoother::Smooth makes (smoother.cpp:110,118).
void ref_hybrid_gvd_cells(void* hh, int32_t* nearestObstacle, int32_t* nearestEdge)
{
	auto* h = (RefHybrid*)hh;
	auto& gvd = *h->planner->m_gvd;
	const int rows = h->world->map->Rows(), cols = h->world->map->Columns();
	for (int r = 0; r < rows; r++)
		for (int c = 0; c < cols; c++) {
			const auto& o = gvd.GetNearestObstacleCell(r, c);
			const auto& e = gvd.GetNearestVoronoiEdgeCell(r, c);
			size_t i = ((size_t)r * cols + c) * 2;
			nearestObstacle[i] = o.row; nearestObstacle[i + 1] = o.col;
			nearestEdge[i] = e.row; nearestEdge[i + 1] = e.col;
		}
}

// N2: Smoother::SetParameters of the planner's smoother; p = the 10 leading fields of ppx_smoother_params as doubles
// (step_tolerance, max_iterations, learning_rate, path/smooth/voronoi/collision/curvature weights, collision_ratio,
// max_curvature).
void ref_hybrid_smoother_set(void* hh, const double* p)
{
	auto* h = (RefHybrid*)hh;
	Smoother::Parameters q((float)p[9]);
	q.stepTolerance = (float)p[0];
	q.maxIterations = (int)p[1];
	q.learningRate = (float)p[2];
	q.pathWeight = (float)p[3];
	q.smoothWeight = (float)p[4];
	q.voronoiWeight = (float)p[5];
	q.collisionWeight = (float)p[6];
	q.curvatureWeight = (float)p[7];
	q.collisionRatio = (float)p[8];
	h->planner->m_smoother.SetParameters(q);
}

// N2: Smoother::Smooth (smoother.cpp:33-158) of the planner's own smoother on a given pose array + cusp flags.
// Returns Smoother::Status; out receives Smoother::GetPath().
int ref_hybrid_smooth(void* hh, long n, const double* poses, const uint8_t* cusp, double* out)
{
	auto* h = (RefHybrid*)hh;
	auto& sm = h->planner->m_smoother;
	std::vector<Pose2d> path;
	std::unordered_set<int> cuspIndices;
	for (long i = 0; i < n; i++) {
		Pose2d pose(poses[3 * i], poses[3 * i + 1], 0.0);
		pose.theta = poses[3 * i + 2]; // as SearchPath hands it over: no re-wrap of the interpolated heading
		path.push_back(pose);
		if (cusp[i])
			cuspIndices.insert((int)i);
	}
	int st = (int)sm.Smooth(path, cuspIndices);
	const auto& res = sm.GetPath();
	for (size_t i = 0; i < res.size() && (long)i < n; i++) {
		out[3 * i] = res[i].x(); out[3 * i + 1] = res[i].y(); out[3 * i + 2] = res[i].theta;
	}
	return st;
}

// Full reference SearchPath (wavefront + GVD + A* + resampling + smoothing): timing only.
int ref_hybrid_search_full(void* hh, const double* start, const double* goal, uint64_t seed, double* seconds, long* path_len)
{
	auto* h = (RefHybrid*)hh;
	auto& pl = *h->planner;
	pl.SetInitState(Pose2d(start[0], start[1], start[2]));
	pl.SetGoalState(Pose2d(goal[0], goal[1], goal[2]));
	SeedRandom(seed);
	h->propagator->record = false;
	auto t0 = std::chrono::steady_clock::now();
	Status st = pl.SearchPath();
	*seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	h->propagator->record = true;
	if (path_len) *path_len = (long)pl.GetPath().size();
	return (int)st;
}

// One call of StatePropagator::GetNeighborStates (A1) on an arbitrary parent, for unit-level
// parity of the successor kernel.  The wavefront for `goal` must already be current
// (call ref_hybrid_prepare_goal).  Children: pose(3), key(3), direction, cost, kind.
void ref_hybrid_prepare_goal(void* hh, const double* goal)
{
	auto* h = (RefHybrid*)hh;
	auto& pl = *h->planner;
	Pose2d g(goal[0], goal[1], goal[2]);
	pl.SetGoalState(g);
	pl.m_obstacleHeuristic->Update(g);
	pl.m_gvd->Update();
	auto goalState = pl.m_propagator->SetGoalState(g);
	pl.m_graphSearch.SetGoalState(goalState);
	pl.m_graphSearch.m_heuristic->SetGoal(goalState);
}

int ref_hybrid_expand(void* hh, const double* parent, uint64_t seed, double* poses /*3*11*/, int* keys /*3*11*/,
	int* directions, double* costs, int* kinds, double* heuristics, double* parent_heuristic)
{
	auto* h = (RefHybrid*)hh;
	auto& pl = *h->planner;
	Pose2d p;
	p.position = Point2d(parent[0], parent[1]);
	p.theta = parent[2];
	auto state = pl.m_propagator->CreateStateFromPose(p);
	SeedRandom(seed);
	bool rec = h->propagator->record;
	h->propagator->record = false;
	auto children = h->propagator->GetNeighborStates(state);
	h->propagator->record = rec;
	if (parent_heuristic)
		*parent_heuristic = pl.m_graphSearch.m_heuristic->GetHeuristicValue(state);
	int n = 0;
	for (auto& [child, action, cost] : children) {
		poses[3 * n + 0] = child.continuousPose.x();
		poses[3 * n + 1] = child.continuousPose.y();
		poses[3 * n + 2] = child.continuousPose.theta;
		keys[3 * n + 0] = child.discretePose.x();
		keys[3 * n + 1] = child.discretePose.y();
		keys[3 * n + 2] = child.discretePose.theta;
		directions[n] = (int)child.direction;
		costs[n] = cost;
		kinds[n] = dynamic_cast<PathReedsShepp*>(action.path.get()) ? 1 : 0;
		if (heuristics)
			heuristics[n] = pl.m_graphSearch.m_heuristic->GetHeuristicValue(child);
		n++;
	}
	return n;
}

// Combined heuristic (max of non-holonomic table and obstacle wavefront) for arbitrary poses.
void ref_hybrid_heuristic(void* hh, long n, const double* poses, double* values, double* nonholo, double* obst)
{
	auto* h = (RefHybrid*)hh;
	auto& pl = *h->planner;
	for (long i = 0; i < n; i++) {
		Pose2d p;
		p.position = Point2d(poses[3 * i], poses[3 * i + 1]);
		p.theta = poses[3 * i + 2];
		auto state = pl.m_propagator->CreateStateFromPose(p);
		if (values) values[i] = pl.m_graphSearch.m_heuristic->GetHeuristicValue(state);
		if (nonholo) nonholo[i] = pl.m_nonHoloHeuristic->GetHeuristicValue(p);
		if (obst) obst[i] = pl.m_obstacleHeuristic->GetHeuristicValue(p);
	}
}

// ---------------------------------------------------------------------------------------------
// Grid A* (BASELINE config C1, CPU-only): Euclidean step cost and heuristic as C++ functors,
// like interfaces/python/scripts/example_a_star_grid.py:46-52 does through Python callbacks.
int ref_astar_n2(void* hw, int r0, int c0, int r1, int c1, double* cost, long* explored, double* seconds,
	long cap, int* path_cells, long* path_len)
{
	auto* w = (RefWorld*)hw;
	auto euclid = [](const GridCellPosition& a, const GridCellPosition& b) {
		double dr = a.row - b.row, dc = a.col - b.col;
		return std::sqrt(dr * dr + dc * dc);
	};
	Ref<OccupancyMap> map = w->map;
	auto propagator = makeRef<AStarStatePropagatorFcnN2>(map, euclid);
	auto heuristic = makeRef<AStarHeuristicFcnN2>(euclid);
	AStarN2 astar;
	astar.Initialize(propagator, heuristic);
	astar.SetInitState(GridCellPosition(r0, c0));
	astar.SetGoalState(GridCellPosition(r1, c1));
	auto t0 = std::chrono::steady_clock::now();
	Status st = astar.SearchPath();
	*seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	*cost = astar.GetOptimalCost();
	*explored = (long)astar.GetExploredStates().size();
	auto path = astar.GetPath();
	*path_len = (long)path.size();
	for (long i = 0; i < (long)path.size() && i < cap; i++) {
		path_cells[2 * i] = path[i].row;
		path_cells[2 * i + 1] = path[i].col;
	}
	return (int)st;
}

// Grid A* with the functor table of ppx_astar_params (same formulas as the device's grid_metric), unidirectional (AStarN2) or
// BidirectionalAStarN2 with one heuristic object per direction.  cellcost: rows*cols floats for cost kind 3.
int ref_astar_n2_ex(void* hw, int r0, int c0, int r1, int c1, int costKind, int heurKind, const float* cellcost, int bidirectional,
	double* cost, long* explored, long cap, int* path_cells, long* path_len)
{
	auto* w = (RefWorld*)hw;
	const int cols = w->map->Columns();
	auto metric = [](int kind, const GridCellPosition& a, const GridCellPosition& b) {
		const double dr = std::fabs((double)(a.row - b.row)), dc = std::fabs((double)(a.col - b.col));
		if (kind == 1)
			return dr + dc;
		if (kind == 2)
			return std::fmax(dr, dc) + (std::sqrt(2.0) - 1.0) * std::fmin(dr, dc);
		return std::sqrt(dr * dr + dc * dc);
	};
	std::function<double(const GridCellPosition&, const GridCellPosition&)> costFcn = [=](const GridCellPosition& a, const GridCellPosition& b) {
		if (costKind == 3)
			return metric(0, a, b) * (double)cellcost[(size_t)b.row * cols + b.col];
		return metric(costKind, a, b);
	};
	std::function<double(const GridCellPosition&, const GridCellPosition&)> heurFcn = [=](const GridCellPosition& s, const GridCellPosition& g) {
		return heurKind == 3 ? 0.0 : metric(heurKind, s, g);
	};
	Ref<OccupancyMap> map = w->map;
	std::vector<GridCellPosition> path;
	Status st;
	if (!bidirectional) {
		AStarN2 astar;
		astar.Initialize(makeRef<AStarStatePropagatorFcnN2>(map, costFcn), makeRef<AStarHeuristicFcnN2>(heurFcn));
		astar.SetInitState(GridCellPosition(r0, c0));
		astar.SetGoalState(GridCellPosition(r1, c1));
		st = astar.SearchPath();
		*cost = astar.GetOptimalCost();
		*explored = (long)astar.GetExploredStates().size();
		path = astar.GetPath();
	} else {
		BidirectionalAStarN2 astar;
		astar.Initialize(makeRef<AStarStatePropagatorFcnN2>(map, costFcn), makeRef<AStarStatePropagatorFcnN2>(map, costFcn),
			makeRef<AStarHeuristicFcnN2>(heurFcn), makeRef<AStarHeuristicFcnN2>(heurFcn));
		astar.SetInitState(GridCellPosition(r0, c0));
		astar.SetGoalState(GridCellPosition(r1, c1));
		st = astar.SearchPath();
		*cost = st == Status::Success ? astar.GetOptimalCost() : std::numeric_limits<double>::infinity();
		auto [fe, re] = astar.GetExploredStates();
		*explored = (long)(fe.size() + re.size());
		if (st == Status::Success)
			path = astar.GetPath();
	}
	*path_len = (long)path.size();
	for (long i = 0; i < (long)path.size() && i < cap; i++) {
		path_cells[2 * i] = path[i].row;
		path_cells[2 * i + 1] = path[i].col;
	}
	return (int)st;
}

// ---------------------------------------------------------------------------------------------
// Tree / kNN (A10) through the reference's own utils/tree.h on top of the brute-force flann
// stand-in (oracle/shim/flann/flann.hpp).  2-D points (the only instantiation that compiles
// unmodified).
void ref_tree_knn_r2(long npoints, const double* points, long nqueries, const double* queries, int k, int* indices)
{
	using VertexTree = Tree<Point2d, 2>;
	VertexTree tree;
	std::unordered_map<VertexTree::Node*, int> index;
	VertexTree::Node* prev = tree.CreateRootNode(Point2d(points[0], points[1]));
	index[prev] = 0;
	for (long i = 1; i < npoints; i++) {
		auto* node = tree.Extend(Point2d(points[2 * i], points[2 * i + 1]), prev);
		if (index.find(node) == index.end())
			index[node] = (int)i;
		prev = node;
	}
	for (long q = 0; q < nqueries; q++) {
		auto near = tree.GetNearestNodes(Point2d(queries[2 * q], queries[2 * q + 1]), k);
		for (int j = 0; j < k; j++)
			indices[q * k + j] = j < (int)near.size() ? index[near[j]] : -1;
	}
}

// RRT / RRT* in R^2 with the reference templates (algo/rrt.h, algo/rrt_star.h) -- the aliases
// the reference itself instantiates (RRTR2, RRTStarR2).  Free-space validator as in
// planner/tests/test_rrt_star.cpp.  Returns status; nodes = tree size.
int ref_rrtstar_r2_free(const double* lb, const double* ub, const double* start, const double* goal, uint64_t seed,
	unsigned maxIteration, unsigned maxNodes, double maxConnectionDistance, double goalBias, int star,
	long* nodes, long cap, double* path_xy, long* path_len)
{
	auto space = makeRef<StateSpaceR2>(std::array<Point2d, 2> { Point2d(lb[0], lb[1]), Point2d(ub[0], ub[1]) });
	Ref<StateValidator<Point2d, 2, double>> validator = makeRef<StateValidatorFree<Point2d, 2, double>>(space);
	Ref<PathConnection<Point2d>> conn = makeRef<PathConnectionR2>();
	SeedRandom(seed);
	std::vector<Point2d> path;
	Status st;
	if (star) {
		RRTStarR2 rrt(space, validator, conn);
		RRTStarParameters p;
		p.maxIteration = maxIteration;
		p.maxNumberTreeNode = maxNodes;
		p.maxConnectionDistance = maxConnectionDistance;
		p.goalBias = goalBias;
		rrt.SetParameters(p);
		rrt.SetInitState(Point2d(start[0], start[1]));
		rrt.SetGoalState(Point2d(goal[0], goal[1]));
		SeedRandom(seed);
		st = rrt.SearchPath();
		*nodes = (long)rrt.m_tree.Size();
		path = rrt.GetPath();
	} else {
		RRTR2 rrt(space, validator, conn);
		RRTParameters p;
		p.maxIteration = maxIteration;
		p.maxNumberTreeNode = maxNodes;
		p.maxConnectionDistance = maxConnectionDistance;
		p.goalBias = goalBias;
		rrt.SetParameters(p);
		rrt.SetInitState(Point2d(start[0], start[1]));
		rrt.SetGoalState(Point2d(goal[0], goal[1]));
		SeedRandom(seed);
		st = rrt.SearchPath();
		*nodes = (long)rrt.m_tree.Size();
		path = rrt.GetPath();
	}
	*path_len = (long)path.size();
	for (long i = 0; i < (long)path.size() && i < cap; i++) {
		path_xy[2 * i] = path[i].x();
		path_xy[2 * i + 1] = path[i].y();
	}
	return (int)st;
}

// Same as above with the whole tree dumped: nodes in flann insertion order, parent indices, RRT*
// costs, iteration count is not observable from outside (the loop counter is a local).
int ref_rrt_r2_run(const double* lb, const double* ub, const double* start, const double* goal, uint64_t seed,
	unsigned maxIteration, unsigned maxNodes, double maxConnectionDistance, double goalBias, int star,
	long cap, long* nodes, double* nodes_xy, int* parents, double* costs, long* path_len, double* path_xy)
{
	auto space = makeRef<StateSpaceR2>(std::array<Point2d, 2> { Point2d(lb[0], lb[1]), Point2d(ub[0], ub[1]) });
	Ref<StateValidator<Point2d, 2, double>> validator = makeRef<StateValidatorFree<Point2d, 2, double>>(space);
	Ref<PathConnection<Point2d>> conn = makeRef<PathConnectionR2>();
	std::vector<Point2d> path;
	Status st;
	auto dump = [&](auto& rrt, bool withCost) {
		auto& tree = rrt.m_tree;
		long n = (long)tree.m_kdTree.size();
		*nodes = (long)tree.Size();
		std::unordered_map<const void*, int> indexOf;
		for (long i = 0; i < n; i++) {
			Point2d pt(tree.m_kdTree.getPoint(i));
			indexOf[tree.m_exploredNodeMap[pt]] = (int)i;
		}
		for (long i = 0; i < n && i < cap; i++) {
			Point2d pt(tree.m_kdTree.getPoint(i));
			auto* node = tree.m_exploredNodeMap[pt];
			nodes_xy[2 * i] = pt.x();
			nodes_xy[2 * i + 1] = pt.y();
			auto* parent = node->GetParent();
			parents[i] = parent ? indexOf[parent] : -1;
			if constexpr (std::is_same_v<std::decay_t<decltype(rrt)>, RRTStarR2>)
				costs[i] = (*node)->info.costFromGoal;
			else
				costs[i] = 0.0;
		}
	};
	if (star) {
		RRTStarR2 rrt(space, validator, conn);
		RRTStarParameters p;
		p.maxIteration = maxIteration;
		p.maxNumberTreeNode = maxNodes;
		p.maxConnectionDistance = maxConnectionDistance;
		p.goalBias = goalBias;
		rrt.SetParameters(p);
		rrt.SetInitState(Point2d(start[0], start[1]));
		rrt.SetGoalState(Point2d(goal[0], goal[1]));
		SeedRandom(seed);
		st = rrt.SearchPath();
		dump(rrt, true);
		path = rrt.GetPath();
	} else {
		RRTR2 rrt(space, validator, conn);
		RRTParameters p;
		p.maxIteration = maxIteration;
		p.maxNumberTreeNode = maxNodes;
		p.maxConnectionDistance = maxConnectionDistance;
		p.goalBias = goalBias;
		rrt.SetParameters(p);
		rrt.SetInitState(Point2d(start[0], start[1]));
		rrt.SetGoalState(Point2d(goal[0], goal[1]));
		SeedRandom(seed);
		st = rrt.SearchPath();
		dump(rrt, false);
		path = rrt.GetPath();
	}
	*path_len = (long)path.size();
	for (long i = 0; i < (long)path.size() && i < cap; i++) {
		path_xy[2 * i] = path[i].x();
		path_xy[2 * i + 1] = path[i].y();
	}
	return (int)st;
}

// RRT* over Pose2d.  The reference's own template does not compile for Pose2d (rrt_star.h:181 is
// commented out; utils/tree.h:88,111 cast double* to State), so this is a RESTATEMENT of the loop of
// rrt_star.h:53-112 and of Tree (utils/tree.h:59-146) -- "port", parity unpinned -- that calls the
// reference's own components for everything else: Random<double>, StateSpaceSE2::SampleUniform,
// PathConnectionReedsShepp / PathConnectionSE2 (Connect, Truncate with trap H9, GetLength,
// GetFinalState), StateValidatorOccupancyMap::IsPathValid or StateValidatorFree, std::equal_to<Pose2d>.
// Nearest neighbours: flann::L2_Simple over (x, y, theta), ties to the lower insertion index.
int ref_rrtstar_se2_restated(void* hw /* null: free-space validator */, const double* lb, const double* ub, const double* start,
	const double* goal, uint64_t seed, unsigned maxIteration, unsigned maxNodes, double maxConnectionDistance, double goalBias,
	int reedsShepp, double rmin, long cap, long* nodes, double* nodes_xyt, int* parents, double* costs, long* iterations,
	long* edges_checked, double* seconds)
{
	auto space = makeRef<StateSpaceSE2>(std::array<Pose2d, 2> { Pose2d(lb[0], lb[1], lb[2]), Pose2d(ub[0], ub[1], ub[2]) });
	Ref<StateValidator<Pose2d, 3, double>> validator;
	if (hw)
		validator = ((RefWorld*)hw)->validator;
	else
		validator = makeRef<StateValidatorFree<Pose2d, 3, double>>(space);
	Ref<PathConnection<Pose2d>> conn;
	if (reedsShepp)
		conn = makeRef<PathConnectionReedsShepp>(rmin);
	else
		conn = makeRef<PathConnectionSE2>();
	Pose2d init, goalState;
	init.position = Point2d(start[0], start[1]);
	init.theta = start[2];
	goalState.position = Point2d(goal[0], goal[1]);
	goalState.theta = goal[2];
	std::vector<Pose2d> states;
	std::vector<int> parent;
	std::vector<double> costFromGoal;
	std::equal_to<Pose2d> equal;
	// m_exploredNodeMap (utils/tree.h:127-129): the first node with an equal state.  A hash on the coordinate bits (-0.0
	// folded onto +0.0, the heading through WrapTheta as equal_to<Pose2d> compares it) keeps the lookup O(1); collisions
	// are resolved with the reference's equal_to.
	auto wrap = [](double t) {
		while (t > M_PI) t -= 2 * M_PI;
		while (t < -M_PI) t += 2 * M_PI;
		return t;
	};
	auto hashState = [&](const Pose2d& s) {
		uint64_t h = 1469598103934665603ULL;
		double v[3] = { s.x() + 0.0, s.y() + 0.0, wrap(s.theta) + 0.0 };
		for (double d : v) {
			uint64_t b;
			memcpy(&b, &d, 8);
			h = (h ^ b) * 1099511628211ULL;
		}
		return h;
	};
	std::unordered_multimap<uint64_t, int> index;
	auto find = [&](const Pose2d& s) {
		int best = -1;
		auto range = index.equal_range(hashState(s));
		for (auto it = range.first; it != range.second; ++it)
			if (equal(states[it->second], s) && (best < 0 || it->second < best))
				best = it->second;
		return best;
	};
	// flann::L2_Simple over (x, y, theta), k nearest by (distance, insertion index): one pass with a sorted top-k list
	auto knn = [&](const Pose2d& q, unsigned k) {
		const size_t n = std::min<size_t>(k, states.size());
		std::vector<std::pair<double, int>> best;
		best.reserve(n + 1);
		const double a[3] = { q.x(), q.y(), q.theta };
		for (size_t i = 0; i < states.size(); i++) {
			const double b[3] = { states[i].x(), states[i].y(), states[i].theta };
			double d = 0;
			for (int c = 0; c < 3; c++) {
				double diff = a[c] - b[c];
				d += diff * diff;
			}
			if (best.size() == n && !(d < best.back().first))
				continue; // a later index never displaces an equal distance
			auto pos = std::upper_bound(best.begin(), best.end(), std::make_pair(d, (int)i));
			best.insert(pos, { d, (int)i });
			if (best.size() > n)
				best.pop_back();
		}
		return best;
	};
	SeedRandom(seed);
	auto t0 = std::chrono::steady_clock::now();
	states.push_back(init);
	index.emplace(hashState(init), 0);
	parent.push_back(-1);
	costFromGoal.push_back(0.0);
	int count = -1, status = -1;
	long edges = 0;
	while (true) {
		count++;
		if (count > (int)maxIteration)
			break;
		if (states.size() > maxNodes)
			break;
		Pose2d randomState = Random<double>::SampleUniform(0, 1) < goalBias ? goalState : space->SampleUniform();
		int nearest = knn(randomState, 1)[0].second;
		auto pathNearToNew = conn->Connect(states[nearest], randomState);
		if (pathNearToNew->GetLength() > 0) {
			double ratio = std::clamp(maxConnectionDistance / pathNearToNew->GetLength(), 0.0, 1.0);
			pathNearToNew->Truncate(ratio);
		}
		edges++;
		if (!validator->IsPathValid(*pathNearToNew))
			continue;
		const Pose2d newState = pathNearToNew->GetFinalState();
		unsigned int nn = std::max<unsigned int>(1, log(states.size()));
		auto near = knn(newState, nn);
		int bestParent = -1;
		double bestCost = std::numeric_limits<double>::infinity();
		for (auto& [d, node] : near) {
			auto pathParentToNew = conn->Connect(states[node], newState);
			double cost = costFromGoal[node] + pathParentToNew->GetLength();
			if (cost < bestCost) {
				edges++;
				if (validator->IsPathValid(*pathParentToNew)) {
					bestParent = node;
					bestCost = cost;
				}
			}
		}
		int node = find(newState);
		if (node < 0) {
			node = (int)states.size();
			states.push_back(newState);
			index.emplace(hashState(newState), node);
			parent.push_back(bestParent >= 0 ? bestParent : near[0].second);
			costFromGoal.push_back(0.0);
		}
		costFromGoal[node] = bestCost;
		if (equal(newState, goalState)) {
			status = 0;
			break;
		}
	}
	if (seconds)
		*seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
	*nodes = (long)states.size();
	*iterations = count;
	*edges_checked = edges;
	for (long i = 0; i < (long)states.size() && i < cap; i++) {
		nodes_xyt[3 * i] = states[i].x();
		nodes_xyt[3 * i + 1] = states[i].y();
		nodes_xyt[3 * i + 2] = states[i].theta;
		parents[i] = parent[i];
		costs[i] = costFromGoal[i];
	}
	return status;
}

// Raw draws from the reference RNG wrapper after seeding (pins H4: one 64-bit draw per sample,
// SampleUniform = lb + (ub-lb)*u).
void ref_random_uniform(uint64_t seed, long n, double lb, double ub, double* out)
{
	SeedRandom(seed);
	for (long i = 0; i < n; i++)
		out[i] = Random<double>::SampleUniform(lb, ub);
}

} // extern "C"
