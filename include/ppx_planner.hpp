// C++ host side of libppx.so: the reference's planner-facing classes for the collision-checked
// expansion path, with the reference's names, argument meaning and error behaviour, implemented
// over the C ABI of ppx.h (no CUDA headers needed; link pathplanning_b200/libppx.so).
//
// Mirrors (reference file:line under planner/src):
//   Planner::Pose2d                       geometry/2dplane.h:17-94 (no Eigen: x/y/theta members + x(), y())
//   Planner::Status                       algo/path_planner.h:9-12
//   Planner::Steer / Direction            paths/path.h:9-19
//   Planner::OccupancyMap + GVD distance  state_validator/occupancy_map.h:101-193, gvd.h:38-39 (grids live on the device)
//   Planner::StateValidatorOccupancyMap   state_validator/state_validator_occupancy_map.h:13-32
//   Planner::PathReedsShepp / PathConnectionReedsShepp   paths/path_reeds_shepp.h:8-54
//   Planner::ObstaclesHeuristic           algo/heuristics.h:47-78
//   Planner::HybridAStar                  algo/hybrid_a_star.h:215-263
//   Planner::NonHolonomicHeuristic        algo/heuristics.h:9-40
//   Planner::AStarN2 / BidirectionalAStarN2  algo/a_star_n2.h:29-42
//   Planner::Tree / RRTR2 / RRTStarR2 / RRTStarSE2  utils/tree.h:27-177, algo/rrt.h:30-155, algo/rrt_star.h:29-181
// Batch entry points (names ending in Batch) are the additions the device makes worthwhile; the
// scalar methods keep the reference's signatures and run a batch of one.
// There is NO CPU fallback: constructing a Device without a CUDA device throws.
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>
#include <unordered_set>
#include "ppx.h"

namespace Planner {

template <typename T>
using Ref = std::shared_ptr<T>;

enum class Status { Success = 0, Failure = -1 }; // algo/path_planner.h:9-12
enum class Steer { Left = 0, Straight = 1, Right = 2 }; // paths/path.h:9-13
enum class Direction { Forward = 0, Backward = 1, NoMotion = 2 }; // paths/path.h:15-19

struct Point2d { // Eigen::Vector2d of geometry/2dplane.h:14
	double x = 0, y = 0;
};

struct Pose2d { // geometry/2dplane.h:17-34; the constructor wraps theta to (-pi, pi] like WrapTheta (:36-45)
	double px = 0, py = 0, theta = 0;
	Pose2d() = default;
	Pose2d(double x, double y, double th) : px(x), py(y), theta(th) { WrapTheta(); }
	double x() const { return px; }
	double y() const { return py; }
	void WrapTheta()
	{
		if (!(std::fabs(theta) < 1e15))
			return;
		while (theta > M_PI)
			theta -= 2 * M_PI;
		while (theta < -M_PI)
			theta += 2 * M_PI;
	}
};

// One GPU, one stream (ppx_ctx).  Shared by the map, the validator and the planners.
class Device {
public:
	explicit Device(int device = 0)
	{
		if (ppx_ctx_create(device, &m_ctx) != PPX_OK)
			throw std::runtime_error("libppx: no usable CUDA device (there is no CPU fallback)");
	}
	~Device() { ppx_ctx_destroy(m_ctx); }
	Device(const Device&) = delete;
	Device& operator=(const Device&) = delete;
	ppx_ctx* Context() const { return m_ctx; }
	/// ppx_ctx_set_option: "rs_filter", "rs_overlap" (include/ppx.h)
	void SetOption(const char* name, long long value) { Check(ppx_ctx_set_option(m_ctx, name, value)); }
	void Check(int rc) const
	{
		if (rc != PPX_OK)
			throw std::runtime_error(std::string("libppx: ") + ppx_last_error(m_ctx));
	}

private:
	ppx_ctx* m_ctx = nullptr;
};

// OccupancyMap (occupancy_map.h) with the GVD obstacle-distance map: occupancy is edited on the host
// (SetOccupied / cells >= 0 are occupied, :129-130); Update() uploads it and builds the squared
// distance map ON THE DEVICE (ppx_map_build, replaces GVD::ObstacleDistanceMap::Update, gvd.cpp:30-89).
class OccupancyMap {
public:
	const float resolution;
	/// InitializeSize(width, height) of occupancy_map.cpp:6-14: a map centred on `position`.
	OccupancyMap(const Ref<Device>& dev, float resolution_, float width, float height, const Pose2d& lb, const Pose2d& ub,
		double originX = 0.0, double originY = 0.0) :
		resolution(resolution_), m_dev(dev)
	{
		m_desc.rows = (int)std::ceil(width / resolution);
		m_desc.cols = (int)std::ceil(height / resolution);
		m_desc.resolution = resolution;
		m_desc.local_origin[0] = originX;
		m_desc.local_origin[1] = originY;
		m_desc.world_origin[0] = originX - (double)(width / 2);
		m_desc.world_origin[1] = originY - (double)(height / 2);
		m_desc.lb[0] = lb.x(); m_desc.lb[1] = lb.y(); m_desc.lb[2] = lb.theta;
		m_desc.ub[0] = ub.x(); m_desc.ub[1] = ub.y(); m_desc.ub[2] = ub.theta;
		m_occupancy.assign((size_t)m_desc.rows * m_desc.cols, -1);
	}
	int Rows() const { return m_desc.rows; }
	int Columns() const { return m_desc.cols; }
	/// WorldPositionToGridCell, occupancy_map.h:106-110 (row <-> x, col <-> y)
	bool WorldPositionToGridCell(double x, double y, int& row, int& col) const
	{
		row = (int)((x - m_desc.world_origin[0]) / (double)resolution);
		col = (int)((y - m_desc.world_origin[1]) / (double)resolution);
		return row >= 0 && row < m_desc.rows && col >= 0 && col < m_desc.cols;
	}
	void SetOccupied(int row, int col, int obstacleId = 0) { m_occupancy[(size_t)row * m_desc.cols + col] = obstacleId; m_dirty = true; }
	void SetFree(int row, int col) { m_occupancy[(size_t)row * m_desc.cols + col] = -1; m_dirty = true; }
	bool IsOccupied(int row, int col) const { return m_occupancy[(size_t)row * m_desc.cols + col] >= 0; }
	int GetOccupancyValue(int row, int col) const { return m_occupancy[(size_t)row * m_desc.cols + col]; }
	/// ObstacleListOccupancyMap::AddObstacle (obstacle_list_occupancy_map.cpp:29-45) for a polygon outline given by its
	/// world-frame vertices (Obstacle::GetBoundaryWorldPosition): rasterised on the device (ppx_map_rasterize); the
	/// obstacle receives the next id.  Returns the id.
	int AddObstacle(const std::vector<Point2d>& boundaryWorldPosition)
	{
		std::vector<double> v;
		for (auto& p : boundaryWorldPosition) {
			v.push_back(p.x);
			v.push_back(p.y);
		}
		const int32_t offsets[2] = { 0, (int32_t)boundaryWorldPosition.size() };
		const int id = m_nextId++;
		m_dev->Check(ppx_map_rasterize(m_dev->Context(), &m_desc, 1, offsets, v.data(), id, m_occupancy.data()));
		m_dirty = true;
		return id;
	}
	/// ObstacleListOccupancyMap::RemoveObstacle (obstacle_list_occupancy_map.cpp:47-61): the outline cells become free
	void RemoveObstacle(const std::vector<Point2d>& boundaryWorldPosition)
	{
		std::vector<double> v;
		for (auto& p : boundaryWorldPosition) {
			v.push_back(p.x);
			v.push_back(p.y);
		}
		const int32_t offsets[2] = { 0, (int32_t)boundaryWorldPosition.size() };
		m_dev->Check(ppx_map_rasterize(m_dev->Context(), &m_desc, 1, offsets, v.data(), -1, m_occupancy.data()));
		m_dirty = true;
	}
	/// OccupancyMap::Update() + GVD::Update() (occupancy_map.cpp:16-19): upload + distance map build.
	void Update()
	{
		m_sqdist.resize(m_occupancy.size());
		m_dev->Check(ppx_map_build(m_dev->Context(), &m_desc, m_occupancy.data(), nullptr, m_sqdist.data()));
		m_dirty = false;
	}
	/// GVD::ObstacleDistanceMap::GetDistanceToNearestObstacle, gvd.h:38-39
	float GetDistanceToNearestObstacle(int row, int col) const
	{
		return (float)(std::sqrt((double)m_sqdist[(size_t)row * m_desc.cols + col]) * (double)resolution);
	}
	const ppx_map_desc& Desc() const { return m_desc; }
	const Ref<Device>& GetDevice() const { return m_dev; }
	bool Dirty() const { return m_dirty; }

private:
	Ref<Device> m_dev;
	ppx_map_desc m_desc {};
	std::vector<int32_t> m_occupancy, m_sqdist;
	bool m_dirty = true;
	int m_nextId = 0;
};

// PathReedsShepp (paths/path_reeds_shepp.h:8-37): a solved segment anchored at its start pose.
class PathReedsShepp {
public:
	PathReedsShepp(const Ref<Device>& dev, const ppx_rs_edge& edge, double minTurningRadius, double length, float cost, int word) :
		m_dev(dev), m_edge(edge), m_minTurningRadius(minTurningRadius), m_length(length), m_cost(cost), m_word(word)
	{
	}
	double GetLength() const { return m_length; }
	float ComputeCost() const { return m_cost; } // with the connection's multipliers
	int GetWord() const { return m_word; } // ReedsShepp::PathWords value, -1 = NoPath
	Pose2d GetInitialState() const { return Pose2d(m_edge.start[0], m_edge.start[1], m_edge.start[2]); }
	/// Path<Pose2d>::Interpolate, path_reeds_shepp.cpp:12-47
	Pose2d Interpolate(double ratio) const { return InterpolateBatch({ ratio })[0]; }
	std::vector<Pose2d> InterpolateBatch(const std::vector<double>& ratios) const
	{
		std::vector<double> poses(3 * ratios.size());
		m_dev->Check(ppx_rs_interpolate(m_dev->Context(), 1, &m_edge, m_minTurningRadius, (int32_t)ratios.size(), ratios.data(), poses.data()));
		std::vector<Pose2d> out(ratios.size());
		for (size_t i = 0; i < ratios.size(); i++) {
			out[i].px = poses[3 * i];
			out[i].py = poses[3 * i + 1];
			out[i].theta = poses[3 * i + 2];
		}
		return out;
	}
	Pose2d GetFinalState() const { return Interpolate(1.0); }
	const ppx_rs_edge& Edge() const { return m_edge; }
	double GetMinTurningRadius() const { return m_minTurningRadius; }

private:
	Ref<Device> m_dev;
	ppx_rs_edge m_edge;
	double m_minTurningRadius, m_length;
	float m_cost;
	int m_word;
};

// PathConnectionReedsShepp (paths/path_reeds_shepp.h:40-54, .cpp:174-179)
class PathConnectionReedsShepp {
public:
	PathConnectionReedsShepp(const Ref<Device>& dev, double minTurningRadius = 1.0, double directionSwitchingCost = 0.0,
		double reverseCostMultiplier = 1.0, double forwardCostMultiplier = 1.0) :
		m_dev(dev)
	{
		m_cp.min_turning_radius = minTurningRadius;
		m_cp.direction_switching_cost = (float)directionSwitchingCost;
		m_cp.reverse_cost_multiplier = (float)reverseCostMultiplier;
		m_cp.forward_cost_multiplier = (float)forwardCostMultiplier;
		m_cp.shortest = 0;
	}
	Ref<PathReedsShepp> Connect(const Pose2d& from, const Pose2d& to) const { return ConnectBatch({ from }, { to })[0]; }
	std::vector<Ref<PathReedsShepp>> ConnectBatch(const std::vector<Pose2d>& from, const std::vector<Pose2d>& to) const
	{
		const size_t n = from.size();
		if (to.size() != n)
			throw std::invalid_argument("ConnectBatch: from/to sizes differ");
		std::vector<double> s(3 * n), g(3 * n), length(n);
		for (size_t i = 0; i < n; i++) {
			s[3 * i] = from[i].x(); s[3 * i + 1] = from[i].y(); s[3 * i + 2] = from[i].theta;
			g[3 * i] = to[i].x(); g[3 * i + 1] = to[i].y(); g[3 * i + 2] = to[i].theta;
		}
		std::vector<ppx_rs_edge> edges(n);
		std::vector<int32_t> word(n);
		std::vector<float> cost(n);
		m_dev->Check(ppx_rs_solve(m_dev->Context(), (int64_t)n, s.data(), g.data(), &m_cp, edges.data(), word.data(), length.data(), cost.data()));
		std::vector<Ref<PathReedsShepp>> out(n);
		for (size_t i = 0; i < n; i++)
			out[i] = std::make_shared<PathReedsShepp>(m_dev, edges[i], m_cp.min_turning_radius, length[i], cost[i], word[i]);
		return out;
	}
	const ppx_rs_cost_params& CostParams() const { return m_cp; }

private:
	Ref<Device> m_dev;
	ppx_rs_cost_params m_cp {};
};

// StateValidatorOccupancyMap (state_validator/state_validator_occupancy_map.h:13-32, .cpp:15-71)
class StateValidatorOccupancyMap {
public:
	float minPathInterpolationDistance = 0.1f; // :27
	float minSafeRadius = 1.0f; // :28
	explicit StateValidatorOccupancyMap(const Ref<OccupancyMap>& map) : m_map(map), m_dev(map->GetDevice()) { }
	bool IsStateValid(const Pose2d& state) const { return IsStateValidBatch({ state })[0] != 0; }
	std::vector<uint8_t> IsStateValidBatch(const std::vector<Pose2d>& states) const
	{
		std::vector<double> p(3 * states.size());
		for (size_t i = 0; i < states.size(); i++) {
			p[3 * i] = states[i].x(); p[3 * i + 1] = states[i].y(); p[3 * i + 2] = states[i].theta;
		}
		std::vector<uint8_t> valid(states.size());
		ppx_validator_params vp = Params();
		m_dev->Check(ppx_is_state_valid(m_dev->Context(), (int64_t)states.size(), p.data(), &vp, valid.data()));
		return valid;
	}
	/// IsPathValid(path, last): the adaptive sphere tracing of .cpp:28-71; *last = ratio of the last valid sample
	bool IsPathValid(const PathReedsShepp& path, float* last = nullptr) const
	{
		uint8_t v;
		float l;
		ppx_validator_params vp = Params();
		m_dev->Check(ppx_check_edges_rs(m_dev->Context(), 1, &path.Edge(), path.GetMinTurningRadius(), &vp, &v, &l, nullptr));
		if (last)
			*last = l;
		return v != 0;
	}
	/// The batched form: all edges share one turning radius (one PathConnectionReedsShepp).
	void IsPathValidBatch(const std::vector<ppx_rs_edge>& edges, double minTurningRadius, std::vector<uint8_t>& valid,
		std::vector<float>& last) const
	{
		valid.resize(edges.size());
		last.resize(edges.size());
		ppx_validator_params vp = Params();
		m_dev->Check(ppx_check_edges_rs(m_dev->Context(), (int64_t)edges.size(), edges.data(), minTurningRadius, &vp, valid.data(),
			last.data(), nullptr));
	}
	ppx_validator_params Params() const { return ppx_validator_params { minSafeRadius, minPathInterpolationDistance }; }
	const Ref<OccupancyMap>& GetMap() const { return m_map; }

private:
	Ref<OccupancyMap> m_map;
	Ref<Device> m_dev;
};

// ObstaclesHeuristic (algo/heuristics.h:47-78, heuristics.cpp:97-165): the wavefront runs on the device,
// GetHeuristicValue is the reference's host lookup on the returned cost grid.
class ObstaclesHeuristic {
public:
	ObstaclesHeuristic(const Ref<OccupancyMap>& map, double reverseCostMultiplier, double forwardCostMultiplier) :
		m_map(map), m_diagonalResolution(std::sqrt(2.0) * map->resolution),
		m_costMultiplier(std::min(reverseCostMultiplier, forwardCostMultiplier) * map->resolution)
	{
	}
	void Update(const Pose2d& goal)
	{
		m_goal = goal;
		m_cost.resize((size_t)m_map->Rows() * m_map->Columns());
		double g[3] = { goal.x(), goal.y(), goal.theta };
		auto& dev = m_map->GetDevice();
		dev->Check(ppx_wavefront(dev->Context(), 1, g, m_cost.data(), nullptr));
	}
	double GetHeuristicValue(const Pose2d& state) const
	{
		double dx = m_goal.x() - state.x(), dy = m_goal.y() - state.y();
		double euclidean = std::sqrt(dx * dx + dy * dy);
		int row, col;
		if (!m_map->WorldPositionToGridCell(state.x(), state.y(), row, col))
			return euclidean;
		float cost = m_cost[(size_t)row * m_map->Columns() + col];
		if (std::isinf(cost))
			return euclidean; // !m_explored
		double heuristic = cost * m_costMultiplier - m_diagonalResolution;
		return std::max(heuristic, euclidean);
	}

private:
	Ref<OccupancyMap> m_map;
	const double m_diagonalResolution, m_costMultiplier;
	Pose2d m_goal;
	std::vector<float> m_cost;
};

// HybridAStar (algo/hybrid_a_star.h:215-263, hybrid_a_star.cpp:206-304): graph search on the device.
// HybridAStar does not run the smoother itself (class Smoother below does, given the GVD cell grids); GetPath() returns the pose array
// SearchPath builds for the smoother (hybrid_a_star.cpp:260-291, resampled on the device every
// pathInterpolation with the cusp points snapped in) -- what the reference's GetPath() returns when
// smoothing does not succeed (:294-297); GetGraphSearchStates() returns the node states of the solution.
class HybridAStar {
public:
	struct SearchParameters { // hybrid_a_star.h:29-38, same defaults
		double wheelbase = 2.6;
		double minTurningRadius = 2.0;
		double directionSwitchingCost = 0.0;
		double reverseCostMultiplier = 1.0;
		double forwardCostMultiplier = 1.0;
		double voronoiCostMultiplier = 1.0;
		unsigned int numGeneratedMotion = 5;
		double spatialResolution = 1.0;
		double angularResolution = 0.0872;
	};
	struct Stats {
		Status graphSearchStatus = Status::Failure;
		int numExpandedNodes = 0, numGeneratedNodes = 0, numReedsSheppAttempts = 0;
	};

	HybridAStar() = default;
	explicit HybridAStar(const SearchParameters& p) : m_params(p) { }
	~HybridAStar()
	{
		if (m_gpu)
			ppx_hybrid_destroy(m_gpu);
	}
	bool Initialize(const Ref<StateValidatorOccupancyMap>& validator)
	{
		m_validator = validator;
		auto& dev = validator->GetMap()->GetDevice();
		if (validator->GetMap()->Dirty())
			validator->GetMap()->Update();
		ppx_hybrid_params hp {};
		hp.wheelbase = m_params.wheelbase;
		hp.min_turning_radius = m_params.minTurningRadius;
		hp.direction_switching_cost = m_params.directionSwitchingCost;
		hp.reverse_cost_multiplier = m_params.reverseCostMultiplier;
		hp.forward_cost_multiplier = m_params.forwardCostMultiplier;
		hp.voronoi_cost_multiplier = m_params.voronoiCostMultiplier;
		hp.num_generated_motion = m_params.numGeneratedMotion;
		hp.spatial_resolution = m_params.spatialResolution;
		hp.angular_resolution = m_params.angularResolution;
		ppx_validator_params vp = validator->Params();
		if (m_gpu) {
			ppx_hybrid_destroy(m_gpu);
			m_gpu = nullptr;
		}
		isInitialized = ppx_hybrid_create(dev->Context(), &hp, &vp, &m_gpu) == PPX_OK;
		return isInitialized;
	}
	void SetInitState(const Pose2d& s) { m_init = s; }
	void SetGoalState(const Pose2d& s) { m_goal = s; }
	const Pose2d& GetInitState() const { return m_init; }
	const Pose2d& GetGoalState() const { return m_goal; }
	/// seed of the std::mt19937_64 that gates the Reeds-Shepp shots (utils/random.h:12-27, trap H4)
	void SetSeed(uint64_t seed) { m_seed = seed; }

	Status SearchPath()
	{
		if (!isInitialized)
			return Status::Failure; // hybrid_a_star.cpp:240-243
		std::vector<Status> st;
		std::vector<std::vector<Pose2d>> paths;
		std::vector<double> costs;
		std::vector<Stats> stats;
		std::vector<std::vector<Pose2d>> resampled;
		std::vector<std::vector<int>> cusps;
		SearchPathBatch({ m_init }, { m_goal }, { m_seed }, st, paths, costs, &stats, 1024, &resampled, &cusps);
		m_states = paths[0];
		m_path = resampled[0];
		m_cuspIndices = cusps[0];
		m_cost = costs[0];
		m_stats = stats[0];
		return st[0];
	}
	/// n independent queries in one device call (BASELINE config C4)
	void SearchPathBatch(const std::vector<Pose2d>& inits, const std::vector<Pose2d>& goals, const std::vector<uint64_t>& seeds,
		std::vector<Status>& status, std::vector<std::vector<Pose2d>>& paths, std::vector<double>& costs,
		std::vector<Stats>* stats = nullptr, int maxPath = 1024, std::vector<std::vector<Pose2d>>* resampled = nullptr,
		std::vector<std::vector<int>>* cuspIndices = nullptr, int maxPoints = 4096)
	{
		if (!isInitialized)
			throw std::runtime_error("HybridAStar: not initialized");
		const size_t n = inits.size();
		if (goals.size() != n || seeds.size() != n)
			throw std::invalid_argument("SearchPathBatch: sizes differ");
		std::vector<double> s(3 * n), g(3 * n);
		for (size_t i = 0; i < n; i++) {
			s[3 * i] = inits[i].x(); s[3 * i + 1] = inits[i].y(); s[3 * i + 2] = inits[i].theta;
			g[3 * i] = goals[i].x(); g[3 * i + 1] = goals[i].y(); g[3 * i + 2] = goals[i].theta;
		}
		ppx_hybrid_limits lim { 0, 0, maxPath, 0 };
		std::vector<ppx_hybrid_result> res(n);
		std::vector<ppx_hybrid_path_state> chain(n * (size_t)maxPath);
		auto& dev = m_validator->GetMap()->GetDevice();
		dev->Check(ppx_hybrid_search(m_gpu, (int64_t)n, s.data(), g.data(), seeds.data(), &lim, res.data(), chain.data(), nullptr));
		status.resize(n);
		paths.assign(n, {});
		costs.resize(n);
		if (stats)
			stats->resize(n);
		for (size_t i = 0; i < n; i++) {
			status[i] = res[i].status == 0 ? Status::Success : Status::Failure;
			costs[i] = res[i].cost;
			int len = std::min(res[i].path_len, maxPath);
			for (int k = 0; k < len; k++) {
				const auto& c = chain[i * (size_t)maxPath + k];
				Pose2d p;
				p.px = c.pose[0]; p.py = c.pose[1]; p.theta = c.pose[2];
				paths[i].push_back(p);
			}
			if (stats) {
				(*stats)[i].graphSearchStatus = status[i];
				(*stats)[i].numExpandedNodes = res[i].n_expanded;
				(*stats)[i].numGeneratedNodes = res[i].n_nodes;
				(*stats)[i].numReedsSheppAttempts = res[i].n_rs_attempts;
			}
		}
		if (!resampled)
			return;
		// "Process graph search path" (hybrid_a_star.cpp:260-291) on the device
		std::vector<double> poses(n * (size_t)maxPoints * 3);
		std::vector<uint8_t> cusp(n * (size_t)maxPoints);
		std::vector<int32_t> count(n);
		dev->Check(ppx_hybrid_resample(m_gpu, (int64_t)n, g.data(), res.data(), chain.data(), maxPath, pathInterpolation, maxPoints,
			poses.data(), cusp.data(), count.data(), nullptr));
		resampled->assign(n, {});
		if (cuspIndices)
			cuspIndices->assign(n, {});
		for (size_t i = 0; i < n; i++) {
			if (count[i] < 0)
				throw std::length_error("HybridAStar: maxPath / maxPoints too small for the solution");
			for (int k = 0; k < count[i]; k++) {
				const double* q = &poses[(i * (size_t)maxPoints + k) * 3];
				Pose2d p;
				p.px = q[0]; p.py = q[1]; p.theta = q[2];
				(*resampled)[i].push_back(p);
				if (cuspIndices && cusp[i * (size_t)maxPoints + k])
					(*cuspIndices)[i].push_back(k);
			}
		}
	}
	std::vector<Pose2d> GetPath() const { return m_path; }
	/// states of the solution's nodes (root .. goal), the vertices of the reference's GetGraphSearchPath()
	std::vector<Pose2d> GetGraphSearchStates() const { return m_states; }
	/// indices into GetPath() of the cusp points (the cuspIndices SearchPath hands to the smoother)
	const std::vector<int>& GetCuspIndices() const { return m_cuspIndices; }
	double GetGraphSearchOptimalCost() const { return m_cost; }
	const Stats& GetStats() const { return m_stats; }
	const SearchParameters& GetSearchParameters() const { return m_params; }
	bool isInitialized = false;
	float pathInterpolation = 0.1f; // hybrid_a_star.h:249

private:
	SearchParameters m_params;
	Ref<StateValidatorOccupancyMap> m_validator;
	ppx_hybrid* m_gpu = nullptr;
	Pose2d m_init, m_goal;
	uint64_t m_seed = 0;
	std::vector<Pose2d> m_path, m_states;
	std::vector<int> m_cuspIndices;
	double m_cost = std::numeric_limits<double>::infinity();
	Stats m_stats;
};

// Smoother (algo/smoother.h:17-92): gradient-descent smoothing of pose arrays on the device (ppx_smoother_*).
// The GVD lookups Smooth makes are cell grids the caller provides (SetGVD) -- the Voronoi half of the GVD is not
// built on the device; with the reference tree they are GVD::GetNearestObstacleCell / GetNearestVoronoiEdgeCell.
class Smoother {
public:
	enum Status { MaxIteration = 0, StepTolerance, PathSize, Failure = -1, Collision = -2 }; // smoother.h:19-25
	struct Parameters { // smoother.h:29-62, same defaults
		float stepTolerance = 1e-3f;
		int maxIterations = 2000;
		float learningRate = 0.01f;
		float pathWeight = 0.0f, smoothWeight = 0.4f, voronoiWeight = 0.02f, collisionWeight = 0.2f, curvatureWeight = 0.4f;
		float collisionRatio = 0.2f;
		float maxCurvature;
		explicit Parameters(float maxCurvature_) : maxCurvature(maxCurvature_) { }
	};
	float alpha = 20.0f, dMax = 30.0f; // GVD::alpha, GVD::dMax (gvd.h:181)

	Smoother() : m_param(1.0f) { }
	~Smoother()
	{
		if (m_gpu)
			ppx_smoother_destroy(m_gpu);
	}
	bool Initialize(const Ref<StateValidatorOccupancyMap>& validator, float maxCurvature)
	{
		if (!validator) {
			isInitialized = false;
			return false;
		}
		m_param = Parameters(maxCurvature);
		m_validator = validator;
		m_dirty = true;
		isInitialized = true;
		return true;
	}
	Parameters GetParameters() const { return m_param; }
	void SetParameters(const Parameters& p)
	{
		m_param = p;
		m_dirty = true;
	}
	/// rows * cols (row, col) pairs each, after GVD::Update()
	void SetGVD(const std::vector<int32_t>& nearestObstacleCell, const std::vector<int32_t>& nearestVoronoiEdgeCell)
	{
		m_nearObstacle = nearestObstacleCell;
		m_nearEdge = nearestVoronoiEdgeCell;
		m_dirty = true;
	}
	Status Smooth(const std::vector<Pose2d>& path, const std::unordered_set<int>& cuspIndices = {})
	{
		if (!isInitialized)
			return Status::Failure; // smoother.cpp:39-42
		std::vector<Status> st;
		std::vector<std::vector<Pose2d>> out;
		SmoothBatch({ path }, { cuspIndices }, st, out);
		m_currentPath = out[0];
		return st[0];
	}
	/// n paths in one device call
	void SmoothBatch(const std::vector<std::vector<Pose2d>>& paths, const std::vector<std::unordered_set<int>>& cuspIndices,
		std::vector<Status>& status, std::vector<std::vector<Pose2d>>& out)
	{
		if (!isInitialized)
			throw std::runtime_error("Smoother: not initialized");
		const size_t n = paths.size();
		if (cuspIndices.size() != n)
			throw std::invalid_argument("SmoothBatch: sizes differ");
		auto& dev = m_validator->GetMap()->GetDevice();
		if (m_dirty) {
			if (m_gpu)
				ppx_smoother_destroy(m_gpu);
			m_gpu = nullptr;
			ppx_smoother_params p { m_param.stepTolerance, m_param.maxIterations, m_param.learningRate, m_param.pathWeight,
				m_param.smoothWeight, m_param.voronoiWeight, m_param.collisionWeight, m_param.curvatureWeight, m_param.collisionRatio,
				m_param.maxCurvature, alpha, dMax, m_validator->minSafeRadius };
			dev->Check(ppx_smoother_create(dev->Context(), &p, &m_gpu));
			dev->Check(ppx_smoother_set_gvd(m_gpu, m_nearObstacle.data(), m_nearEdge.data()));
			m_dirty = false;
		}
		size_t maxPoints = 1;
		for (auto& p : paths)
			maxPoints = std::max(maxPoints, p.size());
		std::vector<double> poses(n * maxPoints * 3), res(n * maxPoints * 3);
		std::vector<uint8_t> cusp(n * maxPoints);
		std::vector<int32_t> count(n), st(n);
		for (size_t i = 0; i < n; i++) {
			count[i] = (int32_t)paths[i].size();
			for (size_t k = 0; k < paths[i].size(); k++) {
				double* q = &poses[(i * maxPoints + k) * 3];
				q[0] = paths[i][k].x(); q[1] = paths[i][k].y(); q[2] = paths[i][k].theta;
				cusp[i * maxPoints + k] = cuspIndices[i].count((int)k) ? 1 : 0;
			}
		}
		dev->Check(ppx_smoother_smooth(m_gpu, (int64_t)n, poses.data(), cusp.data(), count.data(), (int32_t)maxPoints, res.data(),
			st.data(), nullptr));
		status.resize(n);
		out.assign(n, {});
		for (size_t i = 0; i < n; i++) {
			status[i] = (Status)st[i];
			for (int k = 0; k < count[i]; k++) {
				const double* q = &res[(i * maxPoints + k) * 3];
				Pose2d p;
				p.px = q[0]; p.py = q[1]; p.theta = q[2];
				out[i].push_back(p);
			}
		}
	}
	const std::vector<Pose2d>& GetPath() const { return m_currentPath; }
	bool isInitialized = false;

private:
	Parameters m_param;
	Ref<StateValidatorOccupancyMap> m_validator;
	ppx_smoother* m_gpu = nullptr;
	bool m_dirty = true;
	std::vector<int32_t> m_nearObstacle, m_nearEdge;
	std::vector<Pose2d> m_currentPath;
};

// AStarN2 (algo/a_star_n2.h:29-34) over AStarStatePropagatorFcnN2 (a_star_n2.cpp:12-28) with the Euclidean step cost and
// heuristic of example_a_star_grid.py -- the reference takes both as std::function callbacks, which cannot cross the C ABI.
struct GridCellPosition { // utils/grid.h:8-21
	int row = -1, col = -1;
	bool operator==(const GridCellPosition& o) const { return row == o.row && col == o.col; }
};

class AStarN2 {
public:
	bool Initialize(const Ref<OccupancyMap>& map)
	{
		m_map = map;
		isInitialized = (bool)map;
		return isInitialized;
	}
	void SetInitState(const GridCellPosition& c) { m_init = c; }
	void SetGoalState(const GridCellPosition& c) { m_goal = c; }
	Status SearchPath()
	{
		if (!isInitialized)
			return Status::Failure; // a_star.h:330-333
		if (m_map->Dirty())
			m_map->Update();
		std::vector<Status> st;
		std::vector<std::vector<GridCellPosition>> paths;
		std::vector<double> costs;
		std::vector<int> explored;
		SearchPathBatch({ m_init }, { m_goal }, st, paths, costs, &explored);
		m_path = paths[0];
		m_cost = costs[0];
		m_numExplored = explored[0];
		return st[0];
	}
	/// n independent queries in one device call
	void SearchPathBatch(const std::vector<GridCellPosition>& inits, const std::vector<GridCellPosition>& goals, std::vector<Status>& status,
		std::vector<std::vector<GridCellPosition>>& paths, std::vector<double>& costs, std::vector<int>* numExplored = nullptr,
		int maxPath = 4096)
	{
		if (!isInitialized)
			throw std::runtime_error("AStarN2: not initialized");
		const size_t n = inits.size();
		if (goals.size() != n)
			throw std::invalid_argument("SearchPathBatch: sizes differ");
		std::vector<int32_t> s(2 * n), g(2 * n), cells(n * (size_t)maxPath * 2);
		for (size_t i = 0; i < n; i++) {
			s[2 * i] = inits[i].row; s[2 * i + 1] = inits[i].col;
			g[2 * i] = goals[i].row; g[2 * i + 1] = goals[i].col;
		}
		std::vector<ppx_astar_result> res(n);
		auto& dev = m_map->GetDevice();
		dev->Check(ppx_astar_n2(dev->Context(), (int64_t)n, s.data(), g.data(), 0, maxPath, res.data(), cells.data()));
		status.resize(n);
		paths.assign(n, {});
		costs.resize(n);
		if (numExplored)
			numExplored->resize(n);
		for (size_t i = 0; i < n; i++) {
			status[i] = res[i].status == 0 ? Status::Success : Status::Failure;
			costs[i] = res[i].cost;
			if (numExplored)
				(*numExplored)[i] = res[i].n_explored;
			for (int k = 0; k < std::min(res[i].path_len, maxPath); k++)
				paths[i].push_back({ cells[(i * (size_t)maxPath + k) * 2], cells[(i * (size_t)maxPath + k) * 2 + 1] });
		}
	}
	std::vector<GridCellPosition> GetPath() const { return m_path; }
	double GetOptimalCost() const { return m_cost; }
	int GetNumExploredStates() const { return m_numExplored; } // GetExploredStates().size()
	bool isInitialized = false;

private:
	Ref<OccupancyMap> m_map;
	GridCellPosition m_init, m_goal;
	std::vector<GridCellPosition> m_path;
	double m_cost = std::numeric_limits<double>::infinity();
	int m_numExplored = 0;
};

// BidirectionalAStarN2 (algo/a_star_n2.h:36-42, algo/bidirectional_a_star.h:41-204) and AStarN2 with a functor of the
// device's table (ppx_astar_params): the reference's std::function callbacks cannot cross the C ABI.
enum class GridMetric { Euclidean = 0, Manhattan = 1, Octile = 2, ZeroOrCellCost = 3 };

class BidirectionalAStarN2 {
public:
	bool Initialize(const Ref<OccupancyMap>& map, GridMetric pathCost = GridMetric::Euclidean, GridMetric heuristic = GridMetric::Euclidean,
		std::vector<float> cellCost = {})
	{
		if (!map)
			return isInitialized = false;
		m_map = map;
		m_cost = pathCost;
		m_heur = heuristic;
		m_cellCost = std::move(cellCost);
		return isInitialized = true;
	}
	void SetInitState(const GridCellPosition& s) { m_init = s; }
	void SetGoalState(const GridCellPosition& s) { m_goal = s; }
	Status SearchPath()
	{
		if (!isInitialized)
			return Status::Failure;
		if (m_map->Dirty())
			m_map->Update();
		const int maxPath = 8192;
		const int32_t s[2] = { m_init.row, m_init.col }, g[2] = { m_goal.row, m_goal.col };
		std::vector<int32_t> cells((size_t)maxPath * 2);
		ppx_astar_params prm {};
		prm.cost_kind = (int32_t)m_cost;
		prm.heuristic_kind = (int32_t)m_heur;
		prm.cell_cost = m_cellCost.empty() ? nullptr : m_cellCost.data();
		prm.bidirectional = 1;
		ppx_astar_result res {};
		auto& dev = m_map->GetDevice();
		dev->Check(ppx_astar_n2_ex(dev->Context(), 1, s, g, &prm, 0, maxPath, &res, cells.data()));
		m_path.clear();
		for (int k = 0; k < std::min(res.path_len, maxPath); k++)
			m_path.push_back({ cells[2 * k], cells[2 * k + 1] });
		m_optimalCost = res.cost;
		m_numExplored = res.n_explored;
		return res.status == 0 ? Status::Success : Status::Failure;
	}
	std::vector<GridCellPosition> GetPath() const { return m_path; }
	double GetOptimalCost() const { return m_optimalCost; }
	int GetNumExploredStates() const { return m_numExplored; } // forward + reverse explored maps
	bool isInitialized = false;

private:
	Ref<OccupancyMap> m_map;
	GridMetric m_cost = GridMetric::Euclidean, m_heur = GridMetric::Euclidean;
	std::vector<float> m_cellCost;
	GridCellPosition m_init, m_goal;
	std::vector<GridCellPosition> m_path;
	double m_optimalCost = std::numeric_limits<double>::infinity();
	int m_numExplored = 0;
};

// NonHolonomicHeuristic (algo/heuristics.h:9-40, heuristics.cpp:7-95): the Reeds-Shepp cost table over (x, y, theta), built
// on the device by ppx_hybrid_create; GetHeuristicValue is the table lookup of :78-95 (negative heading indices read
// index + numAngular).
class NonHolonomicHeuristic {
public:
	static Ref<NonHolonomicHeuristic> Build(const Ref<OccupancyMap>& map, double spatialResolution = 1.0, double angularResolution = 0.0872,
		double minTurningRadius = 2.0, double reverseCostMultiplier = 1.0, double forwardCostMultiplier = 1.0,
		double directionSwitchingCost = 0.0)
	{
		auto h = Ref<NonHolonomicHeuristic>(new NonHolonomicHeuristic());
		h->m_map = map;
		if (map->Dirty())
			map->Update();
		ppx_hybrid_params hp {};
		hp.wheelbase = 2.6;
		hp.min_turning_radius = minTurningRadius;
		hp.direction_switching_cost = directionSwitchingCost;
		hp.reverse_cost_multiplier = reverseCostMultiplier;
		hp.forward_cost_multiplier = forwardCostMultiplier;
		hp.voronoi_cost_multiplier = 1.0;
		hp.num_generated_motion = 5;
		hp.spatial_resolution = spatialResolution;
		hp.angular_resolution = angularResolution;
		ppx_validator_params vp { 1.0f, 0.1f };
		auto& dev = map->GetDevice();
		dev->Check(ppx_hybrid_create(dev->Context(), &hp, &vp, &h->m_handle));
		int32_t dims[3];
		double off[2];
		dev->Check(ppx_hybrid_table(h->m_handle, dims, off, nullptr));
		h->numSpatialX = (unsigned)dims[0];
		h->numSpatialY = (unsigned)dims[1];
		h->numAngular = (unsigned)dims[2];
		h->offsetX = off[0];
		h->offsetY = off[1];
		return h;
	}
	~NonHolonomicHeuristic()
	{
		if (m_handle)
			ppx_hybrid_destroy(m_handle);
	}
	void SetGoal(const Pose2d& goal) { m_goal = goal; }
	double GetHeuristicValue(const Pose2d& state) { return GetHeuristicValues({ state })[0]; }
	std::vector<double> GetHeuristicValues(const std::vector<Pose2d>& states)
	{
		std::vector<double> p(3 * states.size()), out(states.size());
		for (size_t i = 0; i < states.size(); i++) {
			p[3 * i] = states[i].x();
			p[3 * i + 1] = states[i].y();
			p[3 * i + 2] = states[i].theta;
		}
		const double g[3] = { m_goal.x(), m_goal.y(), m_goal.theta };
		auto& dev = m_map->GetDevice();
		dev->Check(ppx_hybrid_heuristic(m_handle, g, (int64_t)states.size(), p.data(), nullptr, out.data(), nullptr));
		return out;
	}
	/// m_values[i][j][k], row-major
	std::vector<double> GetTable()
	{
		std::vector<double> v((size_t)numSpatialX * numSpatialY * numAngular);
		m_map->GetDevice()->Check(ppx_hybrid_table(m_handle, nullptr, nullptr, v.data()));
		return v;
	}
	unsigned numSpatialX = 0, numSpatialY = 0, numAngular = 0;
	double offsetX = 0, offsetY = 0;

private:
	NonHolonomicHeuristic() = default;
	Ref<OccupancyMap> m_map;
	ppx_hybrid* m_handle = nullptr;
	Pose2d m_goal;
};

// Tree<Vertex, Dim> (utils/tree.h:27-177), the nearest-neighbour and bookkeeping part: nodes are indices in insertion
// order (flann's point order), Extend refuses an existing state like :127-129, the queries are exact brute-force k-NN on
// the device (ppx_knn, replaces flann::Index<L2_Simple<double>>, ties to the lower index).
template <int Dim>
class Tree {
public:
	using State = std::array<double, Dim>;
	explicit Tree(const Ref<Device>& dev) : m_dev(dev) { }
	void Clear()
	{
		m_points.clear();
		m_parent.clear();
	}
	int CreateRootNode(const State& s)
	{
		Clear();
		return Append(s, -1);
	}
	/// node index of the new node, or of the existing node with an equal state (whose parent is left unchanged)
	int Extend(const State& s, int parent)
	{
		for (size_t i = 0; i < m_parent.size(); i++)
			if (std::equal(s.begin(), s.end(), m_points.begin() + i * Dim))
				return (int)i;
		return Append(s, parent);
	}
	int Size() const { return (int)m_parent.size(); }
	int GetParent(int node) const { return m_parent[node]; }
	State GetState(int node) const
	{
		State s;
		std::copy(m_points.begin() + (size_t)node * Dim, m_points.begin() + (size_t)(node + 1) * Dim, s.begin());
		return s;
	}
	int GetNearestNode(const State& s) const { return GetNearestNodes(s, 1)[0]; }
	std::vector<int> GetNearestNodes(const State& s, unsigned int nn) const
	{
		std::vector<int32_t> idx(nn);
		m_dev->Check(ppx_knn(m_dev->Context(), Dim, Size(), m_points.data(), 1, s.data(), (int32_t)nn, idx.data(), nullptr));
		std::vector<int> out;
		for (auto i : idx)
			if (i >= 0)
				out.push_back(i);
		return out;
	}

private:
	int Append(const State& s, int parent)
	{
		m_points.insert(m_points.end(), s.begin(), s.end());
		m_parent.push_back(parent);
		return Size() - 1;
	}
	Ref<Device> m_dev;
	std::vector<double> m_points;
	std::vector<int> m_parent;
};

// RRT / RRTStar (algo/rrt.h:30-152, algo/rrt_star.h:29-178): the whole loop runs on the device (ppx_rrt_run), one tree per
// call here; the reference's template parameters collapse to the state kinds it instantiates or sketches -- R^2 with
// straight connections (RRTR2 / RRTStarR2) and SE(2) with straight or Reeds-Shepp connections (rrt_star.h:181).
struct RRTParameters { // rrt.h:12-21
	unsigned int maxIteration = 100, maxNumberTreeNode = 10000;
	double maxConnectionDistance = 0.1, goalBias = 0.05;
};
struct RRTStarParameters { // rrt_star.h:12-21
	unsigned int maxIteration = 10000, maxNumberTreeNode = 10000;
	double maxConnectionDistance = 0.1, goalBias = 0.05;
};

class RRTBase {
public:
	/// free space: bounds only (StateValidatorFree); with a validator: its map (StateValidatorOccupancyMap, SE(2) only)
	RRTBase(const Ref<Device>& dev, const Pose2d& lb, const Pose2d& ub, int dim, bool star, bool reedsShepp, double minTurningRadius,
		const Ref<StateValidatorOccupancyMap>& validator) :
		m_dev(dev), m_lb(lb), m_ub(ub), m_dim(dim), m_star(star), m_rs(reedsShepp), m_rmin(minTurningRadius), m_validator(validator) { }
	void SetInitState(const Pose2d& s) { m_init = s; }
	void SetGoalState(const Pose2d& s) { m_goal = s; }
	uint64_t seed = 0; // the reference draws from an unseeded process-global engine (utils/random.h:14-16)
	Status SearchPath()
	{
		ppx_rrt_params p {};
		p.dim = m_dim;
		p.connection = m_rs ? 1 : 0;
		p.star = m_star ? 1 : 0;
		p.validator = m_validator ? 1 : 0;
		p.max_iteration = m_maxIteration;
		p.max_nodes = m_maxNodes;
		p.max_connection_distance = m_maxConnectionDistance;
		p.goal_bias = m_goalBias;
		p.min_turning_radius = m_rmin;
		p.lb[0] = m_lb.x(); p.lb[1] = m_lb.y(); p.lb[2] = m_lb.theta;
		p.ub[0] = m_ub.x(); p.ub[1] = m_ub.y(); p.ub[2] = m_ub.theta;
		ppx_validator_params vp { 1.0f, 0.1f };
		if (m_validator) {
			vp.min_safe_radius = m_validator->minSafeRadius;
			vp.min_path_interpolation_distance = m_validator->minPathInterpolationDistance;
		}
		const double s[3] = { m_init.x(), m_init.y(), m_init.theta }, g[3] = { m_goal.x(), m_goal.y(), m_goal.theta };
		const int64_t cap = (int64_t)m_maxNodes + 2;
		std::vector<double> nodes((size_t)cap * 3);
		std::vector<int32_t> parents((size_t)cap);
		m_dev->Check(ppx_rrt_run(m_dev->Context(), &p, m_validator ? &vp : nullptr, 1, s, g, &seed, &m_result, cap, nodes.data(), parents.data(),
			nullptr));
		m_path.clear();
		if (m_result.status == 0 && m_result.solution >= 0) { // GetPath, rrt.h:96-113: the chain root .. solution node
			for (int n = m_result.solution; n >= 0; n = parents[n])
				m_path.insert(m_path.begin(), Pose2d(nodes[3 * (size_t)n], nodes[3 * (size_t)n + 1], nodes[3 * (size_t)n + 2]));
		}
		return m_result.status == 0 ? Status::Success : Status::Failure;
	}
	std::vector<Pose2d> GetPath() const { return m_path; }
	const ppx_rrt_result& GetRecord() const { return m_result; }

protected:
	unsigned int m_maxIteration = 100, m_maxNodes = 10000;
	double m_maxConnectionDistance = 0.1, m_goalBias = 0.05;

private:
	Ref<Device> m_dev;
	Pose2d m_lb, m_ub, m_init, m_goal;
	int m_dim;
	bool m_star, m_rs;
	double m_rmin;
	Ref<StateValidatorOccupancyMap> m_validator;
	ppx_rrt_result m_result {};
	std::vector<Pose2d> m_path;
};

class RRTR2 : public RRTBase { // rrt.h:154
public:
	RRTR2(const Ref<Device>& dev, const Pose2d& lb, const Pose2d& ub) : RRTBase(dev, lb, ub, 2, false, false, 1.0, nullptr) { }
	void SetParameters(const RRTParameters& p)
	{
		m_maxIteration = p.maxIteration; m_maxNodes = p.maxNumberTreeNode;
		m_maxConnectionDistance = p.maxConnectionDistance; m_goalBias = p.goalBias;
	}
};

class RRTStarR2 : public RRTBase { // rrt_star.h:180
public:
	RRTStarR2(const Ref<Device>& dev, const Pose2d& lb, const Pose2d& ub) : RRTBase(dev, lb, ub, 2, true, false, 1.0, nullptr)
	{
		SetParameters(RRTStarParameters());
	}
	void SetParameters(const RRTStarParameters& p)
	{
		m_maxIteration = p.maxIteration; m_maxNodes = p.maxNumberTreeNode;
		m_maxConnectionDistance = p.maxConnectionDistance; m_goalBias = p.goalBias;
	}
};

class RRTStarSE2 : public RRTBase { // rrt_star.h:181 (commented out upstream: it does not compile there)
public:
	RRTStarSE2(const Ref<Device>& dev, const Pose2d& lb, const Pose2d& ub, const Ref<StateValidatorOccupancyMap>& validator,
		bool reedsShepp = true, double minTurningRadius = 1.0) :
		RRTBase(dev, lb, ub, 3, true, reedsShepp, minTurningRadius, validator)
	{
		SetParameters(RRTStarParameters());
	}
	void SetParameters(const RRTStarParameters& p)
	{
		m_maxIteration = p.maxIteration; m_maxNodes = p.maxNumberTreeNode;
		m_maxConnectionDistance = p.maxConnectionDistance; m_goalBias = p.goalBias;
	}
};

} // namespace Planner
