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
// Batch entry points (names ending in Batch) are the additions the device makes worthwhile; the
// scalar methods keep the reference's signatures and run a batch of one.
// There is NO CPU fallback: constructing a Device without a CUDA device throws.
#pragma once
#include <algorithm>
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

} // namespace Planner
