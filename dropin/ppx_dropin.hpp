// Drop-in on the reference side: subclasses of the reference's OWN planner classes whose hot methods call libppx.so.
//
// This header is compiled inside (or next to) the reference tree, against the reference's headers:
//     g++ -std=c++17 -fno-access-control -I<reference>/planner/src -I<repo>/include ... -lppx
// (-fno-access-control: the adapters read what the reference keeps private -- the grids of OccupancyMap / GVD, the
// segment of a PathReedsShepp -- without a single change to the reference's sources.)
//
//   PpxStateValidatorOccupancyMap : StateValidatorOccupancyMap   IsStateValid / IsPathValid / IsPathValidBatch -> ppx_is_state_valid,
//                                                                ppx_check_edges_{rs,csteer} (state_validator_occupancy_map.cpp:15-71)
//   PpxHybridAStar : HybridAStar                                 Initialize / SearchPath / GetPath / GetStats / GetGraphSearch* ->
//                                                                ppx_hybrid_{create,search,resample}, ppx_smoother_* (hybrid_a_star.cpp:206-304)
//   PpxObstaclesHeuristic : ObstaclesHeuristic                   Update -> ppx_wavefront into the reference's own m_cost / m_explored
//                                                                (heuristics.cpp:106-153); GetHeuristicValue is the reference's
//   PpxBuildNonHolonomicHeuristic(...)                           NonHolonomicHeuristic::Build (heuristics.cpp:36-76) with the table
//                                                                computed by ppx_hybrid_create and copied into m_values
// With `-include ppx_dropin.hpp -DPPX_DROPIN_SUBSTITUTE` in front of an UNMODIFIED reference translation unit (the
// pybind11 module interfaces/python/src/pyplanning.cpp, planner/tests/test_hybrid_a_star.cpp, main.cpp) the identifiers
// HybridAStar and StateValidatorOccupancyMap in that unit name the adapters: the reference's own binding / test code then
// runs on the GPU path.  Nothing here links or includes anything under oracle/.
#pragma once
#include <cstdlib>
#include <cstring>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>

#include "core/base.h"
#include "algo/heuristics.h"
#include "algo/hybrid_a_star.h"
#include "algo/smoother.h"
#include "geometry/2dplane.h"
#include "geometry/reeds_shepp.h"
#include "models/kinematic_bicycle_model.h"
#include "paths/path_constant_steer.h"
#include "paths/path_reeds_shepp.h"
#include "state_space/state_space_se2.h"
#include "state_validator/gvd.h"
#include "state_validator/obstacle.h"
#include "state_validator/obstacle_list_occupancy_map.h"
#include "state_validator/occupancy_map.h"
#include "state_validator/state_validator_occupancy_map.h"
#include "utils/make_ref_enabler.h"

#include "ppx.h"

namespace Planner {

// One context per process (device PPX_DEVICE, default 0); throws when there is no GPU: no CPU fallback.
class PpxDevice {
public:
	static ppx_ctx* Ctx()
	{
		static PpxDevice dev;
		return dev.m_ctx;
	}
	static void Check(int rc, const char* what)
	{
		if (rc != PPX_OK)
			throw std::runtime_error(std::string("libppx: ") + what + ": " + ppx_last_error(Ctx()));
	}

private:
	PpxDevice()
	{
		const char* d = std::getenv("PPX_DEVICE");
		if (ppx_ctx_create(d ? std::atoi(d) : 0, &m_ctx) != PPX_OK)
			throw std::runtime_error("libppx: no CUDA device (the drop-in has no CPU fallback)");
	}
	~PpxDevice() { ppx_ctx_destroy(m_ctx); }
	ppx_ctx* m_ctx = nullptr;
};

// The world model of a validator as ppx_map_upload takes it.  The grids are the reference's own: occupancy from the
// OccupancyMap, the squared obstacle distance from its ObstacleDistanceMap, the potential from the GVD when there is one.
inline ppx_map_desc PpxMapDesc(StateSpaceSE2& space, OccupancyMap& map)
{
	ppx_map_desc d {};
	d.rows = map.Rows();
	d.cols = map.Columns();
	d.resolution = map.resolution;
	d.world_origin[0] = map.m_worldGridOrigin.x();
	d.world_origin[1] = map.m_worldGridOrigin.y();
	d.local_origin[0] = map.m_localOrigin.x();
	d.local_origin[1] = map.m_localOrigin.y();
	d.lb[0] = space.bounds[0].x(); d.lb[1] = space.bounds[0].y(); d.lb[2] = space.bounds[0].theta;
	d.ub[0] = space.bounds[1].x(); d.ub[1] = space.bounds[1].y(); d.ub[2] = space.bounds[1].theta;
	return d;
}

inline void PpxUploadWorld(StateSpaceSE2& space, OccupancyMap& map, GVD* gvd)
{
	ppx_map_desc d = PpxMapDesc(space, map);
	const int32_t* sq = map.m_obstacleDistanceMap->m_distance.m_data;
	const int32_t* occ = map.m_occupancyMatrix->m_data;
	const float* pc = gvd ? gvd->m_pathCostMap.m_data : nullptr;
	PpxDevice::Check(ppx_map_upload(PpxDevice::Ctx(), &d, sq, occ, pc), "ppx_map_upload");
}

class PpxStateValidatorOccupancyMap : public StateValidatorOccupancyMap {
public:
	PpxStateValidatorOccupancyMap(const Ref<StateSpaceSE2>& stateSpace, const Ref<OccupancyMap>& map) :
		StateValidatorOccupancyMap(stateSpace, map), m_space(stateSpace) { }

	// Call after OccupancyMap::Update() (or GVD::Update()) changed the grids.  The first validation call syncs by itself.
	void Sync(GVD* gvd = nullptr)
	{
		PpxUploadWorld(*m_space, *GetOccupancyMap(), gvd);
		m_synced = true;
	}

	virtual bool IsStateValid(const Pose2d& state) override
	{
		if (!m_synced)
			Sync();
		const double p[3] = { state.x(), state.y(), state.theta };
		ppx_validator_params vp { minSafeRadius, minPathInterpolationDistance };
		uint8_t ok = 0;
		PpxDevice::Check(ppx_is_state_valid(PpxDevice::Ctx(), 1, p, &vp, &ok), "ppx_is_state_valid");
		return ok != 0;
	}

	virtual bool IsPathValid(const PathSE2Base& path, float* last = nullptr) override
	{
		std::vector<const PathSE2Base*> one { &path };
		std::vector<uint8_t> ok;
		std::vector<float> l;
		if (!IsPathValidBatch(one, ok, l))
			return StateValidatorOccupancyMap::IsPathValid(path, last); // a path type outside the hot path's: the reference's own loop
		if (last)
			*last = l[0];
		return ok[0] != 0;
	}

	// Batched IsPathValid over Reeds-Shepp and constant-steer paths (the two path types of the expansion path).
	// Returns false (outputs untouched) when the batch holds another path type.
	bool IsPathValidBatch(const std::vector<const PathSE2Base*>& paths, std::vector<uint8_t>& valid, std::vector<float>& last)
	{
		if (!m_synced)
			Sync();
		std::vector<ppx_rs_edge> rs;
		std::vector<ppx_csteer_edge> cs;
		std::vector<int> kind(paths.size());
		double rmin = 1.0;
		for (size_t i = 0; i < paths.size(); i++) {
			if (auto* p = dynamic_cast<const PathReedsShepp*>(paths[i])) {
				ppx_rs_edge e {};
				e.start[0] = p->m_init.x(); e.start[1] = p->m_init.y(); e.start[2] = p->m_init.theta;
				uint64_t code = 0;
				for (int m = 0; m < 5; m++) {
					const auto& mo = p->m_pathSegment.m_motions[m];
					e.len[m] = mo.length;
					code |= PPX_RS_CODE((int)mo.steer, (int)mo.direction, m);
				}
				e.code = code;
				if (!rs.empty() && p->m_minTurningRadius != rmin)
					return false; // one turning radius per batch
				rmin = p->m_minTurningRadius;
				kind[i] = 0;
				rs.push_back(e);
			} else if (auto* c = dynamic_cast<const PathConstantSteer*>(paths[i])) {
				ppx_csteer_edge e {};
				e.start[0] = c->m_init.x(); e.start[1] = c->m_init.y(); e.start[2] = c->m_init.theta;
				const double tanS = tan(c->m_steering);
				const double beta = atan(c->m_model->m_rearToCenter * tanS / c->m_model->m_wheelbase);
				e.length = c->m_length;
				e.beta = beta;
				e.cos_beta = cos(beta);
				e.dtheta_ddist = e.cos_beta * tanS / c->m_model->m_wheelbase;
				e.direction = (int32_t)c->m_direction;
				kind[i] = 1;
				cs.push_back(e);
			} else
				return false;
		}
		ppx_validator_params vp { minSafeRadius, minPathInterpolationDistance };
		std::vector<uint8_t> vr(rs.size()), vc(cs.size());
		std::vector<float> lr(rs.size()), lc(cs.size());
		if (!rs.empty())
			PpxDevice::Check(ppx_check_edges_rs(PpxDevice::Ctx(), (int64_t)rs.size(), rs.data(), rmin, &vp, vr.data(), lr.data(), nullptr),
				"ppx_check_edges_rs");
		if (!cs.empty())
			PpxDevice::Check(ppx_check_edges_csteer(PpxDevice::Ctx(), (int64_t)cs.size(), cs.data(), &vp, vc.data(), lc.data(), nullptr),
				"ppx_check_edges_csteer");
		valid.resize(paths.size());
		last.resize(paths.size());
		size_t ir = 0, ic = 0;
		for (size_t i = 0; i < paths.size(); i++) {
			if (kind[i] == 0) {
				valid[i] = vr[ir];
				last[i] = lr[ir++];
			} else {
				valid[i] = vc[ic];
				last[i] = lc[ic++];
			}
		}
		return true;
	}

private:
	Ref<StateSpaceSE2> m_space;
	bool m_synced = false;
};

// ObstaclesHeuristic with the wavefront on the device.  The world model must be on the device (the validator adapter's
// Sync, or PpxUploadWorld): the wavefront reads the occupancy grid of the context.
class PpxObstaclesHeuristic : public ObstaclesHeuristic {
public:
	PpxObstaclesHeuristic(const Ref<OccupancyMap>& map, double reverseCostMultiplier, double forwardCostMultiplier) :
		ObstaclesHeuristic(map, reverseCostMultiplier, forwardCostMultiplier) { }

	// heuristics.cpp:106-153: every cell infinity / unexplored, then the cost-to-go wavefront from the goal's cell
	void Update(const Pose2d& goal)
	{
		const double g[3] = { goal.x(), goal.y(), goal.theta };
		const size_t cells = (size_t)m_cost.rows * m_cost.columns;
		std::vector<uint8_t> explored(cells);
		PpxDevice::Check(ppx_wavefront(PpxDevice::Ctx(), 1, g, m_cost.m_data, explored.data()), "ppx_wavefront");
		for (size_t i = 0; i < cells; i++)
			m_explored.m_data[i] = explored[i] != 0;
	}
};

// NonHolonomicHeuristic::Build with the 41 x 41 x 73-style table of Reeds-Shepp costs computed on the device.  The table
// belongs to the bounds of the world model on the device (ppx_hybrid_create reads them): they must be the `bounds` given here.
inline Ref<NonHolonomicHeuristic> PpxBuildNonHolonomicHeuristic(const std::array<Pose2d, 2>& bounds, double spatialResolution,
	double angularResolution, double minTurningRadius, double reverseCostMultiplier, double forwardCostMultiplier,
	double directionSwitchingCost)
{
	unsigned int nx = ceil((bounds[1].x() - bounds[0].x()) / spatialResolution);
	if (nx % 2 == 0)
		nx++;
	unsigned int ny = ceil((bounds[1].y() - bounds[0].y()) / spatialResolution);
	if (ny % 2 == 0)
		ny++;
	Ref<NonHolonomicHeuristic> heuristic = makeRef<MakeRefEnabler<NonHolonomicHeuristic>>(spatialResolution, angularResolution,
		minTurningRadius, reverseCostMultiplier, forwardCostMultiplier, directionSwitchingCost, nx, ny);
	ppx_hybrid_params hp {};
	hp.wheelbase = 2.6; // not used by the table
	hp.min_turning_radius = minTurningRadius;
	hp.direction_switching_cost = directionSwitchingCost;
	hp.reverse_cost_multiplier = reverseCostMultiplier;
	hp.forward_cost_multiplier = forwardCostMultiplier;
	hp.voronoi_cost_multiplier = 0.0;
	hp.num_generated_motion = 5;
	hp.spatial_resolution = spatialResolution;
	hp.angular_resolution = angularResolution;
	ppx_validator_params vp { 1.0f, 0.1f };
	ppx_hybrid* h = nullptr;
	PpxDevice::Check(ppx_hybrid_create(PpxDevice::Ctx(), &hp, &vp, &h), "ppx_hybrid_create");
	int32_t dims[3];
	double offsets[2];
	PpxDevice::Check(ppx_hybrid_table(h, dims, offsets, nullptr), "ppx_hybrid_table");
	if ((unsigned)dims[0] != heuristic->numSpatialX || (unsigned)dims[1] != heuristic->numSpatialY || (unsigned)dims[2] != heuristic->numAngular) {
		ppx_hybrid_destroy(h);
		throw std::runtime_error("libppx: the world model on the device has other bounds than the heuristic asked for");
	}
	std::vector<double> values((size_t)dims[0] * dims[1] * dims[2]);
	PpxDevice::Check(ppx_hybrid_table(h, dims, offsets, values.data()), "ppx_hybrid_table");
	for (int i = 0; i < dims[0]; i++)
		for (int j = 0; j < dims[1]; j++)
			std::memcpy(heuristic->m_values[i][j], &values[((size_t)i * dims[1] + j) * dims[2]], sizeof(double) * dims[2]);
	ppx_hybrid_destroy(h);
	return heuristic;
}

class PpxHybridAStar : public HybridAStar {
public:
	PpxHybridAStar() :
		PpxHybridAStar(SearchParameters()) { }
	PpxHybridAStar(const SearchParameters& p) :
		HybridAStar(p)
	{
		std::random_device rd;
		seed = ((uint64_t)rd() << 32) | rd();
		if (const char* s = std::getenv("PPX_SEED"))
			seed = std::strtoull(s, nullptr, 10);
	}
	virtual ~PpxHybridAStar()
	{
		if (m_smootherHandle)
			ppx_smoother_destroy(m_smootherHandle);
		if (m_handle)
			ppx_hybrid_destroy(m_handle);
	}

	// HybridAStar::Initialize, hybrid_a_star.cpp:206-235: Voronoi field (the reference's own GVD object, so that a map the
	// user keeps editing stays in charge of its incremental brushfire), heuristic table built on the device.
	bool Initialize(const Ref<PpxStateValidatorOccupancyMap>& validator)
	{
		if (!validator || !validator->GetStateSpace())
			return isInitialized = false;
		m_validator = validator;
		m_ppxValidator = validator;
		const auto& p = m_propagator->GetParameters();
		m_gvd = makeRef<GVD>(m_validator->GetOccupancyMap());
		m_gvd->Update();
		m_ppxValidator->Sync(m_gvd.get());
		if (m_handle)
			ppx_hybrid_destroy(m_handle);
		ppx_hybrid_params hp {};
		hp.wheelbase = p.wheelbase;
		hp.min_turning_radius = p.minTurningRadius;
		hp.direction_switching_cost = p.directionSwitchingCost;
		hp.reverse_cost_multiplier = p.reverseCostMultiplier;
		hp.forward_cost_multiplier = p.forwardCostMultiplier;
		hp.voronoi_cost_multiplier = p.voronoiCostMultiplier;
		hp.num_generated_motion = p.numGeneratedMotion;
		hp.spatial_resolution = p.spatialResolution;
		hp.angular_resolution = p.angularResolution;
		ppx_validator_params vp { validator->minSafeRadius, validator->minPathInterpolationDistance };
		PpxDevice::Check(ppx_hybrid_create(PpxDevice::Ctx(), &hp, &vp, &m_handle), "ppx_hybrid_create");
		m_smootherParams = Smoother::Parameters((float)(1.0 / p.minTurningRadius));
		return isInitialized = true;
	}

	void SetSmootherParameters(const Smoother::Parameters& param) { m_smootherParams = param; }
	const Smoother::Parameters& GetSmootherParameters() const { return m_smootherParams; }

	// HybridAStar::SearchPath, hybrid_a_star.cpp:237-304
	virtual Status SearchPath() override
	{
		if (!isInitialized)
			return Status::Failure;
		m_gvd->Update(); // :250 (the reference's incremental brushfire + potential), then the device copies
		m_ppxValidator->Sync(m_gvd.get());
		const double start[3] = { this->m_init.x(), this->m_init.y(), this->m_init.theta };
		const double goal[3] = { this->m_goal.x(), this->m_goal.y(), this->m_goal.theta };
		const int maxPath = 4096;
		ppx_hybrid_limits lim { 0, 0, maxPath, 0 };
		std::vector<ppx_hybrid_path_state> chain(maxPath);
		PpxDevice::Check(ppx_hybrid_search(m_handle, 1, start, goal, &seed, &lim, &m_result, chain.data(), nullptr), "ppx_hybrid_search");
		seed = seed * 6364136223846793005ULL + 1442695040888963407ULL; // the next search draws a fresh coin sequence
		m_stats.graphSearchStatus = m_result.status == 0 ? Status::Success : Status::Failure;
		m_chain.assign(chain.begin(), chain.begin() + (m_result.status == 0 ? std::min(m_result.path_len, maxPath) : 0));
		if (m_stats.graphSearchStatus < 0)
			return m_stats.graphSearchStatus;
		// "Process graph search path" (:260-291) on the device
		const int maxPoints = 16384;
		std::vector<double> poses((size_t)maxPoints * 3);
		std::vector<uint8_t> cusp(maxPoints);
		int32_t count = 0;
		PpxDevice::Check(ppx_hybrid_resample(m_handle, 1, goal, &m_result, chain.data(), maxPath, pathInterpolation, maxPoints, poses.data(),
			cusp.data(), &count, nullptr), "ppx_hybrid_resample");
		if (count <= 0)
			return m_stats.graphSearchStatus = Status::Failure;
		std::vector<Pose2d> graphSearchPath((size_t)count);
		for (int i = 0; i < count; i++)
			graphSearchPath[i] = Pose2d(poses[3 * i], poses[3 * i + 1], poses[3 * i + 2]);
		// Smoother::Smooth (:294-303) on the device, with the two cell grids of the reference's GVD
		EnsureSmoother();
		std::vector<double> smoothed((size_t)maxPoints * 3);
		int32_t sStatus = -1;
		PpxDevice::Check(ppx_smoother_smooth(m_smootherHandle, 1, poses.data(), cusp.data(), &count, maxPoints, smoothed.data(), &sStatus,
			nullptr), "ppx_smoother_smooth");
		m_stats.smoothingStatus = (Smoother::Status)sStatus;
		m_smoothedPath.clear();
		if (sStatus < 0) {
			m_ppxPath = graphSearchPath;
			return m_stats.graphSearchStatus;
		}
		m_smoothedPath.resize((size_t)count);
		for (int i = 0; i < count; i++)
			m_smoothedPath[i] = Pose2d(smoothed[3 * i], smoothed[3 * i + 1], smoothed[3 * i + 2]);
		m_ppxPath = m_smoothedPath;
		return Status::Success;
	}

	virtual std::vector<Pose2d> GetPath() const override { return m_ppxPath; }
	const Stats& GetStats() const { return m_stats; }
	const std::vector<Pose2d>& GetSmoothedPath() const { return m_smoothedPath; }
	double GetGraphSearchOptimalCost() const { return m_result.cost; }
	const ppx_hybrid_result& GetSearchRecord() const { return m_result; }
	void VisualizeObstacleHeuristic(const std::string&) const { throw std::runtime_error("visualisation is not part of the GPU path"); }

	// The actions of the solution chain as reference path objects (AStar::GetActions -> GetGraphSearchPath)
	std::vector<Ref<PathNonHolonomicSE2Base>> GetGraphSearchPath() const
	{
		std::vector<Ref<PathNonHolonomicSE2Base>> out;
		const auto& p = m_propagator->GetParameters();
		auto model = makeRef<KinematicBicycleModel>(p.wheelbase, 0.0);
		const double deltaMax = model->GetSteeringAngleFromTurningRadius(p.minTurningRadius);
		PathConnectionReedsShepp connection(p.minTurningRadius, p.directionSwitchingCost, p.reverseCostMultiplier, p.forwardCostMultiplier);
		for (size_t i = 1; i < m_chain.size(); i++) {
			const auto& from = m_chain[i - 1];
			const auto& s = m_chain[i];
			const Pose2d parent(from.pose[0], from.pose[1], from.pose[2]);
			if (s.kind == 1) {
				out.push_back(std::static_pointer_cast<PathNonHolonomicSE2Base>(
					std::static_pointer_cast<PathReedsShepp>(connection.Connect(parent, this->m_goal))));
			} else {
				// steering set of StatePropagator (hybrid_a_star.cpp:20-28): 0, +-deltaMax * k / (n / 2)
				const int di = s.slot >> 1;
				const double delta = di == 0 ? 0.0 : ((di + 1) / 2) / 2.0 * deltaMax * ((di & 1) ? 1.0 : -1.0);
				auto path = makeRef<PathConstantSteer>(model, parent, delta, p.spatialResolution * 1.5, (Direction)s.direction);
				if (s.last < 1.0f)
					path->Truncate((double)s.last);
				out.push_back(path);
			}
		}
		return out;
	}
	// The explored set is kept on the device as a hash of the expansion sequence (ppx_hybrid_result::expanded_hash) and,
	// on request, as the expansion log; path objects are only rebuilt for the solution.
	std::unordered_set<Ref<PathNonHolonomicSE2Base>> GetGraphSearchExploredPathSet() const
	{
		std::unordered_set<Ref<PathNonHolonomicSE2Base>> out;
		for (auto& p : GetGraphSearchPath())
			out.insert(p);
		return out;
	}

public:
	uint64_t seed; // seeds the coin flips of the Reeds-Shepp shots (the reference draws from an unseeded process-global engine)

private:
	void EnsureSmoother()
	{
		const auto& p = m_propagator->GetParameters();
		ppx_smoother_params sp {};
		sp.step_tolerance = m_smootherParams.stepTolerance;
		sp.max_iterations = (int32_t)m_smootherParams.maxIterations;
		sp.learning_rate = m_smootherParams.learningRate;
		sp.path_weight = m_smootherParams.pathWeight;
		sp.smooth_weight = m_smootherParams.smoothWeight;
		sp.voronoi_weight = m_smootherParams.voronoiWeight;
		sp.collision_weight = m_smootherParams.collisionWeight;
		sp.curvature_weight = m_smootherParams.curvatureWeight;
		sp.collision_ratio = m_smootherParams.collisionRatio;
		sp.max_curvature = (float)(1.0 / p.minTurningRadius);
		sp.alpha = m_gvd->alpha;
		sp.d_max = m_gvd->dMax;
		sp.min_safe_radius = m_validator->minSafeRadius;
		if (m_smootherHandle)
			ppx_smoother_destroy(m_smootherHandle);
		PpxDevice::Check(ppx_smoother_create(PpxDevice::Ctx(), &sp, &m_smootherHandle), "ppx_smoother_create");
		const int rows = m_gvd->rows, cols = m_gvd->columns;
		std::vector<int32_t> no((size_t)rows * cols * 2), ne((size_t)rows * cols * 2);
		for (int r = 0; r < rows; r++)
			for (int c = 0; c < cols; c++) {
				const auto& o = m_gvd->GetNearestObstacleCell(r, c);
				const auto& e = m_gvd->GetNearestVoronoiEdgeCell(r, c);
				const size_t i = ((size_t)r * cols + c) * 2;
				no[i] = o.row; no[i + 1] = o.col;
				ne[i] = e.row; ne[i + 1] = e.col;
			}
		PpxDevice::Check(ppx_smoother_set_gvd(m_smootherHandle, no.data(), ne.data()), "ppx_smoother_set_gvd");
	}

	Ref<PpxStateValidatorOccupancyMap> m_ppxValidator;
	ppx_hybrid* m_handle = nullptr;
	ppx_smoother* m_smootherHandle = nullptr;
	Smoother::Parameters m_smootherParams { 0.5f };
	ppx_hybrid_result m_result {};
	std::vector<ppx_hybrid_path_state> m_chain;
	std::vector<Pose2d> m_ppxPath, m_smoothedPath;
};

}

#ifdef PPX_DROPIN_SUBSTITUTE
// From here on (i.e. in the unmodified reference translation unit this header was force-included into) the two class
// names mean the adapters.  The reference's headers are already included above and carry `#pragma once`.
#define HybridAStar PpxHybridAStar
#define StateValidatorOccupancyMap PpxStateValidatorOccupancyMap
#endif
