// C++ host-side tests of include/ppx_planner.hpp, written after the reference's own tests
// (planner/tests/test_hybrid_a_star.cpp:9-37, test_reeds_shepp.cpp:7-27) plus a validator case.
// Built by tests/cpp/Makefile against pathplanning_b200/libppx.so; needs a CUDA device to RUN
// (tests/test_cpp_host.py runs it under the gpu marker and only compiles/links it on CPU).
#include <cassert>
#include <cstdio>
#include <cstdlib>
#include "ppx_planner.hpp"

#define CHECK(cond)                                                              \
	do {                                                                         \
		if (!(cond)) {                                                           \
			std::fprintf(stderr, "%s:%d: CHECK failed: %s\n", __FILE__, __LINE__, #cond); \
			std::exit(1);                                                        \
		}                                                                        \
	} while (0)

namespace Planner {

// test_hybrid_a_star.cpp:9-37: empty 20 m x 20 m map at 0.1 m, default parameters
void TestHybridAStar(const Ref<Device>& dev)
{
	Pose2d lb(-10, -10, -M_PI), ub(10, 10, M_PI);
	auto map = std::make_shared<OccupancyMap>(dev, 0.1f, 20.0f, 20.0f, lb, ub);
	auto stateValidator = std::make_shared<StateValidatorOccupancyMap>(map);

	HybridAStar hybridAStar;
	CHECK(hybridAStar.Initialize(stateValidator));

	Pose2d start = { 0.0, 0.0, 0.0 };
	Pose2d goal = { 8.0, 8.0, 0.78 };
	hybridAStar.SetInitState(start);
	hybridAStar.SetGoalState(goal);

	CHECK(hybridAStar.SearchPath() == Status::Success);
	std::vector<Pose2d> path = hybridAStar.GetPath();

	CHECK(!path.empty());
	double spatialTolerance = 1e-1;
	double angularTolerance = 5 * M_PI / 180;
	CHECK(std::hypot(path.front().x() - start.x(), path.front().y() - start.y()) < spatialTolerance);
	CHECK(std::fabs(path.front().theta - start.theta) < angularTolerance);
	CHECK(std::hypot(path.back().x() - goal.x(), path.back().y() - goal.y()) < spatialTolerance);
	CHECK(std::fabs(path.back().theta - goal.theta) < angularTolerance);
	CHECK(hybridAStar.GetGraphSearchOptimalCost() >= std::hypot(8.0, 8.0) - 1e-9);
	// the pose array is sampled every pathInterpolation along the solution, cusp points snapped in (hybrid_a_star.cpp:260-291)
	for (size_t i = 1; i < path.size(); i++)
		CHECK(std::hypot(path[i].x() - path[i - 1].x(), path[i].y() - path[i - 1].y()) <= 0.15 + 1e-9);
	CHECK(!hybridAStar.GetCuspIndices().empty() && hybridAStar.GetCuspIndices().front() == 0);
	CHECK(hybridAStar.GetGraphSearchStates().size() >= 2 && hybridAStar.GetGraphSearchStates().size() < path.size());
	std::printf("hybrid A*: %zu states, cost %.4f, %d expansions\n", path.size(), hybridAStar.GetGraphSearchOptimalCost(),
		hybridAStar.GetStats().numExpandedNodes);

	// Smoother on that pose array (smoother.cpp:33-158) without the curvature term: the descent converges, the end
	// points stay, interior points move by centimetres, the result is collision free
	{
		Smoother smoother;
		CHECK(smoother.Initialize(stateValidator, 1.0f / 2.0f));
		auto sp = smoother.GetParameters();
		sp.curvatureWeight = 0.0f;
		smoother.SetParameters(sp);
		std::vector<int32_t> none(2 * (size_t)map->Rows() * map->Columns(), -1); // no obstacle, no Voronoi edge
		smoother.SetGVD(none, none);
		std::unordered_set<int> cusps(hybridAStar.GetCuspIndices().begin(), hybridAStar.GetCuspIndices().end());
		Smoother::Status st = smoother.Smooth(path, cusps);
		CHECK(st == Smoother::MaxIteration || st == Smoother::StepTolerance);
		const auto& sm = smoother.GetPath();
		CHECK(sm.size() == path.size());
		CHECK(sm.front().x() == path.front().x() && sm.back().y() == path.back().y());
		double moved = 0.0;
		for (size_t i = 0; i < sm.size(); i++)
			moved = std::max(moved, std::hypot(sm[i].x() - path[i].x(), sm[i].y() - path[i].y()));
		CHECK(moved > 1e-6 && moved < 1.0);
		CHECK(smoother.Smooth({ path[0], path[1], path[2] }) == Smoother::PathSize);
		std::printf("smoother: status %d, max displacement %.4f m\n", (int)st, moved);
	}

	// an obstacle outline rasterised on the device (ObstacleListOccupancyMap::AddObstacle): a 2 m x 1 m rectangle
	{
		auto m2 = std::make_shared<OccupancyMap>(dev, 0.1f, 20.0f, 20.0f, lb, ub);
		int id0 = m2->AddObstacle({ { 1.0, 0.5 }, { -1.0, 0.5 }, { -1.0, -0.5 }, { 1.0, -0.5 } });
		int id1 = m2->AddObstacle({ { 5.0, 5.0 }, { 6.0, 5.0 }, { 5.5, 6.0 } });
		CHECK(id0 == 0 && id1 == 1);
		int r, c;
		CHECK(m2->WorldPositionToGridCell(1.0, 0.0, r, c) && m2->GetOccupancyValue(r, c) == 0); // on the right side
		CHECK(m2->WorldPositionToGridCell(0.0, 0.0, r, c) && !m2->IsOccupied(r, c)); // outlines only: the inside stays free
		CHECK(m2->WorldPositionToGridCell(5.5, 5.0, r, c) && m2->GetOccupancyValue(r, c) == 1);
		m2->Update();
		CHECK(m2->WorldPositionToGridCell(0.0, 0.0, r, c) && std::fabs(m2->GetDistanceToNearestObstacle(r, c) - 0.5f) < 0.11f);
		m2->RemoveObstacle({ { 1.0, 0.5 }, { -1.0, 0.5 }, { -1.0, -0.5 }, { 1.0, -0.5 } });
		m2->Update();
		CHECK(m2->WorldPositionToGridCell(1.0, 0.0, r, c) && !m2->IsOccupied(r, c));
		CHECK(m2->WorldPositionToGridCell(0.0, 0.0, r, c) && m2->GetDistanceToNearestObstacle(r, c) > 5.0f); // only the triangle is left
	}

	// test_a_star.cpp:17-35 on the occupancy grid: AStarN2 from (0,0)... here (10,10) to (100,60) on the empty map
	{
		AStarN2 aStar;
		CHECK(aStar.Initialize(map));
		GridCellPosition s0 { 10, 10 }, g0 { 100, 60 };
		aStar.SetInitState(s0);
		aStar.SetGoalState(g0);
		CHECK(aStar.SearchPath() == Status::Success);
		auto cells = aStar.GetPath();
		CHECK(!cells.empty() && cells.front() == s0 && cells.back() == g0);
		// 50 diagonal + 40 straight steps
		CHECK(std::fabs(aStar.GetOptimalCost() - (50 * std::sqrt(2.0) + 40)) < 1e-9 && cells.size() == 91);
		std::printf("grid A*: %zu cells, cost %.4f, %d explored\n", cells.size(), aStar.GetOptimalCost(), aStar.GetNumExploredStates());
	}

	// a wall across the map with no gap: the goal becomes unreachable -> Status::Failure
	int r0, c0;
	for (double y = -10.0; y < 10.0; y += 0.05) {
		if (map->WorldPositionToGridCell(4.0, y, r0, c0))
			map->SetOccupied(r0, c0, 1);
	}
	CHECK(hybridAStar.Initialize(stateValidator)); // re-uploads the edited map
	CHECK(hybridAStar.SearchPath() == Status::Failure);
	CHECK(hybridAStar.GetPath().empty());

	// batch of queries on the walled map: left-side goals stay reachable
	std::vector<Status> st;
	std::vector<std::vector<Pose2d>> paths;
	std::vector<double> costs;
	hybridAStar.SearchPathBatch({ start, start, start }, { Pose2d(-6, 5, 1.0), goal, Pose2d(0, -7, -2.0) }, { 1, 2, 3 }, st, paths, costs);
	CHECK(st[0] == Status::Success && st[1] == Status::Failure && st[2] == Status::Success);
	CHECK(std::isinf(costs[1]) && costs[0] > 0 && costs[2] > 0);
}

// test_reeds_shepp.cpp:7-27: the constructed path leads to the final state
void TestReedsShepp(const Ref<Device>& dev, const Pose2d& start, const Pose2d& goal)
{
	PathConnectionReedsShepp connection(dev, 1.0);
	auto path = connection.Connect(start, goal);
	CHECK(path->GetWord() >= 0);
	double tol = 1e-6;
	Pose2d end = path->GetFinalState();
	CHECK(std::abs(end.x() - goal.x()) < tol);
	CHECK(std::abs(end.y() - goal.y()) < tol);
	CHECK(std::abs(end.theta - goal.theta) < tol * M_PI / 180.0);
	CHECK(path->GetLength() >= std::hypot(goal.x() - start.x(), goal.y() - start.y()) - 1e-9);
}

// StateValidatorOccupancyMap on a map with one obstacle: IsStateValid / IsPathValid / last ratio
void TestValidator(const Ref<Device>& dev)
{
	Pose2d lb(-10, -10, -M_PI), ub(10, 10, M_PI);
	auto map = std::make_shared<OccupancyMap>(dev, 0.1f, 20.0f, 20.0f, lb, ub);
	int r, c;
	CHECK(map->WorldPositionToGridCell(5.0, 0.0, r, c));
	map->SetOccupied(r, c);
	map->Update();
	CHECK(map->GetDistanceToNearestObstacle(r, c) == 0.0f);
	CHECK(std::fabs(map->GetDistanceToNearestObstacle(r - 30, c) - 3.0f) < 1e-5f);
	StateValidatorOccupancyMap validator(map);
	validator.minSafeRadius = 1.0f;
	CHECK(validator.IsStateValid(Pose2d(0, 0, 0)));
	CHECK(!validator.IsStateValid(Pose2d(4.5, 0.2, 0)));
	CHECK(!validator.IsStateValid(Pose2d(10.5, 0, 0))); // out of bounds
	PathConnectionReedsShepp connection(dev, 2.0);
	auto through = connection.Connect(Pose2d(0, 0, 0), Pose2d(9, 0, 0)); // straight through the obstacle
	float last = -1.0f;
	CHECK(!validator.IsPathValid(*through, &last));
	CHECK(last > 0.3f && last < 0.45f); // stops about 1 m before x = 5 on a 9 m path
	auto clear = connection.Connect(Pose2d(0, 3, 0), Pose2d(9, 3, 0));
	CHECK(validator.IsPathValid(*clear, &last) && last == 1.0f);
	// the obstacle heuristic: behind the obstacle the cost-to-go is at least the Euclidean distance
	ObstaclesHeuristic h(map, 1.0, 1.0);
	h.Update(Pose2d(9, 0, 0));
	CHECK(h.GetHeuristicValue(Pose2d(0, 0, 0)) >= 9.0 - 1e-9);
	CHECK(h.GetHeuristicValue(Pose2d(9, 0, 0)) == 0.0);
}

}

namespace Planner {

// planner/tests/test_rrt.cpp:9-38 and test_rrt_star.cpp:10-39: free space [0,5]^2, (0,0) -> (2,2); the reference asserts the
// end points within 1.0 (and can fail by chance on its unseeded RNG: here the seed is fixed); test_tree.cpp:91-130.
void TestRRTAndTree(const Ref<Device>& dev)
{
	Tree<2> tree(dev);
	int a = tree.CreateRootNode({ 0.0, 0.0 });
	int b = tree.Extend({ 1.0, 0.0 }, a);
	int c = tree.Extend({ 2.0, 0.0 }, b);
	int d = tree.Extend({ 3.0, 0.0 }, c);
	CHECK(tree.Size() == 4 && tree.Extend({ 2.0, 0.0 }, a) == c && tree.GetParent(c) == b);
	CHECK(tree.GetNearestNode({ 1.2, 0.0 }) == b);
	auto near = tree.GetNearestNodes({ 2.6, 0.0 }, 2);
	CHECK(near.size() == 2 && ((near[0] == d && near[1] == c) || (near[0] == c && near[1] == d)));
	for (int star = 0; star < 2; star++) {
		Pose2d lb(0, 0, 0), ub(5, 5, 0);
		Status st;
		std::vector<Pose2d> path;
		if (star) {
			RRTStarR2 rrt(dev, lb, ub);
			RRTStarParameters p;
			p.maxIteration = 20000;
			rrt.SetParameters(p);
			rrt.seed = 31;
			rrt.SetInitState(Pose2d(0, 0, 0));
			rrt.SetGoalState(Pose2d(2, 2, 0));
			st = rrt.SearchPath();
			path = rrt.GetPath();
		} else {
			RRTR2 rrt(dev, lb, ub);
			RRTParameters p;
			p.maxIteration = 20000;
			rrt.SetParameters(p);
			rrt.seed = 31;
			rrt.SetInitState(Pose2d(0, 0, 0));
			rrt.SetGoalState(Pose2d(2, 2, 0));
			st = rrt.SearchPath();
			path = rrt.GetPath();
		}
		CHECK(st == Status::Success && path.size() >= 2);
		CHECK(std::hypot(path.front().x(), path.front().y()) < 1.0);
		CHECK(std::hypot(path.back().x() - 2.0, path.back().y() - 2.0) < 1.0);
	}
}

}

int main()
{
	auto dev = std::make_shared<Planner::Device>(0);
	Planner::TestReedsShepp(dev, Planner::Pose2d(0, 0, 0), Planner::Pose2d(5, 5, M_PI / 2));
	Planner::TestReedsShepp(dev, Planner::Pose2d(1, -2, 0.3), Planner::Pose2d(-3, 0.5, -2.0));
	Planner::TestReedsShepp(dev, Planner::Pose2d(0, 0, 0), Planner::Pose2d(0.1, 1.5, M_PI));
	Planner::TestValidator(dev);
	Planner::TestHybridAStar(dev);
	Planner::TestRRTAndTree(dev);
	std::printf("test_planner: all checks passed\n");
	return 0;
}
