// The adapter validator against the reference's own StateValidatorOccupancyMap on the same map: states, Reeds-Shepp
// connections and constant-steer primitives, single calls and one batch.  Exits non-zero on the first difference.
#include "ppx_dropin.hpp"
#include "models/kinematic_bicycle_model.h"
#include <cstdio>
#include <cstring>
#include <random>

using namespace Planner;

int main()
{
	std::array<Pose2d, 2> bounds = { Pose2d(-20, -20, -M_PI), Pose2d(20, 20, M_PI) };
	auto space = makeRef<StateSpaceSE2>(bounds);
	auto map = makeRef<ObstacleListOccupancyMap>(0.1f);
	auto gpu = makeRef<PpxStateValidatorOccupancyMap>(space, map);
	StateValidatorOccupancyMap cpu(space, map); // same map object: the reference's loop on the host
	std::mt19937_64 rng(7);
	std::uniform_real_distribution<double> U(-18.0, 18.0), A(-M_PI, M_PI);
	for (int i = 0; i < 40; i++) {
		auto obstacle = makeRef<Obstacle>();
		obstacle->SetShape(makeRef<RectangleShape>(4.5, 2.0));
		obstacle->SetPose(Pose2d(U(rng), U(rng), A(rng)));
		map->AddObstacle(obstacle);
	}
	map->Update();
	gpu->Sync();
	int bad = 0, nvalid = 0;
	for (int i = 0; i < 300; i++) {
		Pose2d s(U(rng), U(rng), A(rng));
		bool a = cpu.IsStateValid(s), b = gpu->IsStateValid(s);
		bad += a != b;
	}
	PathConnectionReedsShepp connect(2.0);
	auto model = makeRef<KinematicBicycleModel>(2.6, 0.0);
	std::vector<Ref<PathSE2Base>> keep;
	std::vector<const PathSE2Base*> batch;
	std::vector<uint8_t> want;
	std::vector<float> wantLast;
	for (int i = 0; i < 600; i++) {
		Pose2d from(U(rng), U(rng), A(rng));
		Ref<PathSE2Base> path;
		if (i % 3 == 2)
			path = makeRef<PathConstantSteer>(model, from, 0.3 * (i % 5 - 2), 1.5, i % 2 ? Direction::Forward : Direction::Backward);
		else {
			double d = 2.0 + 10.0 * (i % 7) / 7.0, a = A(rng);
			path = connect.Connect(from, Pose2d(from.x() + d * cos(a), from.y() + d * sin(a), A(rng)));
		}
		float la = -1.0f, lb = -1.0f;
		bool a = cpu.IsPathValid(*path, &la), b = gpu->IsPathValid(*path, &lb);
		if (a != b || std::fabs(la - lb) > 1e-6f * std::fmax(1.0f, std::fabs(la))) {
			if (bad < 5)
				std::printf("path %d: reference %d %.9g, adapter %d %.9g\n", i, (int)a, la, (int)b, lb);
			bad++;
		}
		nvalid += a;
		keep.push_back(path);
		batch.push_back(path.get());
		want.push_back(a);
		wantLast.push_back(la);
	}
	std::vector<uint8_t> got;
	std::vector<float> gotLast;
	if (!gpu->IsPathValidBatch(batch, got, gotLast))
		bad++;
	for (size_t i = 0; i < batch.size() && i < got.size(); i++)
		bad += (got[i] != 0) != (want[i] != 0) || std::fabs(gotLast[i] - wantLast[i]) > 1e-6f * std::fmax(1.0f, std::fabs(wantLast[i]));
	std::printf("drop-in validator: 300 states, 600 paths (%d valid), single + batch: %d differences\n", nvalid, bad);
	// heuristic adapters against the reference's own classes on the same world
	int badH = 0;
	{
		ObstaclesHeuristic refObs(map, 1.0, 1.0);
		PpxObstaclesHeuristic gpuObs(map, 1.0, 1.0);
		for (int t = 0; t < 3; t++) {
			Pose2d goal(U(rng), U(rng), A(rng));
			refObs.Update(goal);
			gpuObs.Update(goal);
			for (int r = 0; r < map->Rows(); r++)
				for (int c = 0; c < map->Columns(); c++)
					badH += std::memcmp(&refObs.m_cost[r][c], &gpuObs.m_cost[r][c], sizeof(float)) != 0
						|| refObs.m_explored[r][c] != gpuObs.m_explored[r][c];
		}
		auto refNh = NonHolonomicHeuristic::Build(bounds, 1.0, 2 * M_PI / 72, 2.0, 1.0, 1.0, 0.0);
		auto gpuNh = PpxBuildNonHolonomicHeuristic(bounds, 1.0, 2 * M_PI / 72, 2.0, 1.0, 1.0, 0.0);
		double worst = 0.0;
		for (unsigned i = 0; i < refNh->numSpatialX; i++)
			for (unsigned j = 0; j < refNh->numSpatialY; j++)
				for (unsigned k = 0; k < refNh->numAngular; k++) {
					const double a = refNh->m_values[i][j][k], b = gpuNh->m_values[i][j][k];
					worst = std::fmax(worst, std::fabs(a - b) / std::fmax(1.0, std::fabs(a)));
				}
		badH += worst > 1e-6;
		std::printf("drop-in heuristics: 3 wavefronts (%d x %d cells, float bits), table %u x %u x %u (worst relative difference %.3g): %d differences\n",
			map->Rows(), map->Columns(), refNh->numSpatialX, refNh->numSpatialY, refNh->numAngular, worst, badH);
	}
	return bad || badH ? 1 : 0;
}
