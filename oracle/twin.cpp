// TEST INFRASTRUCTURE ONLY -- never linked into, imported by, or executed from the product path.
//
// "Twin": the arithmetic / state-machine headers of libppx (pathplanning_b200/csrc/*.cuh) compiled
// by g++ for the host (PPX_HOST_TWIN), behind a few C entry points.  There is no GPU in the build
// container, so this is how the control logic of the kernels (micro-step state machines, frontier
// tie orders, the reference's parity traps) is debugged against the compiled reference
// (oracle/_ref/libpp_ref.so) and the golden fixtures before a kernel goes to the B200.
// It is NOT the parity oracle: parity is pinned by the unmodified reference (ref_harness.cpp) and
// by tests/golden; and it is NOT a CPU fallback: nothing under pathplanning_b200/ loads it.
// Note that host libm (glibc) and device libm (CUDA) differ in the last ulp, so twin == reference
// does not by itself prove kernel == reference; the -m gpu tests do that.
#define PPX_HOST_TWIN 1
#include <vector>
#include <cstdio>
#include "ppx_device.cuh"

using namespace ppx;

namespace {
struct TwinWorld {
	ppx_map_desc desc;
	std::vector<float> clearance;
	std::vector<int> occupancy;
	std::vector<float> pathcost;
	MapDev dev() const
	{
		MapDev m;
		m.rows = desc.rows;
		m.cols = desc.cols;
		m.res = desc.resolution;
		m.resd = (double)desc.resolution;
		m.ox = desc.world_origin[0];
		m.oy = desc.world_origin[1];
		m.lox = desc.local_origin[0];
		m.loy = desc.local_origin[1];
		m.lbx = desc.lb[0]; m.lby = desc.lb[1]; m.lbt = desc.lb[2];
		m.ubx = desc.ub[0]; m.uby = desc.ub[1]; m.ubt = desc.ub[2];
		m.clearance = clearance.data();
		m.occupancy = occupancy.data();
		m.pathcost = pathcost.data();
		return m;
	}
};

struct LenArray {
	const double* p;
	double get(int i) const { return p[i]; }
};
}

extern "C" {

void* twin_world_create(const ppx_map_desc* desc, const int32_t* sqdist, const int32_t* occupancy, const float* pathcost)
{
	auto* w = new TwinWorld();
	w->desc = *desc;
	size_t n = (size_t)desc->rows * desc->cols;
	w->clearance.resize(n);
	w->occupancy.assign(n, -1);
	w->pathcost.assign(n, 0.0f);
	for (size_t i = 0; i < n; i++)
		w->clearance[i] = (float)(sqrt((double)sqdist[i]) * (double)desc->resolution);
	if (occupancy)
		w->occupancy.assign(occupancy, occupancy + n);
	if (pathcost)
		w->pathcost.assign(pathcost, pathcost + n);
	return w;
}

void twin_world_destroy(void* h) { delete (TwinWorld*)h; }

// impl 0: thread-per-edge walker (rs_is_path_valid); impl 1: micro-step lane machine (RSLane)
void twin_check_rs(void* h, long n, const ppx_rs_edge* edges, double r, float radius, float minStep, int impl,
	uint8_t* valid, float* last, uint32_t* samples, uint32_t* microsteps)
{
	auto* w = (TwinWorld*)h;
	MapDev map = w->dev();
	for (long i = 0; i < n; i++) {
		Pose start = { edges[i].start[0], edges[i].start[1], edges[i].start[2] };
		if (impl == 0) {
			RSSeg seg;
			for (int j = 0; j < 5; j++)
				seg.len[j] = edges[i].len[j];
			seg.code = (uint32_t)edges[i].code;
			float l;
			unsigned ns;
			valid[i] = rs_is_path_valid(map, start, seg, r, radius, minStep, &l, &ns);
			last[i] = l;
			if (samples) samples[i] = ns;
			if (microsteps) microsteps[i] = 0;
		} else {
			LenArray L { edges[i].len };
			RSLane<LenArray> lane;
			lane.begin(L, start, (uint32_t)edges[i].code, r);
			unsigned steps = 0;
			while (true) {
				double arg = lane.pre(L, r);
				double s, c;
				sincos(arg, &s, &c);
				steps++;
				if (lane.post(map, s, c, r, radius, minStep))
					break;
			}
			valid[i] = lane.verdict;
			last[i] = lane.last;
			if (samples) samples[i] = lane.ns;
			if (microsteps) microsteps[i] = steps;
		}
	}
}

} // extern "C"
