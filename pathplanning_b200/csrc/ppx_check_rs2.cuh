// k_check_rs2: batched IsPathValid on Reeds-Shepp edges (state_validator_occupancy_map.cpp:28-71), second form of the
// lane machine of ppx_fastlane.cuh (included by ppx_edges.cu inside its anonymous namespace).
//
// What changed against k_check_rs (round 1: 78 warp-instructions per edge at 13.4 of 32 active lanes):
//   * PREPROCESS, all lanes active.  A warp takes 32 edges at a time from the global cursor, lane L turns edge L into a
//     100-byte FP32 record in the warp's shared-memory queue: the poses of the <= 5 knots (start of every motion: the
//     reference applies completed motions with ratio exactly 1.0, so they do not depend on the sample,
//     path_reeds_shepp.cpp:12-47) in cell units, the cumulative lengths, the path length as double.  The 32 raw 72-byte
//     records arrive in shared memory by ONE bulk async copy per grab (cp.async.bulk + mbarrier, issued one grab
//     ahead), not by per-lane strided loads.
//   * EVERY ITERATION IS A SAMPLE.  A lane finds the motion of its sample by cumulative length and evaluates the chord
//     form from that knot: no knot-commit iterations (one micro-step in three before).
//   * REFILL PER ITERATION from the queue: an idle lane picks up its next edge by reading the record (a few shared
//     loads), so lanes do not wait for a batch of idle lanes, and the lane state in registers is four scalars.
// The certification rule (kCellEps margin, interior cells, step not limited by the border), the exact FP64 path for
// everything else and therefore the outputs are those of k_check_rs: filter on/off A/B, the device audit and the live
// reference tests are unchanged.
#pragma once

constexpr int kQueueCap = 64; // record slots per warp: <= 32 in use by the lanes + < 32 queued + 32 being added
constexpr int kRecWords = 25; // 100 bytes, odd word stride: conflict-free per-lane access
constexpr int kGrab2 = 32;

struct Rs2Smem { // per warp
	uint32_t q[kQueueCap][kRecWords];
	unsigned char order[kQueueCap]; // FIFO of the queued slots (ring)
	alignas(16) unsigned char stage[kGrab2 * sizeof(ppx_rs_edge)]; // bulk-copy destination
	alignas(8) unsigned long long bar; // mbarrier of the bulk copy
};

// record layout (32-bit words)
//  0..1   pathLength (double, lo / hi)
//  2+4m   knot m: kx, ky (cells, relative to rowb / colb), heading (rad, wrapped), cumulative length at its start (m)
//  22     code (20 bits) | number of motions << 20 | exact-only << 24
//  23     rowb (int16) | colb (int16) << 16
//  24     edge index relative to the launch

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(void* dst, const void* src, uint32_t bytes, unsigned long long* bar)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
		"l"(src), "r"(bytes), "r"(smem_u32(bar))
		: "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity)
{
	asm volatile(
		"{\n"
		".reg .pred p;\n"
		"WAIT_%=:\n"
		"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
		"@p bra DONE_%=;\n"
		"bra WAIT_%=;\n"
		"DONE_%=:\n"
		"}\n" ::"r"(smem_u32(bar)),
		"r"(parity)
		: "memory");
}

// Lane L: raw edge -> record.  The FP32 arithmetic is RSFastLane::begin / advance (full commits) unrolled over the motions.
__device__ __forceinline__ void rs2_preprocess(const double* __restrict__ raw /* 9 doubles */, uint32_t idx, const FastMap& fm,
	const MapDev& map, double r, float inv_r, float rc, uint32_t* __restrict__ rec)
{
	const double sx = raw[0], sy = raw[1], sth = raw[2];
	const uint32_t code = (uint32_t)__double_as_longlong(raw[8]) & 0xfffffu;
	double total = 0.0; // PathSegment::m_length as AddMotion accumulates it
	bool odd = false;
	int nm = 0;
	for (int i = 0; i < 5; i++) {
		const int dir = (code >> (4 * i + 2)) & 3;
		const double l = raw[3 + i];
		if (dir == 2 || l == PPX_INF)
			break;
		odd |= !(l >= 0.0); // negative or NaN lengths: exact path only
		total += fabs(l);
		nm++;
	}
	const double pathLength = total * r;
	const double ux = (sx - map.ox) * fm.inv_resd, uy = (sy - map.oy) * fm.inv_resd;
	odd |= !(fabs(ux) < 32000.0) || !(fabs(uy) < 32000.0) || !(fabs(sth) < 1e4);
	odd |= !(pathLength > 0.0) || !(pathLength * fm.inv_resd < (double)kFastMaxCells) || !(r > 1e-3) || !fm.enabled;
	rec[0] = (uint32_t)__double2loint(pathLength);
	rec[1] = (uint32_t)__double2hiint(pathLength);
	rec[22] = code | ((uint32_t)nm << 20) | (odd ? 1u << 24 : 0u);
	rec[24] = idx;
	if (odd) {
		rec[23] = 0u;
		return;
	}
	const double fx = floor(ux), fy = floor(uy);
	rec[23] = ((uint32_t)(int)fx & 0xffffu) | ((uint32_t)(int)fy << 16);
	float kx = (float)(ux - fx), ky = (float)(uy - fy);
	float thK = (float)(sth - PPX_2PI * rint(sth * (1.0 / PPX_2PI)));
	double lenAcc = 0.0;
	#pragma unroll
	for (int m = 0; m < 5; m++) {
		rec[2 + 4 * m] = __float_as_uint(kx);
		rec[3 + 4 * m] = __float_as_uint(ky);
		rec[4 + 4 * m] = __float_as_uint(thK);
		rec[5 + 4 * m] = __float_as_uint((float)lenAcc);
		if (m >= nm)
			continue;
		// the whole motion m applied to knot m (RSFastLane::advance with `full`)
		const int st = (code >> (4 * m)) & 3, dr = (code >> (4 * m + 2)) & 3;
		const double lm = raw[3 + m] * r; // motionLength, path_reeds_shepp.cpp:25
		const float ml = (float)lm;
		const bool straight = st == 1;
		const float sgnAdv = dr == 1 ? -1.0f : 1.0f, sgnTurn = st == 2 ? -1.0f : 1.0f;
		const float adv = sgnAdv * ml;
		const float a = adv * inv_r * sgnTurn;
		const float phi = 0.5f * a;
		const float ang = fast_wrap(straight ? thK : thK + phi);
		const float chord = (2.0f * rc) * PPX_FSIN(sgnTurn * phi);
		const float Lc = straight ? adv * fm.inv_res : chord;
		kx = kx + Lc * PPX_FCOS(ang);
		ky = ky + Lc * PPX_FSIN(ang);
		if (!straight)
			thK = fast_wrap(thK + a);
		lenAcc += lm;
	}
}

// State of IsPathValid's loop that the exact path reads and writes (by value: stays in registers across the call).
struct ExactIO2 {
	double pathLength, length, lastValidLength;
	unsigned ns;
	float last;
	bool verdict, done;
};

__device__ __noinline__ ExactIO2 exact_sample_lane2(ExactIO2 io, MapDev map, const ppx_rs_edge* __restrict__ e, double r,
	float minSafeRadius, float minStep)
{
	Pose start = { e->start[0], e->start[1], e->start[2] };
	RSSeg seg;
	#pragma unroll
	for (int i = 0; i < 5; i++)
		seg.len[i] = e->len[i];
	seg.code = (uint32_t)e->code;
	double ex, ey;
	io.done = rs_exact_sample(io, map, start, seg, r, minSafeRadius, minStep, &ex, &ey);
	return io;
}

// audit only: exact position of the sample the lane has just certified, compared with (px, py)
__device__ __noinline__ void audit_sample2(float px, float py, int rowb, int colb, double pathLength, const MapDev& map,
	const ppx_rs_edge* __restrict__ e, double r, double lengthAtSample, AuditOut* out)
{
	Pose start = { e->start[0], e->start[1], e->start[2] };
	RSSeg seg;
	for (int i = 0; i < 5; i++)
		seg.len[i] = e->len[i];
	seg.code = (uint32_t)e->code;
	RSWalker w;
	w.init(start);
	Pose p = w.at(seg, r, (lengthAtSample / pathLength) * pathLength);
	int row = trunc_i32((p.x - map.ox) / map.resd), col = trunc_i32((p.y - map.oy) / map.resd);
	int frow = rowb + (int)floorf(px), fcol = colb + (int)floorf(py);
	double dx = (p.x - map.ox) / map.resd - (double)rowb - (double)px;
	double dy = (p.y - map.oy) / map.resd - (double)colb - (double)py;
	float dev = (float)fmax(fabs(dx), fabs(dy));
	atomicMax(&out->max_dev_bits, __float_as_uint(dev));
	atomicAdd(&out->certified, 1ull);
	if (row != frow || col != fcol)
		atomicAdd(&out->cell_mismatch, 1ull);
}

constexpr int kCheck2Ctas = 6;

// kSolve: PathConnectionReedsShepp::Connect + IsPathValid (path_reeds_shepp.cpp:174-179, what GetReedsSheppChild and the
// RRT connections do): the grab is 32 (start, goal) pairs, staged by two bulk copies, solved eight pairs at a time by
// rs_solve_quad; lane L then owns pair L's segment, writes it to `solved` (the exact path re-reads it from there, and it is
// the `edges` output of ppx_rs_solve) and goes on exactly like the pre-solved case.
struct SolveArgs {
	const double* starts;
	const double* goals;
	ppx_rs_edge* solved; // [n] scratch / output
	int32_t* word; // optional outputs
	double* length;
	float* cost;
	float rev, fwd, sw;
	int shortest;
};

template <bool kAudit, bool kSolve, int kCtas = (kSolve ? 3 : kCheck2Ctas)>
__global__ void __launch_bounds__(kThreads, kCtas) k_check_rs2(MapDev map, FastMap fm, int64_t n,
	const ppx_rs_edge* __restrict__ edgesIn, double r, float minSafeRadius, float minStep, uint8_t* __restrict__ valid,
	float* __restrict__ last, uint32_t* __restrict__ samples, unsigned long long* __restrict__ queue, AuditOut* audit,
	unsigned long long* nextQueue, SolveArgs sa)
{
	const ppx_rs_edge* edges = kSolve ? sa.solved : edgesIn;
	__shared__ Rs2Smem s_warp[kThreads / 32];
	if (nextQueue) { // overlapped launches ("rs_overlap"): zero the successor's cursor, then let the successor in
		if (blockIdx.x == 0 && threadIdx.x == 0) {
			atomicExch(nextQueue, 0ull);
			__threadfence();
		}
		__syncthreads();
		asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
	}
	const unsigned lane_id = threadIdx.x & 31;
	const unsigned lt_mask = (1u << lane_id) - 1;
	Rs2Smem& S = s_warp[threadIdx.x >> 5];
	const float inv_r = (float)(1.0 / r);
	const float rc = (float)(r * fm.inv_resd);
	// the bulk copy needs 16-byte aligned sources
	const bool aligned = kSolve ? (((size_t)sa.starts | (size_t)sa.goals) & 15) == 0 : ((size_t)edges & 15) == 0;
	if (lane_id == 0)
		mbar_init(&S.bar, 1);
	__syncwarp();

	// lane state
	int state = LANE_IDLE;
	int slot = 0, m = 0;
	double length = 0.0, lastValidLength = 0.0;
	unsigned ns = 0;
	// warp-uniform state
	int qhead = 0, qcount = 0; // ring of preprocessed records
	int64_t grabBase = 0; // the grab whose raw records are (being) staged
	int grabCnt = 0;
	bool grabBulk = false, exhausted = false;
	uint32_t parity = 0;

	while (true) {
		unsigned idle = __ballot_sync(0xffffffffu, state == LANE_IDLE);
		// ---- A. not enough queued records for the idle lanes: preprocess the staged grab (all lanes), stage the next one.
		// Slots: a lane keeps reading its record while it samples, so new records go to slots that are neither in use nor
		// queued; in use + queued < 32 here, so 32 of the 64 slots are always free.
		if (!exhausted && qcount < __popc(idle)) {
			if (grabCnt > 0) {
				double raw[9];
				if (kSolve) {
					// pairs into the stage (bulk copies, or plain loads), then 4 passes of 8 pairs on the quads
					double* sp = (double*)S.stage; // [32][3] starts, then [32][3] goals
					if (grabBulk) {
						mbar_wait(&S.bar, parity);
						parity ^= 1u;
					} else {
						for (int i = (int)lane_id; i < grabCnt * 3; i += 32) {
							sp[i] = sa.starts[3 * grabBase + i];
							sp[96 + i] = sa.goals[3 * grabBase + i];
						}
						for (int i = grabCnt * 3 + (int)lane_id; i < 96; i += 32) { // pad: the quads of missing pairs solve (0,0,0)->(0,0,0)
							sp[i] = 0.0;
							sp[96 + i] = 0.0;
						}
						__syncwarp();
					}
					int myWord = -1;
					double mt = 0, mu = 0, mv = 0;
					#pragma unroll 1
					for (int pass = 0; pass < 4; pass++) {
						const int j = pass * 8 + (int)(lane_id >> 2);
						const Pose ps = { sp[3 * j], sp[3 * j + 1], sp[3 * j + 2] };
						const Pose pg = { sp[96 + 3 * j], sp[96 + 3 * j + 1], sp[96 + 3 * j + 2] };
						int w;
						double t, u, v;
						rs_solve_quad(ps, pg, r, sa.rev, sa.fwd, sa.sw, sa.shortest, w, t, u, v);
						// pair 8 * pass + k was solved by quad k: lane L of this pass's octet takes its own pair's result
						const int srcLane = (int)(lane_id & 7) << 2;
						const int w2 = __shfl_sync(0xffffffffu, w, srcLane);
						const double t2 = __shfl_sync(0xffffffffu, t, srcLane), u2 = __shfl_sync(0xffffffffu, u, srcLane);
						const double v2 = __shfl_sync(0xffffffffu, v, srcLane);
						if ((int)(lane_id >> 3) == pass) {
							myWord = w2;
							mt = t2; mu = u2; mv = v2;
						}
					}
					raw[0] = sp[3 * lane_id];
					raw[1] = sp[3 * lane_id + 1];
					raw[2] = sp[3 * lane_id + 2];
					__syncwarp(); // every lane has read its pair: the stage may be overwritten by the next copy
					RSSeg seg;
					double mLength = 0.0;
					if (myWord >= 0)
						mLength = rs_build_segment(myWord, mt, mu, mv, seg);
					else { // PathSegment(): all motions invalid
						for (int i = 0; i < 5; i++)
							seg.len[i] = PPX_INF;
						seg.code = 0xAAAAA;
					}
					#pragma unroll
					for (int i = 0; i < 5; i++)
						raw[3 + i] = seg.len[i];
					raw[8] = __longlong_as_double((long long)seg.code);
					if ((int)lane_id < grabCnt) {
						const int64_t idx = grabBase + lane_id;
						double* dst = (double*)(sa.solved + idx);
						#pragma unroll
						for (int i = 0; i < 9; i++)
							dst[i] = raw[i];
						if (sa.word)
							sa.word[idx] = myWord;
						if (sa.length)
							sa.length[idx] = mLength * r;
						if (sa.cost)
							sa.cost[idx] = myWord >= 0 ? rs_compute_cost(seg, mLength, r, sa.rev, sa.fwd, sa.sw) : __int_as_float(0x7f800000);
					}
				} else if (grabBulk) {
					mbar_wait(&S.bar, parity);
					parity ^= 1u;
					const double* src = (const double*)(S.stage + (size_t)lane_id * sizeof(ppx_rs_edge));
					#pragma unroll
					for (int i = 0; i < 9; i++)
						raw[i] = src[i];
				} else if ((int)lane_id < grabCnt) {
					const double* src = (const double*)(edges + grabBase + lane_id);
					#pragma unroll
					for (int i = 0; i < 9; i++)
						raw[i] = src[i];
				}
				unsigned long long mine = 0ull;
				if (state != LANE_IDLE)
					mine = 1ull << slot;
				if ((int)lane_id < qcount)
					mine |= 1ull << S.order[(qhead + lane_id) & (kQueueCap - 1)];
				const unsigned usedLo = __reduce_or_sync(0xffffffffu, (unsigned)mine);
				const unsigned usedHi = __reduce_or_sync(0xffffffffu, (unsigned)(mine >> 32));
				const unsigned freeLo = ~usedLo, freeHi = ~usedHi;
				const int nlo = __popc(freeLo);
				if ((int)lane_id < grabCnt) {
					const int s_new = (int)lane_id < nlo ? (int)__fns(freeLo, 0, (int)lane_id + 1) : 32 + (int)__fns(freeHi, 0, (int)lane_id - nlo + 1);
					S.order[(qhead + qcount + lane_id) & (kQueueCap - 1)] = (unsigned char)s_new;
					rs2_preprocess(raw, (uint32_t)(grabBase + lane_id), fm, map, r, inv_r, rc, S.q[s_new]);
				}
				qcount += grabCnt;
				__syncwarp();
			}
			unsigned long long base = 0;
			if (lane_id == 0)
				base = atomicAdd(queue, (unsigned long long)kGrab2);
			base = __shfl_sync(0xffffffffu, base, 0);
			if ((int64_t)base >= n) {
				exhausted = true;
				grabCnt = 0;
			} else {
				grabBase = (int64_t)base;
				grabCnt = (int)(n - grabBase < kGrab2 ? n - grabBase : kGrab2);
				grabBulk = aligned && grabCnt == kGrab2;
				if (grabBulk && lane_id == 0) {
					asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // the lanes' reads of the stage before the copy's writes
					if (kSolve) {
						asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&S.bar)), "r"(2u * 768u) : "memory");
						asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
										 smem_u32(S.stage)),
							"l"(sa.starts + 3 * grabBase), "r"(768u), "r"(smem_u32(&S.bar))
							: "memory");
						asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
										 smem_u32(S.stage + 768)),
							"l"(sa.goals + 3 * grabBase), "r"(768u), "r"(smem_u32(&S.bar))
							: "memory");
					} else
						bulk_copy_g2s(S.stage, edges + grabBase, kGrab2 * (uint32_t)sizeof(ppx_rs_edge), &S.bar);
				}
			}
		}
		// ---- B. idle lanes pick up their next edge from the queue
		if (idle && qcount > 0) {
			const int rank = __popc(idle & lt_mask);
			if (state == LANE_IDLE && rank < qcount) {
				slot = S.order[(qhead + rank) & (kQueueCap - 1)];
				m = 0;
				length = 0.0;
				lastValidLength = 0.0;
				ns = 0;
				state = (S.q[slot][22] >> 24) & 1u ? LANE_PARKED : LANE_ACTIVE;
			}
			const int take = min(__popc(idle), qcount);
			qhead = (qhead + take) & (kQueueCap - 1);
			qcount -= take;
			idle = __ballot_sync(0xffffffffu, state == LANE_IDLE);
		}
		const unsigned parked = __ballot_sync(0xffffffffu, state == LANE_PARKED);
		const unsigned running = ~(idle | parked);
		const uint32_t* rec = S.q[slot];
		// ---- C. samples the filter could not certify: exact FP64 evaluation, batched
		if (parked && (__popc(parked) >= kParkBatch || !running)) {
			if (state == LANE_PARKED) {
				const double pathLength = __hiloint2double((int)rec[1], (int)rec[0]);
				const int64_t idx = (int64_t)rec[24];
				ExactIO2 io { pathLength, length, lastValidLength, ns, 0.0f, false, false };
				io = exact_sample_lane2(io, map, edges + idx, r, minSafeRadius, minStep);
				length = io.length;
				lastValidLength = io.lastValidLength;
				ns = io.ns;
				if (kAudit)
					atomicAdd(&audit->parked, 1ull);
				state = LANE_ACTIVE;
				if (io.done) {
					valid[idx] = io.verdict ? 1 : 0;
					last[idx] = io.last;
					if (samples)
						samples[idx] = ns;
					state = LANE_IDLE;
				} else if ((rec[22] >> 24) & 1u)
					state = LANE_PARKED;
			}
			continue;
		}
		if (!running) {
			if (exhausted && qcount == 0 && !parked)
				break;
			continue;
		}
		// (A cooperative tail -- eight lanes evaluating the next eight minStep-spaced samples of one remaining chain together,
		// accepting the prefix whose steps really were minStep -- was built and measured: identical results, 255 instead of
		// 192 us per 1M edges.  Long chains skim obstacles with a slack of 0.1-0.3 m, i.e. steps just ABOVE minStep, so the
		// prediction rarely holds and every tail iteration pays the extra shuffles.  Removed.)
		// ---- D. one sample per active lane
		if (state == LANE_ACTIVE) {
			const double pathLength = __hiloint2double((int)rec[1], (int)rec[0]);
			const uint32_t cw = rec[22];
			const int nm = (int)((cw >> 20) & 7u);
			// motion of the sample: the last knot whose cumulative length the sample has passed
			const float sf = (float)length;
			while (m + 1 < nm && sf > __uint_as_float(rec[5 + 4 * (m + 1)]))
				m++;
			const float kx = __uint_as_float(rec[2 + 4 * m]), ky = __uint_as_float(rec[3 + 4 * m]);
			const float thK = __uint_as_float(rec[4 + 4 * m]);
			const float rem = (float)(length - (double)__uint_as_float(rec[5 + 4 * m]));
			const int st = (cw >> (4 * m)) & 3, dr = (cw >> (4 * m + 2)) & 3;
			const bool straight = st == 1;
			const float sgnAdv = dr == 1 ? -1.0f : 1.0f, sgnTurn = st == 2 ? -1.0f : 1.0f;
			const float adv = sgnAdv * rem;
			const float a = adv * inv_r * sgnTurn;
			const float phi = 0.5f * a;
			const float ang = fast_wrap(straight ? thK : thK + phi);
			const float chord = (2.0f * rc) * PPX_FSIN(sgnTurn * phi);
			const float Lc = straight ? adv * fm.inv_res : chord;
			const float px = kx + Lc * PPX_FCOS(ang), py = ky + Lc * PPX_FSIN(ang);
			// certified sample (RSFastLane::sample)
			const int rowb = (int)(short)(rec[23] & 0xffffu), colb = (int)(short)(rec[23] >> 16);
			const float fx = floorf(px), fy = floorf(py);
			const float gx = px - fx, gy = py - fy;
			const bool near = fminf(gx, gy) < kCellEps || fmaxf(gx, gy) > 1.0f - kCellEps;
			const int row = rowb + (int)fx, col = colb + (int)fy;
			const int ib = imin(imin(row - fm.rb0, fm.rb1 - row), imin(col - fm.cb0, fm.cb1 - col));
			int rcode = 2; // 0 continue, 1 finished, 2 park
			bool verdict = false;
			float lastOut = 1.0f;
			if (!near && ib >= 0) {
				const float d = PPX_LDG(map.clearance + (size_t)row * map.cols + col);
				if (!(d >= minSafeRadius)) { // IsStateValid false
					ns++;
					verdict = false;
					lastOut = lastValidLength == 0.0 ? 0.0f : (float)(lastValidLength / pathLength);
					rcode = 1;
				} else {
					const float deltaLength = d - minSafeRadius;
					if (deltaLength <= (float)ib * map.res) { // else the distance to the border may be the smaller one
						const double lengthAtSample = length;
						ns++;
						lastValidLength = length;
						length += (double)std_maxf(deltaLength, minStep);
						rcode = 0;
						if (!(length < pathLength)) {
							verdict = true;
							lastOut = 1.0f;
							rcode = 1;
						}
						if (kAudit)
							audit_sample2(px, py, rowb, colb, pathLength, map, edges + rec[24], r, lengthAtSample, audit);
					}
				}
				if (kAudit && rcode == 1 && !verdict)
					audit_sample2(px, py, rowb, colb, pathLength, map, edges + rec[24], r, length, audit);
			}
			if (rcode == 2)
				state = LANE_PARKED;
			else if (rcode == 1) {
				const int64_t idx = (int64_t)rec[24];
				valid[idx] = verdict ? 1 : 0;
				last[idx] = lastOut;
				if (samples)
					samples[idx] = ns;
				state = LANE_IDLE;
			}
		}
	}
	if (nextQueue) // this launch must not complete before its predecessor has (stream order for whatever follows)
		asm volatile("griddepcontrol.wait;" ::: "memory");
}
