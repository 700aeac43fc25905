// libppx.so: ObstaclesHeuristic::Update (A8), the holonomic-with-obstacles cost-to-go wavefront,
// batched over goals, bit-exact with the reference's order-dependent semantics.
//
// Reference semantics (algo/heuristics.cpp:106-153, utils/frontier.h:39-48,83-91):
//   * cells pop in ascending float cost, LIFO among equal costs (sorted vector, insertion after the
//     run of equal values, pop from the back);
//   * a popped cell pushes every free, never-reached neighbour (enumeration order of
//     utils/grid.cpp:29-47, corner rule heuristics.cpp:130-132) with cost = w + cost[parent] in
//     float32, w in {1.0f, sqrtf(2)};
//   * the decrease-key branch (:142-150) compares two copies of the same value and never fires:
//     the FIRST popped neighbour fixes a cell's cost for good.  It is not Dijkstra.
//
// Parallel formulation ("bucketed frontier"):  every push costs at least parent + 1, and
// fl(c + 1) >= floor(c) + 1, so while the cells with floor(cost) == b are being popped nothing is
// inserted into bucket b any more: the bucket is FROZEN, and its pop order is a pure function of
// its content -- ascending (cost, -insertion sequence).  So one frontier bucket is processed per
// step by a whole CTA:
//   1. sort the bucket by (cost bits, ~seq)  -> pop rank of every parent;
//   2. every (parent rank, neighbour j) pair claims its neighbour cell with
//      atomicMax(tag[cell], 0x7fffffff - (rank*8+j)): the maximum is the pair that the sequential
//      reference reaches first ("first reach wins"); cells reached earlier hold 0x80000000|seq,
//      which beats every claim;
//   3. winners write cost = w + cost[parent], get insertion sequence numbers in (rank, j) order by a
//      block prefix sum (the order the reference pushes them), and are appended to bucket
//      floor(cost) in {b+1, b+2} of a 4-slot ring.
// The frontier is always compact (bucket lists, never a sweep over the grid); one CTA per goal,
// as many goals in flight as CTAs fit on the 148 SMs.
#include <algorithm>
#include <mutex>
#include "ppx_ctx.h"

using namespace ppx;

namespace {

constexpr int kRing = 4;
constexpr uint32_t kUnreached = 0u;
constexpr uint32_t kSeqFlag = 0x80000000u;

struct WfWork {
	uint32_t* tag; // [ctas][cells]   0 unreached; claim (< 2^31) during a step; kSeqFlag|seq once reached
	unsigned long long* keys; // [ctas][kRing][cap]  (cost bits << 32) | ~seq
	uint32_t* cells; // [ctas][kRing][cap]  row << 16 | col
	uint32_t cap; // power of two
};

__device__ __forceinline__ unsigned long long make_key(float cost, uint32_t seq)
{
	return ((unsigned long long)__float_as_uint(cost) << 32) | (unsigned long long)(0xffffffffu - seq);
}

// Claim of neighbour j by the parent with pop rank `rank`: larger = earlier in the reference's order.
__device__ __forceinline__ uint32_t claim_key(uint32_t rank, int j) { return 0x7fffffffu - (rank * 8u + (uint32_t)j); }

// neighbour enumeration order of GridCellPosition::GetNeighbors, utils/grid.cpp:29-47
// { 0, -1, 1, 0, -1, 1, -1, 1 } / { -1, -1, -1, 1, 1, 1, 0, 0 }; functions so that they fold in the unrolled loops over j
__host__ __device__ constexpr int wf_dr(int j) { return j == 1 || j == 4 || j == 6 ? -1 : j == 2 || j == 5 || j == 7 ? 1 : 0; }
__host__ __device__ constexpr int wf_dc(int j) { return j <= 2 ? -1 : j <= 5 ? 1 : 0; }

// In-place bitonic sort of n (power of two) (key, cell) pairs by the whole CTA.
__device__ void cta_bitonic_sort(unsigned long long* keys, uint32_t* cells, uint32_t n)
{
	for (uint32_t k = 2; k <= n; k <<= 1) {
		for (uint32_t j = k >> 1; j > 0; j >>= 1) {
			for (uint32_t p = threadIdx.x; p < (n >> 1); p += blockDim.x) {
				uint32_t i = ((p & ~(j - 1)) << 1) | (p & (j - 1));
				uint32_t q = i | j;
				unsigned long long a = keys[i], b = keys[q];
				bool up = (i & k) == 0;
				if ((a > b) == up) {
					keys[i] = b;
					keys[q] = a;
					uint32_t t = cells[i];
					cells[i] = cells[q];
					cells[q] = t;
				}
			}
			__syncthreads();
		}
	}
}

// Pop order of a bucket of up to sortCap / 4 entries.  Every warp sorts runs of 64 entries in registers (bitonic network
// over shuffles, two entries per lane, straight from the ring in global memory), then every entry finds its rank: its
// position in its own run plus, for every other run, the number of smaller keys there (binary search; keys are unique).
// O(L (log 64)^2 / 32 + L (L / 64) log 64) against the O(L^2) of ranking by counting, which was 58% of the kernel's
// instructions (profiles/r2k_k_wavefront_4096goals_ncu_full_summary.csv).  Padding keys are ~0: they sort to the end of
// the last run and are never counted.
__device__ void cta_run_merge_sort(const unsigned long long* gk, const uint32_t* gc, uint32_t L,
	unsigned long long* runKeys, uint32_t* runCells, unsigned long long* outKeys, uint32_t* outCells)
{
	const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
	const uint32_t nruns = (L + 63) >> 6;
	unsigned long long* dstK = nruns == 1 ? outKeys : runKeys;
	uint32_t* dstC = nruns == 1 ? outCells : runCells;
	for (uint32_t run = warp; run < nruns; run += nwarps) {
		const uint32_t base = run << 6, i0 = base + 2 * lane;
		// entry e = 2 * lane + t of the run; the ring slot is a multiple of 1024 entries long: the vector loads stay inside
		const ulonglong2 kk = ((const ulonglong2*)gk)[(base >> 1) + lane];
		const uint2 cc = ((const uint2*)gc)[(base >> 1) + lane];
		unsigned long long k0 = i0 < L ? kk.x : ~0ull, k1 = i0 + 1 < L ? kk.y : ~0ull;
		uint32_t c0 = cc.x, c1 = cc.y;
		#pragma unroll
		for (int k = 2; k <= 64; k <<= 1) {
			const bool asc = ((2 * lane) & k) == 0; // k = 2: bit 1 of e is bit 0 of the lane; k = 64: always ascending
			#pragma unroll
			for (int j = k >> 1; j > 0; j >>= 1) {
				if (j == 1) {
					if ((k0 > k1) == asc) {
						const unsigned long long tk = k0;
						k0 = k1;
						k1 = tk;
						const uint32_t tc = c0;
						c0 = c1;
						c1 = tc;
					}
				} else {
					const int pl = j >> 1;
					const unsigned long long o0 = __shfl_xor_sync(0xffffffffu, k0, pl), o1 = __shfl_xor_sync(0xffffffffu, k1, pl);
					const uint32_t oc0 = __shfl_xor_sync(0xffffffffu, c0, pl), oc1 = __shfl_xor_sync(0xffffffffu, c1, pl);
					const bool keepMin = ((lane & pl) == 0) == asc;
					if (keepMin ? o0 < k0 : o0 > k0) {
						k0 = o0;
						c0 = oc0;
					}
					if (keepMin ? o1 < k1 : o1 > k1) {
						k1 = o1;
						c1 = oc1;
					}
				}
			}
		}
		((ulonglong2*)dstK)[(base >> 1) + lane] = make_ulonglong2(k0, k1);
		((uint2*)dstC)[(base >> 1) + lane] = make_uint2(c0, c1);
	}
	__syncthreads();
	if (nruns == 1)
		return;
	for (uint32_t i = threadIdx.x; i < L; i += blockDim.x) {
		const unsigned long long mine = runKeys[i];
		const uint32_t own = i >> 6;
		uint32_t rank = i & 63u;
		for (uint32_t r = 0; r < nruns; r++) {
			if (r == own)
				continue;
			const unsigned long long* R = runKeys + (r << 6);
			uint32_t cnt = 0;
			#pragma unroll
			for (uint32_t s = 32; s > 0; s >>= 1)
				cnt += R[cnt + s - 1] < mine ? s : 0u;
			cnt += R[cnt] < mine ? 1u : 0u;
			rank += cnt;
		}
		outKeys[rank] = mine;
		outCells[rank] = runCells[i];
	}
	__syncthreads();
}

// Exclusive prefix sum of one value per thread over the CTA; returns the exclusive prefix and the
// CTA total through *total.  s_warp must hold 32 entries.  The caller must pass a barrier before the next call.
__device__ uint32_t cta_exclusive_scan(uint32_t v, uint32_t* s_warp, uint32_t* total)
{
	const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const unsigned nwarps = (blockDim.x + 31) >> 5;
	uint32_t incl = v;
	#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
		if (lane >= (unsigned)o)
			incl += t;
	}
	if (lane == 31)
		s_warp[warp] = incl;
	__syncthreads();
	if (warp == 0) {
		uint32_t w = lane < nwarps ? s_warp[lane] : 0;
		uint32_t wi = w;
		#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
			if (lane >= (unsigned)o)
				wi += t;
		}
		s_warp[lane] = wi - w; // exclusive prefix of the warp sums
		if (lane == 31)
			*total = wi;
	}
	__syncthreads();
	return s_warp[warp] + incl - v; // s_warp / *total are rewritten by the next call only: the caller has a barrier before it
}

// One goal per CTA at a time; CTAs take goals from a global cursor.  kThreads / kMinBlocks only set the register budget
// (the kernel is bound by barrier and load latency per bucket: more resident CTAs per SM hide it).
template <int kThreads, int kMinBlocks>
__global__ void __launch_bounds__(kThreads, kMinBlocks) k_wavefront(MapDev map, const uint8_t* __restrict__ occ8, int64_t nGoals, const double* __restrict__ goals,
	float* __restrict__ costOut, uint8_t* __restrict__ exploredOut, WfWork work, uint32_t sortCap,
	unsigned long long* __restrict__ queue, int* __restrict__ status)
{
	extern __shared__ __align__(16) unsigned long long s_dyn[];
	unsigned long long* s_keys = s_dyn; // [sortCap]
	unsigned long long* s_keys2 = s_dyn + sortCap; // [sortCap/4] scratch for the rank sort input
	uint32_t* s_cells = (uint32_t*)(s_dyn + sortCap + sortCap / 4); // [sortCap]
	uint32_t* s_cells2 = s_cells + sortCap; // [sortCap/4]
	__shared__ uint32_t s_cnt[kRing];
	__shared__ uint32_t s_warp[32];
	__shared__ uint32_t s_total;
	__shared__ long long s_goal;
	__shared__ int s_err;

	const int rows = map.rows, cols = map.cols;
	const size_t ncells = (size_t)rows * cols;
	uint32_t* tag = work.tag + (size_t)blockIdx.x * ncells;
	unsigned long long* ringKeys = work.keys + (size_t)blockIdx.x * kRing * work.cap;
	uint32_t* ringCells = work.cells + (size_t)blockIdx.x * kRing * work.cap;
	const float sqrt2 = sqrtf(2.0f); // std::sqrt(2.0f), heuristics.cpp:134

	while (true) {
		if (threadIdx.x == 0)
			s_goal = (long long)atomicAdd(queue, 1ull);
		__syncthreads();
		const long long g = s_goal;
		if (g >= nGoals)
			return;
		float* cost = costOut + (size_t)g * ncells;

		// heuristics.cpp:108-113: every cell infinity / not explored
		for (size_t i = threadIdx.x; i < ncells; i += blockDim.x) {
			cost[i] = __int_as_float(0x7f800000);
			tag[i] = kUnreached;
		}
		if (threadIdx.x < kRing)
			s_cnt[threadIdx.x] = 0;
		if (threadIdx.x == 0)
			s_err = 0;
		// goal cell, bounded WorldPositionToGridCell (occupancy_map.h:106-117,180-183)
		const int r0 = trunc_i32((goals[3 * g] - map.ox) / map.resd);
		const int c0 = trunc_i32((goals[3 * g + 1] - map.oy) / map.resd);
		const bool startValid = r0 >= 0 && r0 < rows && c0 >= 0 && c0 < cols;
		__syncthreads();
		uint32_t pending = 0, counter = 1;
		int b = 0;
		const int maxBucket = (int)min((size_t)0x7ffffff0, 2 * ncells + 8);
		if (startValid) {
			if (threadIdx.x == 0) {
				cost[(size_t)r0 * cols + c0] = 0.0f;
				tag[(size_t)r0 * cols + c0] = kSeqFlag | 0u;
				ringKeys[0] = make_key(0.0f, 0);
				ringCells[0] = ((uint32_t)r0 << 16) | (uint32_t)c0;
				s_cnt[0] = 1;
			}
			pending = 1;
		}
		__syncthreads();

		while (pending > 0) {
			const int slot = b & (kRing - 1);
			const uint32_t L = s_cnt[slot];
			if (L == 0) {
				// cannot happen while the bookkeeping is right (every pending cell sits in bucket b+1 or b+2);
				// the guard turns a would-be endless scan into an error status
				if (++b > maxBucket) {
					if (threadIdx.x == 0) {
						atomicExch(status, 2);
						printf("ppx wavefront: goal %lld lost %u pending cells at bucket %d (seq %u)\n", g, pending, b, counter);
					}
					break;
				}
				continue;
			}
			unsigned long long* gk = ringKeys + (size_t)slot * work.cap;
			uint32_t* gc = ringCells + (size_t)slot * work.cap;
			// ---- 1. pop order of the frozen bucket -------------------------------------------------
			const unsigned long long* sk;
			const uint32_t* sc;
			if (L <= sortCap / 4) {
				cta_run_merge_sort(gk, gc, L, s_keys2, s_cells2, s_keys, s_cells);
				sk = s_keys;
				sc = s_cells;
			} else {
				uint32_t n = 2;
				while (n < L)
					n <<= 1;
				if (n <= sortCap) {
					for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
						s_keys[i] = i < L ? gk[i] : ~0ull;
						s_cells[i] = i < L ? gc[i] : 0u;
					}
					__syncthreads();
					cta_bitonic_sort(s_keys, s_cells, n);
					sk = s_keys;
					sc = s_cells;
				} else { // oversized bucket: sort it where it lies (n <= cap is guaranteed by the append check)
					for (uint32_t i = L + threadIdx.x; i < n; i += blockDim.x) {
						gk[i] = ~0ull;
						gc[i] = 0u;
					}
					__syncthreads();
					cta_bitonic_sort(gk, gc, n);
					sk = gk;
					sc = gc;
				}
			}
			// ---- 2. claims: contiguous block of parents per thread -----------------------------------
			const uint32_t per = (L + blockDim.x - 1) / blockDim.x;
			const uint32_t first = threadIdx.x * per;
			const uint32_t lastp = first + per < L ? first + per : L;
			// The eight occupancy bytes of a parent's neighbours are loaded together (one latency, not up to ten dependent
			// ones: the corner rule reads the 4-neighbours, which are among the eight), and so are the tags in step 3, which
			// looks only at the neighbours this thread claimed (claimMask: bit 8 * (rank - first) + j, first eight parents).
			unsigned long long claimMask = 0ull;
			for (uint32_t rank = first; rank < lastp; rank++) {
				const uint32_t cell = sc[rank];
				const int r = (int)(cell >> 16), c = (int)(cell & 0xffffu);
				uint8_t o[8];
				#pragma unroll
				for (int j = 0; j < 8; j++) {
					const int rr = r + wf_dr(j), cc = c + wf_dc(j);
					const bool in = rr >= 0 && rr < rows && cc >= 0 && cc < cols;
					o[j] = in ? occ8[(size_t)rr * cols + cc] : (uint8_t)1; // outside the map: never a neighbour
				}
				#pragma unroll
				for (int j = 0; j < 8; j++) {
					if (o[j])
						continue;
					// corner rule (heuristics.cpp:130-132): (rr, c) is neighbour 6 / 7 (dr = -1 / +1), (r, cc) is neighbour 0 / 3
					if (wf_dr(j) != 0 && wf_dc(j) != 0 && o[wf_dr(j) < 0 ? 6 : 7] && o[wf_dc(j) < 0 ? 0 : 3])
						continue;
					// reached cells hold kSeqFlag|seq >= 2^31 > any claim: the atomicMax leaves them alone
					atomicMax(&tag[(size_t)(r + wf_dr(j)) * cols + (c + wf_dc(j))], claim_key(rank, j));
					if (rank - first < 8)
						claimMask |= 1ull << (8 * (rank - first) + j);
				}
			}
			__syncthreads();
			// ---- 3. winners ------------------------------------------------------------------------
			uint32_t wins = 0;
			unsigned long long wonMask = 0ull; // bit 8 * (rank - first) + j for the first eight parents of the thread
			for (uint32_t rank = first; rank < lastp; rank++) {
				const uint32_t cell = sc[rank];
				const int r = (int)(cell >> 16), c = (int)(cell & 0xffffu);
				const bool cachedClaims = rank - first < 8;
				const uint32_t cbits = cachedClaims ? (uint32_t)(claimMask >> (8 * (rank - first))) & 0xffu : 0xffu;
				if (cbits == 0)
					continue;
				uint32_t t[8];
				#pragma unroll
				for (int j = 0; j < 8; j++) {
					const int rr = r + wf_dr(j), cc = c + wf_dc(j);
					const bool look = ((cbits >> j) & 1u) && rr >= 0 && rr < rows && cc >= 0 && cc < cols;
					t[j] = look ? tag[(size_t)rr * cols + cc] : 0u; // 0 is never a claim key
				}
				#pragma unroll
				for (int j = 0; j < 8; j++) {
					const bool won = t[j] == claim_key(rank, j);
					wins += won ? 1u : 0u;
					if (won && rank - first < 8)
						wonMask |= 1ull << (8 * (rank - first) + j);
				}
			}
			uint32_t seq = counter + cta_exclusive_scan(wins, s_warp, &s_total);
			const uint32_t total = s_total;
			for (uint32_t rank = first; rank < lastp; rank++) {
				const uint32_t cell = sc[rank];
				const int r = (int)(cell >> 16), c = (int)(cell & 0xffffu);
				const float parentCost = __uint_as_float((uint32_t)(sk[rank] >> 32));
				const bool cached = rank - first < 8;
				const uint32_t bits = cached ? (uint32_t)(wonMask >> (8 * (rank - first))) & 0xffu : 0xffu;
				if (bits == 0)
					continue;
				#pragma unroll
				for (int j = 0; j < 8; j++) {
					if (!((bits >> j) & 1u))
						continue;
					const int rr = r + wf_dr(j), cc = c + wf_dc(j);
					if (rr < 0 || rr >= rows || cc < 0 || cc >= cols)
						continue;
					const size_t ni = (size_t)rr * cols + cc;
					if (!cached && tag[ni] != claim_key(rank, j))
						continue;
					const float w = (wf_dr(j) == 0 || wf_dc(j) == 0) ? 1.0f : sqrt2;
					const float pathCost = w + parentCost;
					cost[ni] = pathCost;
					tag[ni] = kSeqFlag | seq;
					const int nb = (int)pathCost;
					const int ns = nb & (kRing - 1);
					const uint32_t pos = atomicAdd(&s_cnt[ns], 1u);
					if (pos < work.cap && nb > b && nb <= b + 2) {
						ringKeys[(size_t)ns * work.cap + pos] = make_key(pathCost, seq);
						ringCells[(size_t)ns * work.cap + pos] = ((uint32_t)rr << 16) | (uint32_t)cc;
					} else {
						s_err = 1;
					}
					seq++;
				}
			}
			__syncthreads();
			// slot b is appended to again in step b + 2 at the earliest (bucket b + 4): barriers in between
			if (threadIdx.x == 0)
				s_cnt[slot] = 0;
			pending = pending - L + total;
			counter += total;
			b++;
			if (s_err) { // written before the barrier above; the next write is behind the barriers of the next step
				pending = 0;
				if (threadIdx.x == 0)
					atomicExch(status, 1);
			}
		}
		// every reached cell has been popped when the frontier drains: explored <=> cost finite
		if (exploredOut) {
			uint8_t* ex = exploredOut + (size_t)g * ncells;
			for (size_t i = threadIdx.x; i < ncells; i += blockDim.x)
				ex[i] = cost[i] != __int_as_float(0x7f800000) ? 1 : 0;
		}
		__syncthreads();
	}
}

}

// ---- ppx_wavefront_dev_static: the same kernel with its scratch in a file-scope device array (no allocation at call time) ----
// Every GPU that loads the library holds these 64 MiB until the process exits.  In host code `g_wf_static` is a host-side
// stand-in for the symbol, not a device address: the address a kernel can use comes from cudaGetSymbolAddress, once per
// device, with that device current (wf_static_base).
constexpr size_t kWfStaticBytes = (size_t)PPX_WAVEFRONT_STATIC_BYTES;
constexpr size_t kWfStaticHeader = 256; // control words: work-queue cursor at +0, status word at +64
static __device__ __align__(256) unsigned char g_wf_static[kWfStaticBytes];

namespace {

constexpr int kWfMaxDevices = 64;
std::mutex g_wfStaticMutex;
unsigned char* g_wfStaticPtr[kWfMaxDevices] = {};

// Device address of g_wf_static on ctx's device (the caller holds PPX_GUARD: that device is current).
int wf_static_base(ppx_ctx* ctx, unsigned char** out)
{
	if (ctx->device < 0 || ctx->device >= kWfMaxDevices)
		return ppx_fail(ctx, PPX_ERR_INVALID, "wavefront (static scratch): device ordinal %d out of range", ctx->device);
	std::lock_guard<std::mutex> lock(g_wfStaticMutex);
	if (!g_wfStaticPtr[ctx->device]) {
		void* p = nullptr;
		cudaError_t err = cudaGetSymbolAddress(&p, g_wf_static);
		if (err != cudaSuccess || !p)
			return ppx_fail(ctx, PPX_ERR_CUDA, "wavefront (static scratch): cudaGetSymbolAddress: %s", cudaGetErrorString(err));
		g_wfStaticPtr[ctx->device] = (unsigned char*)p;
	}
	*out = g_wfStaticPtr[ctx->device];
	return PPX_OK;
}

}

// `staticBase`: nullptr = scratch from the context's grow-only allocation (ppx_scratch), else the device address of
// g_wf_static: nothing is allocated, the number of goals in flight is what fits into the array, and a map of which not even
// one goal fits is rejected.  `d_statusOut`: caller-owned device status word (static form only; may be null).
static int wavefront_launch_impl(ppx_ctx* ctx, int64_t n, const double* d_goals, float* d_cost, uint8_t* d_explored,
	unsigned char* staticBase, int* d_statusOut)
{
	const int rows = ctx->desc.rows, cols = ctx->desc.cols;
	if (rows > 65535 || cols > 65535)
		return ppx_fail(ctx, PPX_ERR_INVALID, "wavefront: map larger than 65535 cells per side");
	const size_t ncells = (size_t)rows * cols;
	// team size and sort capacity from the map perimeter (a wave front is a closed curve through the map)
	const int perimeter = 2 * (rows + cols);
	int threads = perimeter <= 1024 ? 128 : perimeter <= 2048 ? 256 : perimeter <= 4096 ? 512 : 1024;
	uint32_t sortCap = (uint32_t)threads * 8;
	// few goals: twice the team shortens a single wavefront (C2: 5.5 -> 4.3 ms); for batches the smaller team wins
	// (4096 goals: 45.9 ms with 256 threads, 59.5 ms with 512, 52.9 ms with 128)
	if (threads == 256 && n <= (int64_t)ctx->num_sms) {
		threads = 512;
		sortCap = 4096;
	}
	size_t smem = (size_t)sortCap * 8 + sortCap / 4 * 8 + (size_t)sortCap * 4 + sortCap / 4 * 4;
	using Kernel = void (*)(MapDev, const uint8_t*, int64_t, const double*, float*, uint8_t*, WfWork, uint32_t, unsigned long long*, int*);
	// 256 threads: 4 CTAs per SM (51 registers) for a batch the GPU holds at once -- shortest single wavefront --, 5 (48
	// registers) for larger batches: 4096 goals 46.5 against 48.8 ms; 6 and 7 CTAs per SM spill and are slower (51, 79 ms)
	static const int envBlocks = getenv("PPX_WF_BLOCKS") ? atoi(getenv("PPX_WF_BLOCKS")) : 0;
	const int wantBlocks = envBlocks ? envBlocks : n > (int64_t)ctx->num_sms * 4 ? 5 : 4;
	Kernel kernel = threads == 128 ? k_wavefront<128, 8>
		: threads == 256 ? (wantBlocks <= 4 ? k_wavefront<256, 4> : k_wavefront<256, 5>)
		: threads == 512 ? k_wavefront<512, 2>
		: k_wavefront<1024, 1>;
	PPX_CUDA(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
	int perSm = 1;
	PPX_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kernel, threads, smem));
	if (perSm < 1)
		perSm = 1;
	int ctas = (int)std::min<int64_t>(n, (int64_t)ctx->num_sms * perSm);
	uint32_t cap = 1024;
	while (cap < (uint32_t)(8 * (rows + cols)))
		cap <<= 1;
	while ((size_t)cap > 2 * ncells && cap > 1024)
		cap >>= 1;
	WfWork work;
	work.cap = cap;
	if (staticBase) {
		// per goal in flight: tag grid + ring keys + ring cells, each part rounded to 256 bytes
		const size_t perCta = (((size_t)ncells * 4 + 255) & ~(size_t)255) + (size_t)kRing * cap * 8 + (((size_t)kRing * cap * 4 + 255) & ~(size_t)255);
		const size_t room = kWfStaticBytes - kWfStaticHeader;
		if (perCta > room)
			return ppx_fail(ctx, PPX_ERR_CAPACITY, "wavefront (static scratch): a %d x %d map needs %zu bytes of scratch per goal in flight, the "
				"static array holds %zu (maps of up to %d cells with rows + cols <= 65536 always fit); use ppx_wavefront_dev", rows, cols, perCta,
				room, PPX_WAVEFRONT_STATIC_MAX_CELLS);
		ctas = (int)std::min<size_t>((size_t)ctas, room / perCta);
	}
	char* base;
	size_t bytesTag = (size_t)ctas * ncells * 4, bytesKeys = (size_t)ctas * kRing * cap * 8, bytesCells = (size_t)ctas * kRing * cap * 4;
	size_t offKeys = (bytesTag + 255) & ~(size_t)255, offCells = offKeys + bytesKeys, offStatus = offCells + ((bytesCells + 255) & ~(size_t)255);
	unsigned long long* d_cursor = ctx->d_queue;
	int* d_status;
	if (staticBase) {
		// the array keeps the last call's contents: the cursor and the status word are cleared here, the tag grid of a goal by
		// the kernel itself before it starts that goal; the ring slots are written before they are read
		base = (char*)staticBase + kWfStaticHeader;
		d_cursor = (unsigned long long*)staticBase;
		d_status = d_statusOut ? d_statusOut : (int*)(staticBase + 64);
	} else {
		PPX_CHECK(ppx_scratch(ctx, 4, offStatus + 256, (void**)&base));
		d_status = (int*)(base + offStatus);
	}
	work.tag = (uint32_t*)base;
	work.keys = (unsigned long long*)(base + offKeys);
	work.cells = (uint32_t*)(base + offCells);
	PPX_CUDA(ctx, cudaMemsetAsync(d_status, 0, sizeof(int), ctx->stream));
	PPX_CUDA(ctx, cudaMemsetAsync(d_cursor, 0, sizeof(unsigned long long), ctx->stream));
	kernel<<<ctas, threads, smem, ctx->stream>>>(ctx->map_dev(), ctx->d_occ8, n, d_goals, d_cost, d_explored, work, sortCap,
		d_cursor, d_status);
	PPX_LAUNCHED(ctx);
	ctx->d_wf_status = d_status;
	return PPX_OK;
}

int ppx_wavefront_launch(ppx_ctx* ctx, int64_t n, const double* d_goals, float* d_cost, uint8_t* d_explored)
{
	return wavefront_launch_impl(ctx, n, d_goals, d_cost, d_explored, nullptr, nullptr);
}

extern "C" {

int ppx_wavefront_dev(ppx_ctx* ctx, int64_t n, const double* d_goals, float* d_cost, uint8_t* d_explored)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	if (!ctx->has_map)
		return ppx_fail(ctx, PPX_ERR_STATE, "no map uploaded (ppx_map_upload)");
	PPX_REQUIRE(ctx, n >= 0 && (n == 0 || (d_goals && d_cost)), "bad arguments");
	if (n == 0)
		return PPX_OK;
	return ppx_wavefront_launch(ctx, n, d_goals, d_cost, d_explored);
}

int ppx_wavefront_dev_static(ppx_ctx* ctx, int64_t n, const double* d_goals, float* d_cost, uint8_t* d_explored, int32_t* d_status)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	if (!ctx->has_map)
		return ppx_fail(ctx, PPX_ERR_STATE, "no map uploaded (ppx_map_upload)");
	PPX_REQUIRE(ctx, n >= 0 && (n == 0 || (d_goals && d_cost)), "bad arguments");
	if (n == 0)
		return PPX_OK;
	unsigned char* base = nullptr;
	PPX_CHECK(wf_static_base(ctx, &base));
	return wavefront_launch_impl(ctx, n, d_goals, d_cost, d_explored, base, (int*)d_status);
}

int ppx_wavefront(ppx_ctx* ctx, int64_t n, const double* goals, float* cost, uint8_t* explored)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	if (!ctx->has_map)
		return ppx_fail(ctx, PPX_ERR_STATE, "no map uploaded (ppx_map_upload)");
	PPX_REQUIRE(ctx, n >= 0 && (n == 0 || (goals && cost)), "bad arguments");
	if (n == 0)
		return PPX_OK;
	const size_t ncells = ctx->cells;
	double* d_goals;
	float* d_cost;
	uint8_t* d_explored = nullptr;
	PPX_CHECK(ppx_scratch(ctx, 0, n * 3 * sizeof(double), (void**)&d_goals));
	PPX_CHECK(ppx_scratch(ctx, 1, n * ncells * sizeof(float), (void**)&d_cost));
	if (explored)
		PPX_CHECK(ppx_scratch(ctx, 2, n * ncells, (void**)&d_explored));
	PPX_CUDA(ctx, cudaMemcpyAsync(d_goals, goals, n * 3 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
	PPX_CHECK(ppx_wavefront_launch(ctx, n, d_goals, d_cost, d_explored));
	int status = 0;
	PPX_CUDA(ctx, cudaMemcpyAsync(&status, ctx->d_wf_status, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaMemcpyAsync(cost, d_cost, n * ncells * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
	if (explored)
		PPX_CUDA(ctx, cudaMemcpyAsync(explored, d_explored, n * ncells, cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	if (status)
		return ppx_fail(ctx, PPX_ERR_CAPACITY, "wavefront: device status %d (1: a frontier bucket overflowed its ring slot, 2: internal)", status);
	return PPX_OK;
}

} // extern "C"
