// libppx.so: multi-GPU plumbing of the path -- one context (= one GPU, one process or thread) per rank,
// NCCL over NVLink for exactly two things: broadcasting the replicated world model once per map update, and gathering
// the per-query results (fixed-size records, variable-length pose arrays by counts -> offsets).  No kernel of the path
// has an exchange step, so no collective is fused into any kernel.
//
// NCCL is resolved at run time (dlopen of libnccl.so.2, the copy the process already has -- e.g. torch's -- or the system
// one): libppx.so has no link-time dependency on it, and a single-GPU user never loads it.
#include <dlfcn.h>
#include <nccl.h>
#include <algorithm>
#include <cstring>
#include <vector>
#include "ppx_ctx.h"

namespace {

struct NcclApi {
	void* lib = nullptr;
	ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
	ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
	ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
	ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
	ncclResult_t (*GroupStart)() = nullptr;
	ncclResult_t (*GroupEnd)() = nullptr;
	const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

NcclApi* nccl_api(ppx_ctx* ctx)
{
	static NcclApi api;
	static bool tried = false;
	if (!tried) {
		tried = true;
		const char* names[] = { "libnccl.so.2", "libnccl.so" };
		for (const char* n : names) {
			api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
			if (api.lib)
				break;
		}
		if (api.lib) {
			auto sym = [&](const char* s) { return dlsym(api.lib, s); };
			api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
			api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
			api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
			api.Broadcast = (decltype(api.Broadcast))sym("ncclBroadcast");
			api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
			api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
			api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
			api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
		}
	}
	if (!api.lib || !api.GetUniqueId || !api.CommInitRank || !api.Broadcast || !api.AllGather || !api.GroupStart || !api.GroupEnd) {
		ppx_fail(ctx, PPX_ERR_STATE, "NCCL is not available (dlopen libnccl.so.2: %s)", api.lib ? "missing symbols" : dlerror());
		return nullptr;
	}
	return &api;
}

#define PPX_NCCL(ctx, api, call)                                                                                  \
	do {                                                                                                          \
		ncclResult_t r_ = (call);                                                                                 \
		if (r_ != ncclSuccess)                                                                                    \
			return ppx_fail(ctx, PPX_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call,                          \
				(api)->GetErrorString ? (api)->GetErrorString(r_) : "nccl error");                                \
	} while (0)

int require_comm(ppx_ctx* ctx)
{
	if (!ctx)
		return PPX_ERR_INVALID;
	if (!ctx->comm)
		return ppx_fail(ctx, PPX_ERR_STATE, "no communicator (ppx_comm_init)");
	return PPX_OK;
}

// counts[r] bytes from every rank's `d_local` into d_all at the exclusive prefix of the counts: one broadcast per rank,
// grouped (uneven shards need no padding)
int allgather_var(ppx_ctx* ctx, NcclApi* api, const void* d_local, const std::vector<int64_t>& bytes, void* d_all)
{
	int64_t off = 0;
	PPX_NCCL(ctx, api, api->GroupStart());
	for (int r = 0; r < ctx->comm_world; r++) {
		if (bytes[r] > 0)
			PPX_NCCL(ctx, api, api->Broadcast(r == ctx->comm_rank ? d_local : (const char*)d_all + off, (char*)d_all + off, (size_t)bytes[r],
				ncclUint8, r, (ncclComm_t)ctx->comm, ctx->stream));
		off += bytes[r];
	}
	PPX_NCCL(ctx, api, api->GroupEnd());
	return PPX_OK;
}

// every rank's value of one int64 -> host vector
int allgather_i64(ppx_ctx* ctx, NcclApi* api, int64_t mine, std::vector<int64_t>& out)
{
	int64_t* d;
	PPX_CHECK(ppx_scratch(ctx, 5, (size_t)(ctx->comm_world + 1) * sizeof(int64_t), (void**)&d));
	PPX_CUDA(ctx, cudaMemcpyAsync(d + ctx->comm_world, &mine, sizeof(int64_t), cudaMemcpyHostToDevice, ctx->stream));
	PPX_NCCL(ctx, api, api->AllGather(d + ctx->comm_world, d, 1, ncclInt64, (ncclComm_t)ctx->comm, ctx->stream));
	out.resize(ctx->comm_world);
	PPX_CUDA(ctx, cudaMemcpyAsync(out.data(), d, (size_t)ctx->comm_world * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return PPX_OK;
}

// packs the first count[q] poses of every query (max_points slots each) back to back; offsets = exclusive prefix of counts
__global__ void __launch_bounds__(256) k_pack_paths(int64_t n, const int32_t* __restrict__ count, const int64_t* __restrict__ offsets,
	const double* __restrict__ poses, int maxPoints, double* __restrict__ packed)
{
	for (int64_t q = blockIdx.x; q < n; q += gridDim.x) {
		const int c = max(0, min(count[q], maxPoints));
		const double* src = poses + (size_t)q * maxPoints * 3;
		double* dst = packed + (size_t)offsets[q] * 3;
		for (int i = threadIdx.x; i < 3 * c; i += blockDim.x)
			dst[i] = src[i];
	}
}

}

extern "C" {

int ppx_comm_unique_id(void* id_out)
{
	if (!id_out)
		return PPX_ERR_INVALID;
	NcclApi* api = nccl_api(nullptr);
	if (!api)
		return PPX_ERR_STATE;
	ncclUniqueId id;
	if (api->GetUniqueId(&id) != ncclSuccess)
		return PPX_ERR_CUDA;
	static_assert(sizeof(ncclUniqueId) == PPX_COMM_ID_BYTES, "ncclUniqueId size");
	memcpy(id_out, &id, sizeof(id));
	return PPX_OK;
}

int ppx_comm_init(ppx_ctx* ctx, const void* id, int32_t rank, int32_t world)
{
	PPX_GUARD(ctx);
	if (!ctx || !id)
		return PPX_ERR_INVALID;
	PPX_REQUIRE(ctx, world >= 1 && rank >= 0 && rank < world, "bad rank / world size");
	PPX_REQUIRE(ctx, !ctx->comm, "the context already has a communicator");
	NcclApi* api = nccl_api(ctx);
	if (!api)
		return PPX_ERR_STATE;
	ncclUniqueId uid;
	memcpy(&uid, id, sizeof(uid));
	ncclComm_t comm;
	PPX_NCCL(ctx, api, api->CommInitRank(&comm, world, uid, rank));
	ctx->comm = comm;
	ctx->comm_rank = rank;
	ctx->comm_world = world;
	return PPX_OK;
}

int ppx_comm_destroy(ppx_ctx* ctx)
{
	PPX_GUARD(ctx);
	if (!ctx)
		return PPX_ERR_INVALID;
	if (ctx->comm) {
		NcclApi* api = nccl_api(ctx);
		cudaStreamSynchronize(ctx->stream);
		if (api && api->CommDestroy)
			api->CommDestroy((ncclComm_t)ctx->comm);
		ctx->comm = nullptr;
	}
	return PPX_OK;
}

int ppx_comm_broadcast_map(ppx_ctx* ctx, int32_t root)
{
	PPX_GUARD(ctx);
	PPX_CHECK(require_comm(ctx));
	NcclApi* api = nccl_api(ctx);
	PPX_REQUIRE(ctx, root >= 0 && root < ctx->comm_world, "bad root");
	if (ctx->comm_rank == root && !ctx->has_map)
		return ppx_fail(ctx, PPX_ERR_STATE, "the root has no map (ppx_map_upload / ppx_gvd_build)");
	// 1. the description (rows, cols, frames, bounds)
	ppx_map_desc* d_desc;
	PPX_CHECK(ppx_scratch(ctx, 5, sizeof(ppx_map_desc), (void**)&d_desc));
	if (ctx->comm_rank == root)
		PPX_CUDA(ctx, cudaMemcpyAsync(d_desc, &ctx->desc, sizeof(ppx_map_desc), cudaMemcpyHostToDevice, ctx->stream));
	PPX_NCCL(ctx, api, api->Broadcast(d_desc, d_desc, sizeof(ppx_map_desc), ncclUint8, root, (ncclComm_t)ctx->comm, ctx->stream));
	ppx_map_desc desc;
	PPX_CUDA(ctx, cudaMemcpyAsync(&desc, d_desc, sizeof(ppx_map_desc), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	PPX_REQUIRE(ctx, desc.rows > 0 && desc.cols > 0 && desc.resolution > 0.0f, "broadcast map description is invalid");
	// 2. the three grids every kernel reads, into this rank's own allocation
	PPX_CHECK(ppx_map_alloc(ctx, &desc));
	const size_t cells = (size_t)desc.rows * desc.cols;
	PPX_NCCL(ctx, api, api->GroupStart());
	PPX_NCCL(ctx, api, api->Broadcast(ctx->d_sqdist, ctx->d_sqdist, cells, ncclInt32, root, (ncclComm_t)ctx->comm, ctx->stream));
	PPX_NCCL(ctx, api, api->Broadcast(ctx->d_occupancy, ctx->d_occupancy, cells, ncclInt32, root, (ncclComm_t)ctx->comm, ctx->stream));
	PPX_NCCL(ctx, api, api->Broadcast(ctx->d_pathcost, ctx->d_pathcost, cells, ncclFloat32, root, (ncclComm_t)ctx->comm, ctx->stream));
	PPX_NCCL(ctx, api, api->GroupEnd());
	if (ctx->comm_rank != root)
		ctx->has_gvd = false; // the Voronoi cell grids are not replicated (the smoother's caller builds or uploads them)
	return ppx_map_derive(ctx, &desc, true, true);
}

int ppx_comm_gather_results(ppx_ctx* ctx, const void* d_local, int64_t n_local, int32_t record_bytes, void* d_all, int64_t capacity,
	int64_t* counts)
{
	PPX_GUARD(ctx);
	PPX_CHECK(require_comm(ctx));
	NcclApi* api = nccl_api(ctx);
	PPX_REQUIRE(ctx, n_local >= 0 && record_bytes > 0 && d_all && (n_local == 0 || d_local), "bad arguments");
	std::vector<int64_t> cnt;
	PPX_CHECK(allgather_i64(ctx, api, n_local, cnt));
	int64_t total = 0;
	for (int64_t c : cnt)
		total += c;
	if (counts)
		for (int r = 0; r < ctx->comm_world; r++)
			counts[r] = cnt[r];
	PPX_REQUIRE(ctx, total <= capacity, "ppx_comm_gather_results: d_all is too small for the gathered records");
	for (auto& c : cnt)
		c *= record_bytes;
	return allgather_var(ctx, api, d_local, cnt, d_all);
}

int ppx_comm_gather_paths(ppx_ctx* ctx, int64_t n_local, const int32_t* d_count, const double* d_poses, int32_t max_points,
	int64_t* query_counts, int64_t* d_offsets_all, int64_t offsets_capacity, double* d_poses_all, int64_t poses_capacity,
	int64_t* total_poses)
{
	PPX_GUARD(ctx);
	PPX_CHECK(require_comm(ctx));
	NcclApi* api = nccl_api(ctx);
	PPX_REQUIRE(ctx, n_local >= 0 && max_points > 0 && d_offsets_all && d_poses_all && (n_local == 0 || (d_count && d_poses)),
		"bad arguments");
	// local counts -> local offsets (host prefix: this is result assembly, once per batch) -> packed local poses
	std::vector<int32_t> cnt((size_t)n_local);
	if (n_local)
		PPX_CUDA(ctx, cudaMemcpyAsync(cnt.data(), d_count, (size_t)n_local * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	std::vector<int64_t> off((size_t)n_local + 1, 0);
	for (int64_t q = 0; q < n_local; q++)
		off[q + 1] = off[q] + std::max(0, std::min(cnt[q], max_points));
	const int64_t localPoses = off[n_local];
	std::vector<int64_t> nq, np;
	PPX_CHECK(allgather_i64(ctx, api, n_local, nq));
	PPX_CHECK(allgather_i64(ctx, api, localPoses, np));
	int64_t totalQ = 0, totalP = 0, myQ0 = 0, myP0 = 0;
	for (int r = 0; r < ctx->comm_world; r++) {
		if (r == ctx->comm_rank) {
			myQ0 = totalQ;
			myP0 = totalP;
		}
		totalQ += nq[r];
		totalP += np[r];
		if (query_counts)
			query_counts[r] = nq[r];
	}
	if (total_poses)
		*total_poses = totalP;
	PPX_REQUIRE(ctx, totalQ + 1 <= offsets_capacity && totalP <= poses_capacity, "ppx_comm_gather_paths: output buffers too small");
	// global offsets of the local queries, and the local poses packed at their place in the global array
	for (auto& o : off)
		o += myP0;
	int64_t* d_off;
	double* d_packed;
	PPX_CHECK(ppx_scratch(ctx, 6, (size_t)(n_local + 1) * sizeof(int64_t) + 256 + (size_t)std::max<int64_t>(1, localPoses) * 24, (void**)&d_off));
	d_packed = (double*)((char*)d_off + (((size_t)(n_local + 1) * sizeof(int64_t) + 255) & ~(size_t)255));
	PPX_CUDA(ctx, cudaMemcpyAsync(d_off, off.data(), (size_t)(n_local + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, ctx->stream));
	if (n_local) {
		// k_pack_paths writes relative to `packed`: hand it the array shifted back by this rank's first pose
		k_pack_paths<<<(unsigned)std::min<int64_t>(n_local, (int64_t)ctx->num_sms * 8), 256, 0, ctx->stream>>>(n_local, d_count, d_off, d_poses,
			max_points, d_packed - (size_t)myP0 * 3);
		PPX_LAUNCHED(ctx);
	}
	std::vector<int64_t> qbytes(ctx->comm_world), pbytes(ctx->comm_world);
	for (int r = 0; r < ctx->comm_world; r++) {
		qbytes[r] = nq[r] * (int64_t)sizeof(int64_t);
		pbytes[r] = np[r] * 24;
	}
	PPX_CHECK(allgather_var(ctx, api, d_off, qbytes, d_offsets_all));
	PPX_CHECK(allgather_var(ctx, api, d_packed, pbytes, d_poses_all));
	// the closing offset (total number of poses) after the last query
	PPX_CUDA(ctx, cudaMemcpyAsync(d_offsets_all + totalQ, &totalP, sizeof(int64_t), cudaMemcpyHostToDevice, ctx->stream));
	PPX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	(void)myQ0;
	return PPX_OK;
}

} // extern "C"
