// mgb_comm.cu -- multi-GPU exchange: origin partition + in-place ncclAllGather of the J / Lambda slices.
// Replaces the reference's vestigial ray-distributed MPI collectives (Radiation::MPI_reduce_J,
// src/model/radiation/radiation.cpp:76-91; Lambda::MPI_gather, lambda.tpp:121-206) with the owner-computes
// origin split of SURVEY.md 8(e). NCCL is bound at run time (dlopen of libnccl.so.2) so that a host
// application which already carries its own NCCL shares that copy.
#include <dlfcn.h>

#include "mgb_ctx.h"

namespace {
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclUint8 = 1, ncclFloat64 = 8, ncclUint32 = 3 };

struct Nccl {
    void* h = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*)                                                   = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int)                            = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t)                                                      = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t)         = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t)    = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t)    = nullptr;
    ncclResult_t (*GroupStart)()                                                                 = nullptr;
    ncclResult_t (*GroupEnd)()                                                                   = nullptr;
    const char* (*GetErrorString)(ncclResult_t)                                                  = nullptr;
    bool load(std::string& err) {
        if (h) return true;
        h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) { err = std::string("dlopen(libnccl.so.2): ") + dlerror(); return false; }
#define SYM(f, name)                                               \
        *(void**)(&f) = dlsym(h, name);                            \
        if (!f) { err = std::string("dlsym ") + name; return false; }
        SYM(GetUniqueId, "ncclGetUniqueId") SYM(CommInitRank, "ncclCommInitRank") SYM(CommDestroy, "ncclCommDestroy")
        SYM(AllGather, "ncclAllGather") SYM(Broadcast, "ncclBroadcast") SYM(AllReduce, "ncclAllReduce") SYM(GroupStart, "ncclGroupStart") SYM(GroupEnd, "ncclGroupEnd")
        SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
        return true;
    }
} g_nccl;
} // namespace

extern "C" int mgb_nccl_unique_id(void* id128) {
    if (!id128) return MGB_EINVAL;
    std::string err;
    if (!g_nccl.load(err)) return MGB_ENCCL;
    ncclUniqueId id;
    if (g_nccl.GetUniqueId(&id) != 0) return MGB_ENCCL;
    memcpy(id128, &id, 128);
    return MGB_OK;
}

extern "C" int mgb_comm_init(mgb_ctx* ctx, int nranks, int rank, const void* id128) {
    if (!ctx || !id128 || nranks < 1 || rank < 0 || rank >= nranks) return MGB_EINVAL;
    std::string err;
    if (!g_nccl.load(err)) return ctx->fail(MGB_ENCCL, "%s", err.c_str());
    cudaSetDevice(ctx->device);
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    ncclComm_t comm = nullptr;
    const ncclResult_t r = g_nccl.CommInitRank(&comm, nranks, id, rank);
    if (r != 0) return ctx->fail(MGB_ENCCL, "ncclCommInitRank: %s", g_nccl.GetErrorString(r));
    if (ctx->comm) g_nccl.CommDestroy((ncclComm_t)ctx->comm);
    ctx->comm = comm; ctx->nranks = nranks; ctx->rank = rank;
    return MGB_OK;
}

// mgb_destroy calls this; also exported so that a host can drop the communicator and keep the context.
extern "C" int mgb_comm_destroy(mgb_ctx* ctx) {
    if (!ctx) return MGB_EINVAL;
    if (ctx->comm && g_nccl.h) {
        cudaSetDevice(ctx->device);
        g_nccl.CommDestroy((ncclComm_t)ctx->comm);
    }
    ctx->comm = nullptr; ctx->nranks = 1; ctx->rank = 0;
    ctx->part_bounds.clear();
    return MGB_OK;
}

extern "C" int mgb_comm_agree(mgb_ctx* ctx, int status, int* min_status) {
    if (!ctx || !min_status) return MGB_EINVAL;
    *min_status = status;
    if (ctx->nranks == 1) return MGB_OK;
    if (!ctx->comm) return ctx->fail(MGB_ESTATE, "mgb_comm_agree: mgb_comm_init was not called");
    cudaSetDevice(ctx->device);
    MGB_CUDA(ctx, ctx->counters.ensure(64));
    int* d = reinterpret_cast<int*>(ctx->counters.as<unsigned long long>() + 7);
    MGB_CUDA(ctx, cudaMemcpyAsync(d, &status, 4, cudaMemcpyHostToDevice, ctx->stream));
    const ncclResult_t r = g_nccl.AllReduce(d, d, 1, /* ncclInt32 */ 2, /* ncclMin */ 3, (ncclComm_t)ctx->comm, ctx->stream);
    if (r) return ctx->fail(MGB_ENCCL, "ncclAllReduce: %s", g_nccl.GetErrorString(r));
    MGB_CUDA(ctx, cudaMemcpyAsync(min_status, d, 4, cudaMemcpyDeviceToHost, ctx->stream));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return MGB_OK;
}

static uint32_t rows_per_rank(uint32_t npoints, int nranks) { return (npoints + (uint32_t)nranks - 1) / (uint32_t)nranks; }

extern "C" int mgb_partition_rows(uint32_t npoints, int nranks, int rank, uint32_t* o_begin, uint32_t* o_end, uint32_t* rows_per_rank_out) {
    if (!o_begin || !o_end || nranks < 1 || rank < 0 || rank >= nranks) return MGB_EINVAL;
    const uint32_t N = npoints, per = rows_per_rank(N, nranks);
    const uint64_t b = (uint64_t)per * rank, e = b + per;
    *o_begin = (uint32_t)(b < N ? b : N);
    *o_end   = (uint32_t)(e < N ? e : N);
    if (rows_per_rank_out) *rows_per_rank_out = per;
    return MGB_OK;
}

extern "C" int mgb_partition(mgb_ctx* ctx, int nranks, int rank, uint32_t* o_begin, uint32_t* o_end) {
    if (!ctx || !ctx->have_geometry || nranks < 1) return MGB_EINVAL;
    ctx->part_bounds.assign(nranks + 1, ctx->m.npoints);
    for (int r = 0; r < nranks; r++) {
        uint32_t a, b;
        mgb_partition_rows(ctx->m.npoints, nranks, r, &a, &b, nullptr);
        ctx->part_bounds[r] = a;
    }
    return mgb_partition_rows(ctx->m.npoints, nranks, rank, o_begin, o_end, nullptr);
}

extern "C" int mgb_allgather_results(mgb_ctx* ctx) {
    if (ctx && ctx->params.n_off_diag) return ctx->fail(MGB_EUNSUPPORTED, "mgb_allgather_results: the banded Lambda lists (n_off_diag > 0) are gathered by the host (mgb_get_lambda_lists per rank)");
    if (!ctx) return MGB_EINVAL;
    if (ctx->nranks == 1) return MGB_OK;
    if (!ctx->comm) return ctx->fail(MGB_ESTATE, "mgb_allgather_results: mgb_comm_init was not called");
    cudaSetDevice(ctx->device);
    /* Every rank must enter the collective: a rank without results of its own (empty origin range) still owns the buffers. A rank whose
       run FAILED must not come here alone -- callers agree on the error first (the host Model reduces the status over the ranks). */
    if (!ctx->J.p) { const int rc = mgb_internal_alloc_results(ctx); if (rc != MGB_OK) return rc; }
    const mgb::DevModel& m = ctx->m;
    if ((int)ctx->part_bounds.size() != ctx->nranks + 1)
        return ctx->fail(MGB_ESTATE, "mgb_allgather_results: call mgb_partition or mgb_partition_balanced with nranks = %d first", ctx->nranks);
    ncclComm_t comm = (ncclComm_t)ctx->comm;
    mgb::NvtxRange nvtx("mgb_allgather_results (NCCL)");
    // all-gather with per-rank row counts, in place: every rank's slice already sits at its global rows, one broadcast per owner
    ncclResult_t r = g_nccl.GroupStart();
    char* J = (char*)ctx->J.p; char* lv = (char*)ctx->lam_val.p; char* lc = (char*)ctx->lam_cnt.p; char* jl = (char*)ctx->J_lines.p;
    for (int root = 0; root < ctx->nranks && !r; root++) {
        const size_t o0 = ctx->part_bounds[root], rows = ctx->part_bounds[root + 1] - o0;
        if (!rows) continue;
        const size_t nJ = rows * m.nfreqs, nL = rows * m.nlines;
        r = g_nccl.Broadcast(J + 8 * o0 * m.nfreqs, J + 8 * o0 * m.nfreqs, nJ, ncclFloat64, root, comm, ctx->stream);
        if (!r) r = g_nccl.Broadcast(lv + 8 * o0 * m.nlines, lv + 8 * o0 * m.nlines, nL, ncclFloat64, root, comm, ctx->stream);
        if (!r) r = g_nccl.Broadcast(jl + 8 * o0 * m.nlines, jl + 8 * o0 * m.nlines, nL, ncclFloat64, root, comm, ctx->stream);
        if (!r) r = g_nccl.Broadcast(lc + 4 * o0 * m.nlines, lc + 4 * o0 * m.nlines, nL, ncclUint32, root, comm, ctx->stream);
    }
    const ncclResult_t r2 = g_nccl.GroupEnd();
    if (r || r2) return ctx->fail(MGB_ENCCL, "ncclBroadcast (all-gather of origin slices): %s", g_nccl.GetErrorString(r ? r : r2));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->res_o0 = 0; ctx->res_o1 = m.npoints; ctx->res_valid = true;
    return MGB_OK;
}
