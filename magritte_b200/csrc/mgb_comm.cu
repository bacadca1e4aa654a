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
        SYM(AllGather, "ncclAllGather") SYM(GroupStart, "ncclGroupStart") SYM(GroupEnd, "ncclGroupEnd")
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
    ctx->comm = comm; ctx->nranks = nranks; ctx->rank = rank;
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

extern "C" int mgb_partition(const mgb_ctx* ctx, int nranks, int rank, uint32_t* o_begin, uint32_t* o_end) {
    if (!ctx || !ctx->have_geometry) return MGB_EINVAL;
    return mgb_partition_rows(ctx->m.npoints, nranks, rank, o_begin, o_end, nullptr);
}

extern "C" int mgb_allgather_results(mgb_ctx* ctx) {
    if (!ctx) return MGB_EINVAL;
    if (ctx->nranks == 1) return MGB_OK;
    if (!ctx->comm) return ctx->fail(MGB_ESTATE, "mgb_allgather_results: mgb_comm_init was not called");
    if (!ctx->J.p) return ctx->fail(MGB_ESTATE, "mgb_allgather_results: no results");
    cudaSetDevice(ctx->device);
    const mgb::DevModel& m = ctx->m;
    const uint32_t per = rows_per_rank(m.npoints, ctx->nranks);
    const size_t need_rows = (size_t)per * ctx->nranks;
    if (ctx->J.bytes < 8 * need_rows * m.nfreqs || ctx->lam_val.bytes < 8 * need_rows * m.nlines || ctx->lam_cnt.bytes < 4 * need_rows * m.nlines)
        return ctx->fail(MGB_ESTATE, "mgb_allgather_results: result buffers are not padded for %d ranks (run after mgb_comm_init)", ctx->nranks);
    ncclComm_t comm = (ncclComm_t)ctx->comm;
    // in place: every rank's slice already sits at rank*per rows of the full-size buffers
    ncclResult_t r = g_nccl.GroupStart();
    char* J = (char*)ctx->J.p; char* lv = (char*)ctx->lam_val.p; char* lc = (char*)ctx->lam_cnt.p;
    const size_t nJ = (size_t)per * m.nfreqs, nL = (size_t)per * m.nlines;
    if (!r) r = g_nccl.AllGather(J + 8 * nJ * ctx->rank, J, nJ, ncclFloat64, comm, ctx->stream);
    if (!r) r = g_nccl.AllGather(lv + 8 * nL * ctx->rank, lv, nL, ncclFloat64, comm, ctx->stream);
    if (!r) r = g_nccl.AllGather(lc + 4 * nL * ctx->rank, lc, nL, ncclUint32, comm, ctx->stream);
    const ncclResult_t r2 = g_nccl.GroupEnd();
    if (r || r2) return ctx->fail(MGB_ENCCL, "ncclAllGather: %s", g_nccl.GetErrorString(r ? r : r2));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->res_o0 = 0; ctx->res_o1 = m.npoints;
    return MGB_OK;
}
