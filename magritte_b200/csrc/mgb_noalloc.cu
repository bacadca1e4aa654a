// mgb_noalloc.cu -- mgb_set_column_noalloc: Model::set_column without any device allocation.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc (code under
// CUDA graph capture, tight serving loops, harnesses that forbid it). mgb_set_column (mgb_api.cu) stages the density and the column
// block through two device buffers it allocates and frees per call; this variant stages them through ONE file-scope device array
// of fixed size, so the call allocates nothing on the device: no cudaMalloc / cudaFree, no allocating library routine, no events.
//
//   * g_column_scratch: kNoAllocMaxDoubles doubles (32 MiB) of static device memory. Every GPU that loads the library holds it until
//     the process exits. Layout of one call: density [npoints] at offset 0, column [nrays][o_end-o_begin] after it (256-byte aligned).
//     A call whose npoints + nrays * (o_end - o_begin) (+ alignment padding) exceeds the array is rejected with MGB_EINVAL.
//   * In host code `g_column_scratch` is a host-side stand-in for the symbol, not a device address: the device address is fetched
//     with cudaGetSymbolAddress once per device (that device current) and cached; kernels and copies only ever see that pointer.
//   * Every call shares the array and finds the previous call's contents in it. The density block is overwritten by the upload and
//     every element of the column block in use is written by the kernel (one store per half-ray); the column block is nevertheless
//     cleared first, so a half-ray the kernel did not reach could never return a stale value.
//   * Calls that overlap on the GPU (two contexts or two host threads on the same device) RACE on the array: serialise them.
#include <mutex>

#include "mgb_ctx.h"
#include "mgb_device.cuh"

using namespace mgb;

namespace {

constexpr size_t kNoAllocMaxDoubles = (size_t)4 << 20; // 4 Mi doubles = 32 MiB per GPU
static __device__ __align__(256) double g_column_scratch[kNoAllocMaxDoubles];

constexpr int kMaxDevices = 64;
std::mutex g_scratch_mutex;
double* g_scratch_ptr[kMaxDevices] = {nullptr}; // device address of g_column_scratch on each device, fetched once

// Device address of the scratch array on `device` (which must be current). cudaGetSymbolAddress allocates nothing: the array is
// part of the module image.
cudaError_t scratch_address(int device, double** out) {
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_ptr[device]) {
        double* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_column_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_scratch_ptr[device] = p;
    }
    *out = g_scratch_ptr[device];
    return cudaSuccess;
}

// k_column (mgb_kernels.cuh) on explicit pointers into the static array: Solver::get_column (solver.tpp:2968-2991), one warp per
// half-ray (o, r), lane 0 keeps the trapezoid sum in the reference's order (outward from the origin). column is [nrays][no].
__global__ void __launch_bounds__(128) k_column_static(DevModel m, uint32_t o0, uint32_t no, const double* __restrict__ density,
                                                       double* __restrict__ column) {
    const unsigned lane = threadIdx.x & 31;
    const size_t w      = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= (size_t)no * m.nrays) return;
    const uint32_t ol = (uint32_t)(w / m.nrays), r = (uint32_t)(w % m.nrays);
    const uint32_t o  = o0 + ol;
    const RayFrame fr = make_frame(m, o, r);
    double Z = 0.0, dZ = 0.0, col = 0.0;
    uint32_t ns;
    uint32_t nxt = warp_get_next(m, fr, o, Z, dZ, ns, lane);
    if (nxt < m.npoints) {
        double rho_crt = density[o];
        for (;;) {
            const double rho_nxt = density[nxt];
            col = __dadd_rn(col, __dmul_rn(__dmul_rn(0.5, __dadd_rn(rho_crt, rho_nxt)), dZ));
            if (__ldg(&m.p2b[nxt]) != m.npoints) break; /* reached the boundary */
            rho_crt = rho_nxt;
            nxt     = warp_get_next(m, fr, nxt, Z, dZ, ns, lane);
            if (!(nxt < m.npoints)) break; /* dead end, as in k_column */
        }
    }
    if (lane == 0) column[(size_t)r * no + ol] = col;
}

} // namespace

extern "C" int mgb_set_column_noalloc(mgb_ctx* ctx, const double* density, uint32_t o_begin, uint32_t o_end, double* column) {
    if (!ctx || !density || !column) return MGB_EINVAL;
    if (!ctx->have_geometry) return ctx->fail(MGB_ESTATE, "mgb_set_column_noalloc before mgb_set_geometry");
    const DevModel& m = ctx->m;
    if (o_begin >= o_end || o_end > m.npoints) return ctx->fail(MGB_EINVAL, "mgb_set_column_noalloc: bad origin range [%u,%u)", o_begin, o_end);
    const uint32_t no  = o_end - o_begin;
    const size_t nw    = (size_t)no * m.nrays;
    const size_t off_c = ((size_t)m.npoints + 31) / 32 * 32; /* column block starts 256-byte aligned */
    if (off_c + nw > kNoAllocMaxDoubles)
        return ctx->fail(MGB_EINVAL,
                         "mgb_set_column_noalloc: npoints (%u, padded to %zu) + nrays * origins (%u * %u) = %zu doubles exceed the static scratch of %zu "
                         "doubles (32 MiB); split the origin range or use mgb_set_column",
                         m.npoints, off_c, m.nrays, no, off_c + nw, kNoAllocMaxDoubles);
    MGB_CUDA(ctx, cudaSetDevice(ctx->device));
    double* scratch = nullptr;
    MGB_CUDA(ctx, scratch_address(ctx->device, &scratch));
    double* rho = scratch;
    double* col = scratch + off_c;
    MGB_CUDA(ctx, cudaMemcpyAsync(rho, density, 8 * (size_t)m.npoints, cudaMemcpyHostToDevice, ctx->stream));
    MGB_CUDA(ctx, cudaMemsetAsync(col, 0, 8 * nw, ctx->stream)); /* the array keeps the previous call's contents */
    k_column_static<<<(unsigned)((nw * 32 + 127) / 128), 128, 0, ctx->stream>>>(m, o_begin, no, rho, col);
    MGB_CUDA(ctx, cudaGetLastError());
    MGB_CUDA(ctx, cudaMemcpyAsync(column, col, 8 * nw, cudaMemcpyDeviceToHost, ctx->stream));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return MGB_OK;
}
