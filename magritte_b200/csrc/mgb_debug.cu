// mgb_debug.cu -- test hooks of the C ABI: ray dump, device-libm self-test, FP64 peak probe.
#include "mgb_ctx.h"

using namespace mgb;

#define MGB_TRY(x)            \
    do {                      \
        int rc_ = (x);        \
        if (rc_ != MGB_OK) return rc_; \
    } while (0)

namespace {
#ifndef MGB_NO_TEST_HOOKS

// One warp traces half-ray (o, r) and expands every geometric step into the sub-steps of line l
// (l == nlines: the all-lines criterion of Solver::trace_ray / InterpHelper::get_n_interp,
// solver_interp_helper.tpp:10-46). Entries are emitted nearest-first, as Solver::set_data_for_line
// generates them (solver.tpp:1057-1097).
__global__ void k_trace_dump(DevModel m, uint32_t o, uint32_t r, uint32_t l, uint32_t cap, uint32_t* nr, uint32_t* nri, double* fac,
                             double* sh, double* dz, uint32_t* n_out) {
    const unsigned lane = threadIdx.x & 31;
    stage_exp_table();
    const RayFrame fr   = make_frame(m, o, r);
    double Z = 0.0, dZ = 0.0;
    uint32_t ns, n_emit = 0;
    uint32_t crt = o;
    double shift_crt = get_shift_comoving(m, fr, o, 0.0);
    uint32_t nxt = warp_get_next(m, fr, o, Z, dZ, ns, lane);
    while (nxt < m.npoints) {
        const double shift_nxt = get_shift_comoving(m, fr, nxt, Z);
        uint32_t n;
        if (l < m.nlines) {
            n = n_interp_from_criteria(m, m.lrow[(size_t)crt * m.nlines + l].criterion, m.lrow[(size_t)nxt * m.nlines + l].criterion);
        } else {
            double mx = 0.0;
            for (uint32_t k = 0; k < m.nlines; k++) {
                const double x = m.lrow[(size_t)nxt * m.nlines + k].criterion, y = m.lrow[(size_t)crt * m.nlines + k].criterion;
                const double d = fabs((double)mgb_logf((float)(x / y)));
                if (k == 0 || mx < d) mx = d;
            }
            n = to_size(ceil(mx / m.log_max_interp_diff));
            n = n > 1 ? n : 1;
        }
        if (lane == 0) {
            const double dshift = shift_nxt - shift_crt;
            for (uint32_t i = 1; i <= n; i++) {
                if (n_emit < cap) {
                    nr[n_emit]  = nxt;
                    nri[n_emit] = crt;
                    fac[n_emit] = (n > 1) ? 1.0 - ((double)i) / n : 0.0;
                    sh[n_emit]  = (n > 1) ? shift_crt + i * (dshift / n) : shift_nxt;
                    dz[n_emit]  = (n > 1) ? dZ / n : dZ;
                }
                n_emit++;
            }
        } else {
            n_emit += n;
        }
        if (m.p2b[nxt] != m.npoints) break;
        crt = nxt; shift_crt = shift_nxt;
        nxt = warp_get_next(m, fr, crt, Z, dZ, ns, lane);
    }
    if (lane == 0) *n_out = n_emit;
}

__global__ void k_libm(int fn, unsigned long long n, const float* x, const float* y, float* out) {
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    stage_exp_table();
    if (i >= n) return;
    float r;
    switch (fn) {
    case 0: r = mgb_expf(x[i]); break;
    case 1: r = mgb_logf(x[i]); break;
    case 2: r = mgb_erff(x[i]); break;
    case 3: r = mgb_expm1f(x[i]); break;
    default: r = mgb_powf(x[i], y[i]); break;
    }
    out[i] = r;
}

__global__ void k_ddiv(unsigned long long n, const double* a, const double* b, double* out) {
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = (a[i] == 1.0 && !signbit(a[i])) ? drcp(b[i]) : ddiv(a[i], b[i]); /* 1/b call sites use drcp */
    /* ddiv2 (two numerators, one reciprocal refinement) must give ddiv's bits for both; a mismatch poisons the output */
    const unsigned long long j = (i ^ 1ull) < n ? (i ^ 1ull) : i;
    const Div2 d = ddiv2(a[i], a[j], b[i]);
    const double w1 = ddiv_scaled(a[i], b[i]), w2 = ddiv_scaled(a[j], b[i]);
    const bool same1 = __double_as_longlong(d.q1) == __double_as_longlong(w1) || (d.q1 != d.q1 && w1 != w1);
    const bool same2 = __double_as_longlong(d.q2) == __double_as_longlong(w2) || (d.q2 != d.q2 && w2 != w2);
    if (!(same1 && same2)) out[i] = __longlong_as_double(0x4242424242424242ll);
    /* ddiv_raw (k_solve2's inline division on the unscaled operands) must give ddiv's bits too */
    const double w3 = ddiv_raw(a[i], b[i]);
    if (!(__double_as_longlong(w3) == __double_as_longlong(w1) || (w3 != w3 && w1 != w1))) out[i] = __longlong_as_double(0x4343434343434343ll);
}

#endif // MGB_NO_TEST_HOOKS
// 8 independent DFMA chains per thread; explicit fma() because the file is compiled with -fmad=false.
__global__ void k_dfma(double* out, int iters) {
    double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double b = 1.0000001, c = 1e-7;
    for (int i = 0; i < iters; i++) {
        a0 = fma(a0, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c);
        a4 = fma(a4, b, c); a5 = fma(a5, b, c); a6 = fma(a6, b, c); a7 = fma(a7, b, c);
    }
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

} // namespace

extern "C" int mgb_trace_dump(mgb_ctx* ctx, uint32_t o, uint32_t rr, uint32_t l, uint32_t cap, uint32_t* nr, uint32_t* nr_interp,
                              double* interp_factor, double* shift, double* dZ, uint32_t* n_out, uint32_t* centre_out) {
#ifdef MGB_NO_TEST_HOOKS
    return ctx ? ctx->fail(MGB_EUNSUPPORTED, "mgb_trace_dump: this build has no test hooks (MGB_NO_TEST_HOOKS)") : MGB_EINVAL;
#else
    if (!ctx || !nr || !nr_interp || !interp_factor || !shift || !dZ || !n_out || !centre_out) return MGB_EINVAL;
    if (!(ctx->have_geometry && ctx->have_lines && ctx->have_state)) return ctx->fail(MGB_ESTATE, "mgb_trace_dump: model incomplete");
    const DevModel& m = ctx->m;
    if (o >= m.npoints || rr >= m.hnrays || l > m.nlines) return ctx->fail(MGB_EINVAL, "mgb_trace_dump: index out of range");
    cudaSetDevice(ctx->device);
    std::vector<uint32_t> ar(m.nrays);
    MGB_CUDA(ctx, cudaMemcpy(ar.data(), m.antipod, 4 * (size_t)m.nrays, cudaMemcpyDeviceToHost));
    DevBuf b_nr, b_nri, b_fac, b_sh, b_dz, b_n;
    std::vector<uint32_t> h_nr[2], h_nri[2];
    std::vector<double> h_fac[2], h_sh[2], h_dz[2];
    uint32_t cnt[2];
    for (int side = 0; side < 2; side++) {
        MGB_CUDA(ctx, b_nr.ensure(4 * (size_t)cap)); MGB_CUDA(ctx, b_nri.ensure(4 * (size_t)cap));
        MGB_CUDA(ctx, b_fac.ensure(8 * (size_t)cap)); MGB_CUDA(ctx, b_sh.ensure(8 * (size_t)cap));
        MGB_CUDA(ctx, b_dz.ensure(8 * (size_t)cap)); MGB_CUDA(ctx, b_n.ensure(4));
        k_trace_dump<<<1, 32, 0, ctx->stream>>>(m, o, side ? ar[rr] : rr, l, cap, b_nr.as<uint32_t>(), b_nri.as<uint32_t>(),
                                                b_fac.as<double>(), b_sh.as<double>(), b_dz.as<double>(), b_n.as<uint32_t>());
        MGB_CUDA(ctx, cudaGetLastError());
        MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        MGB_CUDA(ctx, cudaMemcpy(&cnt[side], b_n.p, 4, cudaMemcpyDeviceToHost));
        if (cnt[side] > cap) return ctx->fail(MGB_EINVAL, "mgb_trace_dump: half-ray has %u entries, cap = %u", cnt[side], cap);
        const uint32_t k = cnt[side];
        h_nr[side].resize(k); h_nri[side].resize(k); h_fac[side].resize(k); h_sh[side].resize(k); h_dz[side].resize(k);
        if (k) {
            MGB_CUDA(ctx, cudaMemcpy(h_nr[side].data(), b_nr.p, 4 * (size_t)k, cudaMemcpyDeviceToHost));
            MGB_CUDA(ctx, cudaMemcpy(h_nri[side].data(), b_nri.p, 4 * (size_t)k, cudaMemcpyDeviceToHost));
            MGB_CUDA(ctx, cudaMemcpy(h_fac[side].data(), b_fac.p, 8 * (size_t)k, cudaMemcpyDeviceToHost));
            MGB_CUDA(ctx, cudaMemcpy(h_sh[side].data(), b_sh.p, 8 * (size_t)k, cudaMemcpyDeviceToHost));
            MGB_CUDA(ctx, cudaMemcpy(h_dz[side].data(), b_dz.p, 8 * (size_t)k, cudaMemcpyDeviceToHost));
        }
    }
    const uint32_t n_tot = cnt[0] + cnt[1] + 1;
    if (n_tot > cap) return ctx->fail(MGB_EINVAL, "mgb_trace_dump: ray has %u entries, cap = %u", n_tot, cap);
    uint32_t w = 0;
    for (uint32_t i = cnt[0]; i-- > 0;) { /* backward half: dZ[idx] is the entry's own sub-step (id2 == id1, solver.tpp:490) */
        nr[w] = h_nr[0][i]; nr_interp[w] = h_nri[0][i]; interp_factor[w] = h_fac[0][i]; shift[w] = h_sh[0][i]; dZ[w] = h_dz[0][i];
        w++;
    }
    nr[w] = o; nr_interp[w] = o; interp_factor[w] = 0.0; shift[w] = 1.0;
    *centre_out = w;
    w++;
    for (uint32_t i = 0; i < cnt[1]; i++) { /* forward half: dZ[idx-1] (id2 == id1 - 1, solver.tpp:493) */
        nr[w] = h_nr[1][i]; nr_interp[w] = h_nri[1][i]; interp_factor[w] = h_fac[1][i]; shift[w] = h_sh[1][i]; dZ[w - 1] = h_dz[1][i];
        w++;
    }
    *n_out = n_tot;
    return MGB_OK;
#endif
}

extern "C" int mgb_selftest_libm(mgb_ctx* ctx, int fn, uint64_t n, const float* x, const float* y, float* out) {
#ifdef MGB_NO_TEST_HOOKS
    return ctx ? ctx->fail(MGB_EUNSUPPORTED, "mgb_selftest_libm: this build has no test hooks (MGB_NO_TEST_HOOKS)") : MGB_EINVAL;
#else
    if (!ctx || !x || !out || fn < 0 || fn > 4 || (fn == 4 && !y)) return MGB_EINVAL;
    cudaSetDevice(ctx->device);
    DevBuf dx, dy, dout;
    MGB_CUDA(ctx, dx.ensure(4 * n)); MGB_CUDA(ctx, dout.ensure(4 * n));
    MGB_CUDA(ctx, cudaMemcpy(dx.p, x, 4 * n, cudaMemcpyHostToDevice));
    if (fn == 4) { MGB_CUDA(ctx, dy.ensure(4 * n)); MGB_CUDA(ctx, cudaMemcpy(dy.p, y, 4 * n, cudaMemcpyHostToDevice)); }
    k_libm<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(fn, n, dx.as<float>(), dy.as<float>(), dout.as<float>());
    MGB_CUDA(ctx, cudaGetLastError());
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    MGB_CUDA(ctx, cudaMemcpy(out, dout.p, 4 * n, cudaMemcpyDeviceToHost));
    return MGB_OK;
#endif
}

extern "C" int mgb_measure_fp64_peak(mgb_ctx* ctx, double* tflops) {
    if (!ctx || !tflops) return MGB_EINVAL;
    cudaSetDevice(ctx->device);
    cudaDeviceProp prop;
    MGB_CUDA(ctx, cudaGetDeviceProperties(&prop, ctx->device));
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 14;
    DevBuf out;
    MGB_CUDA(ctx, out.ensure(8 * (size_t)blocks * threads));
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    double best = 0.0;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(a, ctx->stream);
        k_dfma<<<blocks, threads, 0, ctx->stream>>>(out.as<double>(), iters);
        cudaEventRecord(b, ctx->stream);
        cudaEventSynchronize(b);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, a, b);
        const double tf = 2.0 * 8.0 * (double)iters * blocks * threads / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(a); cudaEventDestroy(b);
    MGB_CUDA(ctx, cudaGetLastError());
    *tflops = best;
    return MGB_OK;
}

extern "C" int mgb_selftest_ddiv(mgb_ctx* ctx, uint64_t n, const double* a, const double* b, double* out) {
#ifdef MGB_NO_TEST_HOOKS
    return ctx ? ctx->fail(MGB_EUNSUPPORTED, "mgb_selftest_ddiv: this build has no test hooks (MGB_NO_TEST_HOOKS)") : MGB_EINVAL;
#else
    if (!ctx || !a || !b || !out) return MGB_EINVAL;
    cudaSetDevice(ctx->device);
    DevBuf da, db, dout;
    MGB_CUDA(ctx, da.ensure(8 * n)); MGB_CUDA(ctx, db.ensure(8 * n)); MGB_CUDA(ctx, dout.ensure(8 * n));
    MGB_CUDA(ctx, cudaMemcpy(da.p, a, 8 * n, cudaMemcpyHostToDevice));
    MGB_CUDA(ctx, cudaMemcpy(db.p, b, 8 * n, cudaMemcpyHostToDevice));
    k_ddiv<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(n, da.as<double>(), db.as<double>(), dout.as<double>());
    MGB_CUDA(ctx, cudaGetLastError());
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    MGB_CUDA(ctx, cudaMemcpy(out, dout.p, 8 * n, cudaMemcpyDeviceToHost));
    return MGB_OK;
#endif
}
