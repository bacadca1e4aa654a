// mgb_shortchar.cuh -- k_shortchar: short characteristics, order 0 (reference Solver::solve_shortchar_order_0, solver.tpp:150-235, and
// solve_shortchar_order_0_ray_forward / _backward, :1564-1654 / :1664-1757).
//
// Same work decomposition, ray records and window logic as k_solve (mgb_solve.cuh): one warp = (origin, ray pair, <= 32 frequency
// bins), phase A expands the traced path of each half ray into per-(entry, line) records in shared memory, phase B walks them from
// the boundary end towards the origin with lanes = frequency bins. What differs is the per-segment operation: instead of one row of
// the Feautrier elimination, the specific intensity is carried along the half ray,
//     I <- expf(-dtau) I + expm1f(-dtau) / dtau (S_c - S_n (1 + dtau)) + S_c - S_n          (solver.tpp:1617, 1720)
// with (dtau, S_c, S_n) of compute_source_dtau (:2139-2186), dtau clamped from below by Parameters::min_dtau, float expf / expm1f
// (glibc's, bit for bit: mgb_libm.h). BOTH half rays are walked in the reference's orientation curr = far entry, next = near entry
// (k_solve's side 1 is mirrored; this kernel's is not). Outputs per (rr, o, f): u = (I_f + I_b) / 2 and the two factors
// expm1f(-dtau_last) / dtau_last of the local Lambda element, one per half ray (:1621, 1724); k_reduce_lambda_sc forms the elements.
// A half ray without a segment has no dtau: the reference reads an uninitialised variable there; here (and in the oracle) the factor
// is 0.
#pragma once
#include "mgb_solve.cuh"

namespace mgb {

struct ScState {
    double chiX, SX; // eta/chi evaluation at the far entry X of the open segment
    double I;        // intensity carried to X
    double dtau;     // (clamped) optical depth of the last closed segment
    bool haveX, any;
};

// One entry of the walk: evaluate chi, S at the record Y = q[0]; close the segment X = q[-1] -> Y and carry I across it.
template <int APPROX>
__device__ __forceinline__ void sc_step(const DevModel& m, const LaneLine& L, ScState& st, const double* q, uint32_t ll, double freq, double min_dtau,
                                        unsigned& nlarge) {
    const double shY = MGB_REC(q, R_SHIFT);
    const double frY = freq * shY;
    double chiY, SY;
    {
        int cls = WIN_OTHER;
        if (APPROX == APPROX_CLOSELINES) {
            const double ub = m.max_dist * (frY * MGB_REC(q, R_SQW));
            cls = window_class(L, frY - ub, frY + ub);
        } else if (APPROX == APPROX_ONELINE) {
            cls = L.own ? WIN_OWN : WIN_OTHER;
        }
        if (cls == WIN_OWN) {
            const double prof = gaussian(MGB_REC(q, R_IW), frY - L.s_own);
            const double eta = 0.0 + prof * MGB_REC(q, R_IE);
            if (APPROX == APPROX_ONELINE) chiY = (0.0 + prof * MGB_REC(q, R_IO)) + m.min_opacity;
            else chiY = m.min_opacity + prof * MGB_REC(q, R_IO);
            SY = ddiv(eta, chiY);
        } else if (cls == WIN_EMPTY) {
            chiY = m.min_opacity;
            SY   = zero_over(chiY);
        } else {
            const uint2 pp = *reinterpret_cast<const uint2*>(&MGB_REC(q, R_PTS));
            double eta;
            eta_chi_generic<APPROX>(m, pp.x, pp.y, MGB_REC(q, R_FACTOR), ll, frY, eta, chiY);
            SY = ddiv(eta, chiY);
        }
    }
    if (st.haveX) {
        const double* x = q - 1;
        const double shX = MGB_REC(x, R_SHIFT), dzX = MGB_REC(x, R_DZ);
        const double dshift_abs = fabs(shY - shX);
        const double dX = MGB_REC(x, R_DSM), dY = MGB_REC(q, R_DSM);
        const double dshift_max = (dY < dX) ? dY : dX;
        double dtau, S_c, S_n;
        if (dshift_abs > dshift_max) {
            nlarge++;
            const double frX = freq * shX;
            int cls = WIN_OTHER;
            if (APPROX == APPROX_CLOSELINES) {
                const double left_freq = (frX < frY) ? frX : frY, right_freq = (frX < frY) ? frY : frX;
                const double bX = m.max_dist * (right_freq * MGB_REC(x, R_SQW));
                const double bY = m.max_dist * (right_freq * MGB_REC(q, R_SQW));
                const double ub = (bX < bY) ? bY : bX;
                cls = window_class(L, left_freq - ub, right_freq + ub);
            } else if (APPROX == APPROX_ONELINE) {
                cls = L.own ? WIN_OWN : WIN_OTHER;
            }
            if (cls == WIN_OWN) {
                const LineIntegral li = line_integrated_own<APPROX>(m.min_opacity, L.s_own, frX, frY, dzX, MGB_REC(x, R_IW), MGB_REC(q, R_IW),
                                                                    MGB_REC(x, R_IO), MGB_REC(q, R_IO), MGB_REC(x, R_SLINE), MGB_REC(q, R_SLINE));
                dtau = li.dtau; S_c = li.S_atX; S_n = li.S_atY;
            } else if (cls == WIN_EMPTY) {
                double d = 0.0;
                const double bound_min_dtau = m.min_opacity * dzX;
                if (-bound_min_dtau < d) d = (bound_min_dtau < d) ? d : bound_min_dtau;
                dtau = d; S_c = zero_over(d); S_n = S_c;
            } else {
                const uint2 pX = *reinterpret_cast<const uint2*>(&MGB_REC(x, R_PTS)), pY = *reinterpret_cast<const uint2*>(&MGB_REC(q, R_PTS));
                S_dtau_generic<APPROX>(m, pX.x, pX.y, MGB_REC(x, R_FACTOR), pY.x, pY.y, MGB_REC(q, R_FACTOR), ll, frX, frY, dzX, dtau, S_c, S_n);
            }
        } else {
            dtau = 0.5 * (st.chiX + chiY) * dzX;
            S_c  = st.SX;
            S_n  = SY;
        }
        dtau = (min_dtau < dtau) ? dtau : min_dtau; /* std::max(min_dtau, dtau) */
        const double ex  = (double)expf_ool((float)(-dtau));
        const double em1 = (double)expm1f_ool((float)(-dtau));
        st.I    = ex * st.I + ddiv(em1, dtau) * (S_c - S_n * (1.0 + dtau)) + S_c - S_n;
        st.dtau = dtau;
        st.any  = true;
    } else {
        const uint2 pY = *reinterpret_cast<const uint2*>(&MGB_REC(q, R_PTS));
        st.I     = boundary_intensity(m, pY.x, frY); /* I = boundary_intensity(nr[first], freq * shift[first]) (solver.tpp:1586, 1687) */
        st.haveX = true;
    }
    st.chiX = chiY; st.SX = SY;
}

template <int APPROX>
__global__ void __launch_bounds__(32 * kSolveWarps, kSolveCtas / 2) k_shortchar(const __grid_constant__ DevModel m, const __grid_constant__ PathView pv,
                                                                                 ScratchView sc, const FreqChunk* __restrict__ chunks,
                                                                                 uint32_t nchunks_f, uint32_t lpw, double min_dtau) {
    const unsigned lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const size_t item   = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const size_t nitems = (size_t)pv.no * m.hnrays * nchunks_f;
    stage_exp_table();
    if (item >= nitems) return; /* whole warps only */
    WarpSmem ws;
    ws.lpw  = lpw;
    ws.base = wib * (uint32_t)(solve_smem_per_warp(lpw) / 8);
    const uint32_t cf = (uint32_t)(item % nchunks_f);
    const size_t t    = item / nchunks_f;
    const uint32_t rr = (uint32_t)(t % m.hnrays);
    const uint32_t ol = (uint32_t)(t / m.hnrays);
    const uint32_t o  = pv.o0 + ol;
    const FreqChunk ch = chunks[cf];
    const bool active  = lane < ch.nf;
    const uint32_t f   = ch.f0 + (active ? lane : ch.nf - 1);
    const uint32_t l   = f / m.nquads;
    const uint32_t z   = f - l * m.nquads;
    const uint32_t ks  = l - ch.l0;
    const uint32_t ll  = (APPROX == APPROX_ONELINE) ? __ldg(&m.corr_line[f]) : l;
    const uint32_t ar  = __ldg(&m.antipod[rr]);
    const double freq  = __ldg(&m.nu[(size_t)o * m.nfreqs + f]);
    const size_t out   = ((size_t)rr * pv.no + ol) * m.nfreqs + f;

    const uint32_t len_b = __ldg(&pv.len[(size_t)o * m.nrays + rr]);
    const uint32_t len_f = __ldg(&pv.len[(size_t)o * m.nrays + ar]);
    if (len_b + len_f == 0) {
        // n_tot == 1: the boundary intensity, no Lambda element (solver.tpp:213-224)
        if (active) { sc.u[out] = boundary_intensity(m, o, freq); sc.offV[out * 2] = 0.0; sc.offV[out * 2 + 1] = 0.0; }
        const unsigned e1 = __popc(__ballot_sync(0xffffffffu, active && z == 0));
        if (lane == 0 && e1) atomicAdd(&sc.counters[0], (unsigned long long)e1);
        return;
    }

    LaneLine L;
    {
        const uint32_t pos = __ldg(&m.sorted_pos[l]);
        L.s_own  = __ldg(&m.line[l]);
        L.s_prev = pos ? __ldg(&m.sorted_line[pos - 1]) : -CUDART_INF;
        L.s_next = (pos + 1 < m.nlines) ? __ldg(&m.sorted_line[pos + 1]) : CUDART_INF;
        L.own    = (APPROX == APPROX_CLOSELINES) || (APPROX == APPROX_ONELINE && ll == l);
    }

    unsigned nlarge = 0, nentries = 1;
    double I_side[2], fac_side[2];
    const double* rb   = ws.rec(ks) + 1;
    const uint32_t* eo = ws.eoff(ks);
    for (int side = 0; side < 2; side++) {
        const uint32_t r   = side ? ar : rr;
        const uint32_t len = side ? len_f : len_b;
        if (len == 0) {
            /* the origin is this half ray's end: I = boundary_intensity(nr[centre], freq * 1), no segment */
            I_side[side]   = boundary_intensity(m, o, freq * 1.0);
            fac_side[side] = 0.0;
            continue;
        }
        const size_t hr = (size_t)ol * m.nrays + r;
        const unsigned long long off = __ldg(&pv.poff[hr]);
        const double shift0 = __ldg(&pv.shift0[hr]);
        ScState st;
        st.haveX = false; st.any = false;
        st.chiX = st.SX = st.I = st.dtau = 0.0;
        __syncwarp();
        if (lane < ch.nl) { *ws.isub(lane) = 0; *ws.last(lane) = 0; }
        __syncwarp();
        uint32_t j = len;
        for (;;) {
            uint32_t S = 1;
            const bool centre = (j == 0);
            carry_last_records(ws, ch.nl, lane);
            if (!centre) S = produce_window(m, pv, ws, ch.l0, ch.nl, o, off, shift0, j, lane);
            else produce_centre(m, ws, ch.l0, ch.nl, o, lane);
            __syncwarp();
            if (!centre) nentries += eo[S];
            for (uint32_t s = 0; s < S; s++) {
                const uint32_t e0 = eo[s], ns = eo[s + 1] - e0;
                for (uint32_t it = 0; it < ns; it++) sc_step<APPROX>(m, L, st, rb + e0 + it, ll, freq, min_dtau, nlarge);
            }
            __syncwarp();
            if (centre) break;
        }
        I_side[side]   = st.I;
        fac_side[side] = st.any ? ddiv((double)expm1f_ool((float)(-st.dtau)), st.dtau) : 0.0;
    }
    if (active) {
        sc.u[out] = 0.5 * (I_side[0] + I_side[1]);
        sc.offV[out * 2]     = fac_side[0];
        sc.offV[out * 2 + 1] = fac_side[1];
    }
    unsigned e = (active && z == 0) ? nentries : 0u;
    unsigned g = active ? nlarge : 0u;
#pragma unroll
    for (int off = 16; off; off >>= 1) {
        e += __shfl_xor_sync(0xffffffffu, e, off);
        g += __shfl_xor_sync(0xffffffffu, g, off);
    }
    if (lane == 0) {
        atomicAdd(&sc.counters[0], (unsigned long long)e);
        if (g) atomicAdd(&sc.counters[1], (unsigned long long)g);
    }
}

// Lambda(o, lid) of the short-characteristics solver: per (rr, z) the two half-ray elements
//   L = constante * freq * phi * (factor + 1.0) * inverse_chi,  constante = A_k * w_z * weight(o, rr)   (the ray's weight, not twice it)
// (solver.tpp:1628-1651, 1731-1754), forward ray first, accumulated in the order Lambda::add_element saw them (lambda.tpp:79-89):
// rr outermost, then z. chi(o, freq_line) and phi(o, f) depend on the origin only. One thread per (o, line).
template <int APPROX>
__global__ void __launch_bounds__(128) k_reduce_lambda_sc(const __grid_constant__ DevModel m, uint32_t o0, uint32_t no, const uint32_t* __restrict__ lens,
                                                          const double* __restrict__ fac2 /* [hnrays][no][nfreqs][2] */, double* __restrict__ lam_val,
                                                          uint32_t* __restrict__ lam_cnt) {
    stage_exp_table();
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)no * m.nlines) return;
    const uint32_t ol = (uint32_t)(i / m.nlines), l = (uint32_t)(i % m.nlines);
    const uint32_t o  = o0 + ol;
    const double freq_line = __ldg(&m.line[l]);
    double eta_l, chi_l;
    // NOTE: the reference passes the within-species transition index k as OneLine's line index here (solver.tpp:1638, 1741)
    eta_chi_generic<APPROX>(m, o, o, 0.0, __ldg(&m.line_k[l]), freq_line, eta_l, chi_l);
    const double inverse_chi = 1.0 / chi_l;
    const double invmass = __ldg(&m.line_invmass[l]);
    const double A_l = __ldg(&m.line_A[l]);
    double acc = 0.0;
    uint32_t cnt = 0;
    for (uint32_t rr = 0; rr < m.hnrays; rr++) {
        const uint32_t ar = __ldg(&m.antipod[rr]);
        if (lens[(size_t)o * m.nrays + rr] + lens[(size_t)o * m.nrays + ar] == 0) continue; /* n_tot <= 1: no add_element */
        const double w_ang = ray_weight(m, o, rr);
        for (uint32_t z = 0; z < m.nquads; z++) {
            const uint32_t f = l * m.nquads + z;
            const double freq = __ldg(&m.nu[(size_t)o * m.nfreqs + f]);
            const double phi  = thermo_profile(m, invmass, o, freq_line, freq);
            const double constante = A_l * __ldg(&m.line_quadw[(size_t)l * m.nquads + z]) * w_ang;
            const size_t at = (((size_t)rr * no + ol) * m.nfreqs + f) * 2;
            const double vf = constante * freq * phi * (fac2[at] + 1.0) * inverse_chi;
            const double vb = constante * freq * phi * (fac2[at + 1] + 1.0) * inverse_chi;
            if (cnt == 0) acc = vf; else acc += vf;
            acc += vb;
            cnt = 1;
        }
    }
    lam_val[(size_t)o * m.nlines + l] = acc;
    lam_cnt[(size_t)o * m.nlines + l] = cnt;
}

} // namespace mgb
