// mgb_solve3.cuh -- k_solve3: the hot kernel for the diagonal Lambda operator (n_off_diag = 0), third generation.
//
// Same algorithm, same ARITHMETIC and same work items as k_solve (mgb_solve.cuh: one warp = (origin, ray pair, <= 32 frequency bins),
// phase A records per (entry, line), phase B lanes = bins, two-sided elimination meeting at the origin): every result bit is k_solve's,
// which tests/test_gpu_parity.py pins to the oracle. What changed is the instruction stream, from the two ncu comparisons under
// profiles/ (ncu_k_solve_r1g_c3.txt, ncu_k_solve2_vs_k_solve_r2.txt):
//
//   * the common step is straight-line code with the Gaussian's expf and the IEEE division / reciprocal inline on the raw operands
//     (k_solve2's step, 172 warp instructions per bin against ~330 with k_solve's calls), for ONE bin per lane: k_solve2's two bins
//     per lane cost a third of the resident warps (168 registers, 19.6 KB of records per warp) and lost more than the ILP gave;
//   * the large-shift test of a segment (|dshift| > dshift_max, compute_source_dtau solver.tpp:2150-2153) does not depend on the bin:
//     phase A leaves it as the sign of the pre-scaled inverse width inverse_width / sqrt(pi), which takes the record slot of dshift_max
//     (still 10 fields: 8 CTAs of 2 warps per SM);
//   * the line-integrated branch (two float erff per segment and bin, solver.tpp:2374-2455) ran at 12 of 32 lanes because the lanes of
//     a warp fall into different |x| ranges of fdlibm's erff: it now uses the range-free evaluation mgb_erff_main (mgb_libm.h; one
//     Horner pass over a per-range coefficient row from shared memory, both arguments of the difference interleaved), bit-identical
//     to the branchy routine over all 2^32 inputs (tests/test_libm_host.py);
//   * lanes walk their own line's entries linearly (no per-step max-reduction and offset loads).
// The banded operator (n_off_diag > 0) stays with k_solve<APPROX, true>.
#pragma once
#include "mgb_solve2.cuh"

namespace mgb {

enum { R3_IWN = R_DSM }; // inverse_width / sqrt(pi), negative when the segment (previous entry -> this entry) takes the large-shift branch

// dshift_max of the last entry of the previous round (its record slot now holds R3_IWN): the two spare words of WarpSmem's line block
__device__ __forceinline__ double* dsm_last(const WarpSmem& ws, uint32_t k) { return reinterpret_cast<double*>(ws.words(k) + 102); }

// After make_record filled slots 1 .. total of line k: replace every slot's dshift_max by the flagged, pre-scaled inverse width.
// Reads of all lanes come first (lane e reads the dshift_max of slots e and e + 1), then the writes.
__device__ __forceinline__ void flag_large3(const WarpSmem& ws, uint32_t k, uint32_t total, unsigned lane) {
    double dY = 0.0, dX = 0.0, dsh = 0.0, iw = 0.0;
    double* y = ws.rec(k) + 1 + lane;
    if (lane < total) {
        const double* x = y - 1;
        dY  = MGB_REC(y, R_DSM);
        dX  = lane ? MGB_REC(x, R_DSM) : *dsm_last(ws, k);
        dsh = fabs(MGB_REC(y, R_SHIFT) - MGB_REC(x, R_SHIFT));
        iw  = MGB_REC(y, R_IW);
    }
    __syncwarp();
    if (lane < total) {
        const double dshift_max = (dY < dX) ? dY : dX; /* std::min(dshift_max[curr], dshift_max[next]) */
        const double iwn = iw * kINV_SQRT_PI;          /* gaussian(): inverse_width * INVERSE_SQRT_PI * expf(..), left to right */
        MGB_REC(y, R3_IWN) = (dsh > dshift_max) ? -iwn : iwn;
        if (lane == total - 1) *dsm_last(ws, k) = dY;
    }
    __syncwarp();
}

// produce_window of mgb_solve.cuh followed by the flag pass; *ws.last(k) = entries of line k now in slots 1 .. last.
__device__ __forceinline__ void produce_window3(const DevModel& m, const PathView& pv, const WarpSmem& ws, uint32_t l0, uint32_t nl,
                                                uint32_t o, unsigned long long off, double shift0, uint32_t& j, unsigned lane) {
    const int js = (int)j - 1 - (int)lane;
    StepData sd;
    sd.nxt = o; sd.crt = o; sd.dZ = 0.0; sd.sh_nxt = 1.0; sd.sh_crt = 1.0; sd.dsm = 0.0; sd.sqw = 0.0;
    if (js >= 0) {
        sd.nxt    = __ldg(&pv.nr[off + js]);
        sd.crt    = js ? __ldg(&pv.nr[off + js - 1]) : o;
        sd.dZ     = __ldg(&pv.dZ[off + js]);
        sd.sh_nxt = __ldg(&pv.shift[off + js]);
        sd.sh_crt = js ? __ldg(&pv.shift[off + js - 1]) : shift0;
        sd.dsm    = __ldg(&m.dshift_max[sd.nxt]);
        sd.sqw    = __ldg(&m.sqrtw_max[sd.nxt]);
    }
    uint32_t S = j < 32u ? j : 32u;
    for (uint32_t k = 0; k < nl; k++) {
        uint32_t ni = 0;
        if (js >= 0) {
            const double crit_nxt = __ldg(&m.lrow[(size_t)sd.nxt * m.nlines + l0 + k].criterion);
            const double crit_crt = __ldg(&m.lrow[(size_t)sd.crt * m.nlines + l0 + k].criterion);
            ni = n_interp_from_criteria(m, crit_crt, crit_nxt);
        }
        uint32_t cum = ni;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, cum, d);
            if ((int)lane >= d) cum += v;
        }
        ws.ni(k)[lane]   = ni;
        ws.excl(k)[lane] = cum - ni;
        const uint32_t fits = __ballot_sync(0xffffffffu, cum <= 32u);
        const uint32_t Sk   = (fits == 0xffffffffu) ? 32u : (uint32_t)(__ffs(~fits) - 1);
        S = Sk < S ? Sk : S;
    }
    __syncwarp();
    if (S > 0) {
        for (uint32_t k = 0; k < nl; k++) {
            const uint32_t ni = (lane < S) ? ws.ni(k)[lane] : 0u, excl = ws.excl(k)[lane];
            const uint32_t total = __shfl_sync(0xffffffffu, excl + ni, S - 1);
            uint32_t* mark = ws.eoff(k); /* entry e -> the step that owns it: scatter step ids at their first entry, running maximum */
            mark[lane] = 0;
            __syncwarp();
            if (lane < S) mark[excl] = lane + 1;
            __syncwarp();
            uint32_t te1 = mark[lane];
            __syncwarp();
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, te1, d);
                if ((int)lane >= d && v > te1) te1 = v;
            }
            const int te = te1 ? (int)te1 - 1 : 0;
            StepData se;
            se.nxt = __shfl_sync(0xffffffffu, sd.nxt, te); se.crt = __shfl_sync(0xffffffffu, sd.crt, te);
            se.dZ = __shfl_sync(0xffffffffu, sd.dZ, te); se.sh_nxt = __shfl_sync(0xffffffffu, sd.sh_nxt, te);
            se.sh_crt = __shfl_sync(0xffffffffu, sd.sh_crt, te); se.dsm = __shfl_sync(0xffffffffu, sd.dsm, te);
            se.sqw = __shfl_sync(0xffffffffu, sd.sqw, te);
            const uint32_t ni_e = __shfl_sync(0xffffffffu, ni, te), excl_e = __shfl_sync(0xffffffffu, excl, te);
            if (lane < total) make_record(m, ws.rec(k) + 1 + lane, l0 + k, se, ni_e - (lane - excl_e), ni_e);
            if (lane == 0) *ws.last(k) = total;
            __syncwarp();
            flag_large3(ws, k, total, lane);
        }
        j -= S;
        return;
    }
    // a step with more than 32 sub-steps on some line: lanes = sub-entries of step j-1, 32 at a time per line
    StepData s0;
    s0.nxt = __shfl_sync(0xffffffffu, sd.nxt, 0); s0.crt = __shfl_sync(0xffffffffu, sd.crt, 0);
    s0.dZ = __shfl_sync(0xffffffffu, sd.dZ, 0); s0.sh_nxt = __shfl_sync(0xffffffffu, sd.sh_nxt, 0);
    s0.sh_crt = __shfl_sync(0xffffffffu, sd.sh_crt, 0); s0.dsm = __shfl_sync(0xffffffffu, sd.dsm, 0);
    s0.sqw = __shfl_sync(0xffffffffu, sd.sqw, 0);
    bool all_done = true;
    for (uint32_t k = 0; k < nl; k++) {
        const uint32_t ni = ws.ni(k)[0], is = *ws.isub(k);
        const uint32_t rem = ni - is, nemit = rem < 32u ? rem : 32u;
        __syncwarp();
        if (lane < nemit) make_record(m, ws.rec(k) + 1 + lane, l0 + k, s0, ni - (is + lane), ni);
        if (lane == 0) { *ws.isub(k) = is + nemit; *ws.last(k) = nemit; }
        __syncwarp();
        flag_large3(ws, k, nemit, lane);
        all_done = all_done && (is + nemit == ni);
    }
    if (all_done) {
        __syncwarp();
        if (lane < nl) *ws.isub(lane) = 0;
        j -= 1;
    }
}

// The origin's own entry (nr_[centre] = o, shift_[centre] = 1; solver.tpp:484-485) as the single record of the last round.
__device__ __forceinline__ void produce_centre3(const DevModel& m, const WarpSmem& ws, uint32_t l0, uint32_t nl, uint32_t o, unsigned lane) {
    if (lane < nl) {
        StepData sd;
        sd.nxt = o; sd.crt = o; sd.dZ = 0.0; sd.sh_nxt = 1.0; sd.sh_crt = 1.0;
        sd.dsm = __ldg(&m.dshift_max[o]); sd.sqw = __ldg(&m.sqrtw_max[o]);
        double* y = ws.rec(lane) + 1;
        make_record(m, y, l0 + lane, sd, 1, 1);
        *ws.last(lane) = 1;
        const double dX = *dsm_last(ws, lane), dY = sd.dsm;
        const double dsh = fabs(MGB_REC(y, R_SHIFT) - MGB_REC(y - 1, R_SHIFT));
        const double iwn = MGB_REC(y, R_IW) * kINV_SQRT_PI;
        MGB_REC(y, R3_IWN) = (dsh > ((dY < dX) ? dY : dX)) ? -iwn : iwn;
        *dsm_last(ws, lane) = dY;
    }
}

// ---------------------------------------------------------------------------------------------------
// the line-integrated branch on the range-free erff
// ---------------------------------------------------------------------------------------------------
// erff(pY) - erff(pX), a float subtraction (solver.tpp:2450); both evaluations are straight-line and interleave.
__device__ __forceinline__ float erff_diff(float pY, float pX) {
    const int iy = __float_as_int(pY) & 0x7fffffff, ix = __float_as_int(pX) & 0x7fffffff;
    /* inf, NaN and |x| < 2^-28 keep the branchy routine (mgb_erff_u's rule) */
    const bool weird = ((unsigned)(iy - 0x31800000) >= (unsigned)(0x7f800000 - 0x31800000)) | ((unsigned)(ix - 0x31800000) >= (unsigned)(0x7f800000 - 0x31800000));
    float eY = mgb_erff_main(pY), eX = mgb_erff_main(pX);
    if (weird) { eY = erff_ool(pY); eX = erff_ool(pX); }
    return eY - eX;
}

// line_integrated_own of mgb_solve.cuh (compute_S_dtau_line_integrated for a window holding exactly the own line, solver.tpp:1878-1912,
// 2012-2111, 2374-2455) with erff_diff and the inline division; same operations in the same order.
template <int APPROX>
__device__ __noinline__ LineIntegral line_integrated_own3(double min_opacity, double linefreq, double freqX, double freqY, double dz, double iwX,
                                                          double iwY, double ioX, double ioY, double slX, double slY) {
    const double avg_iw = (iwX + iwY) / 2.0;
    const double avg_op = (ioX + ioY) / 2.0;
    double ld;
    if (freqX == freqY) {
        const double prof = gaussian(avg_iw, freqX - linefreq);
        ld = dz * (prof * avg_op + min_opacity);
    } else {
        const double posY = (linefreq - freqY) * avg_iw;
        const double posX = (linefreq - freqX) * avg_iw;
        const double diff_pos = posY - posX;
        const float erfdiff   = erff_diff((float)posY, (float)posX);
        const double erfterm  = ddiv_raw(avg_op, diff_pos) / 2.0 * (double)erfdiff;
        ld = dz * (avg_iw * erfterm + min_opacity);
    }
    if (APPROX == APPROX_ONELINE) return LineIntegral{ld, slX, slY};
    const double sum_dtau = 0.0 + ld, sum_x = 0.0 + ld * slX, sum_y = 0.0 + ld * slY;
    double d = sum_dtau;
    const double bound_min_dtau = min_opacity * dz;
    if (-bound_min_dtau < d) d = (bound_min_dtau < d) ? d : bound_min_dtau;
    const Div2 dv = ddiv2(sum_x, sum_y, d);
    return LineIntegral{d, dv.q1, dv.q2};
}

// ---------------------------------------------------------------------------------------------------
// phase B (lanes = frequency bins)
// ---------------------------------------------------------------------------------------------------
struct Lane3 {
    double chiX, SX;   // eta/chi evaluation at the far entry X of the open segment
    double dtau_far;   // optical depth of X's far segment
    double SatY_far;   // line-integrated source of that segment at its near end (= at X)
    double F, a;       // Rybicki-Hummer F (FF or GG), eliminated source carried as b = Su * (1 + F)
    bool haveX, X_is_end, large_far;
    unsigned nlarge;
};

// The general step (k_solve's sweep_step: boundary rows, windows holding other lines, ApproximationType None) with the segment's
// large-shift flag read from the record. Out of line, state by value: a small minority of the entries.
template <int APPROX>
__device__ __noinline__ Lane3 step3_general(const DevModel& m, LaneLine L, Lane3 st, double freq, int side, const double* q, uint32_t ll) {
    const double shY = MGB_REC(q, R_SHIFT);
    const double frY = freq * shY;
    const double iwns = MGB_REC(q, R3_IWN);
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
            const double eta  = 0.0 + prof * MGB_REC(q, R_IE);
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
        double dtau, S_atX = 0.0, S_atY = 0.0;
        const bool large = __double2hiint(iwns) < 0;
        if (large) {
            st.nlarge++;
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
                const LineIntegral li = line_integrated_own3<APPROX>(m.min_opacity, L.s_own, frX, frY, dzX, MGB_REC(x, R_IW), MGB_REC(q, R_IW),
                                                                     MGB_REC(x, R_IO), MGB_REC(q, R_IO), MGB_REC(x, R_SLINE), MGB_REC(q, R_SLINE));
                dtau = li.dtau; S_atX = li.S_atX; S_atY = li.S_atY;
            } else if (cls == WIN_EMPTY) {
                double d = 0.0;
                const double bound_min_dtau = m.min_opacity * dzX;
                if (-bound_min_dtau < d) d = (bound_min_dtau < d) ? d : bound_min_dtau;
                dtau = d; S_atX = zero_over(d); S_atY = S_atX;
            } else {
                const uint2 pX = *reinterpret_cast<const uint2*>(&MGB_REC(x, R_PTS)), pY = *reinterpret_cast<const uint2*>(&MGB_REC(q, R_PTS));
                const double fX = MGB_REC(x, R_FACTOR), fY = MGB_REC(q, R_FACTOR);
                if (side == 0) S_dtau_generic<APPROX>(m, pX.x, pX.y, fX, pY.x, pY.y, fY, ll, frX, frY, dzX, dtau, S_atX, S_atY);
                else S_dtau_generic<APPROX>(m, pY.x, pY.y, fY, pX.x, pX.y, fX, ll, frY, frX, dzX, dtau, S_atY, S_atX);
            }
        } else {
            dtau = 0.5 * (st.chiX + chiY) * dzX;
        }
        if (st.X_is_end) { // boundary row (solver.tpp:2240-2262 for `first`, 2283-2297 mirrored for `last`)
            const double S0  = large ? S_atX : st.SX;
            const double inv = drcp(dtau);
            const double Cb  = 2.0 * inv * inv;
            const double Bm  = 1.0 + 2.0 * inv;
            const double B   = Bm + Cb;
            const uint2 pX   = *reinterpret_cast<const uint2*>(&MGB_REC(x, R_PTS));
            const double Ib  = boundary_intensity(m, pX.x, freq * shX);
            const double Su0 = ddiv(S0 + 2.0 * Ib * inv, B);
            st.F = 0.5 * Bm * dtau * dtau;
            st.a = Su0 * (1.0 + st.F);
        } else { // interior row (solver.tpp:2264-2281; mirrored 2306-2311)
            const double Sx      = side ? (st.large_far ? st.SatY_far : st.SX) : (large ? S_atX : st.SX);
            const double avg     = 0.5 * (st.dtau_far + dtau);
            const double invFar  = avg * st.dtau_far;
            const double invNear = avg * dtau;
            const double P       = drcp(invFar * (1.0 + st.F));
            st.a = (st.a * P + Sx) * invNear;
            st.F = (st.F * P + 1.0) * invNear;
        }
        st.dtau_far  = dtau;
        st.large_far = large;
        st.SatY_far  = S_atY;
        st.X_is_end  = false;
    } else {
        st.haveX    = true;
        st.X_is_end = true;
    }
    st.chiX = chiY; st.SX = SY;
    return st;
}

// The common step: an interior row whose windows hold the own line or no line. Returns false (nothing done) when a participating
// lane of the warp needs the general step.
template <int APPROX>
__device__ __forceinline__ bool step3_common(const DevModel& m, const LaneLine& L, Lane3& st, double freq, int side, const double* q, bool part) {
    if (APPROX == APPROX_NONE) return false;
    const double* x  = q - 1;
    const double shY = MGB_REC(q, R_SHIFT), sqwY = MGB_REC(q, R_SQW);
    const double iwns = MGB_REC(q, R3_IWN);
    const bool large  = __double2hiint(iwns) < 0;
    const double frY  = freq * shY;
    int cls;
    if (APPROX == APPROX_CLOSELINES) {
        const double ub = m.max_dist * (frY * sqwY);
        cls = window_class(L, frY - ub, frY + ub);
    } else {
        cls = L.own ? WIN_OWN : WIN_OTHER;
    }
    bool ok = st.haveX & !st.X_is_end & (cls != WIN_OTHER);
    int clsS = cls;
    double frX = 0.0;
    if (__any_sync(0xffffffffu, large & part)) { /* the segment's own window matters to large shifts only */
        frX = freq * MGB_REC(x, R_SHIFT);
        if (APPROX == APPROX_CLOSELINES) {
            const double sqwX = MGB_REC(x, R_SQW);
            const double lf = (frX < frY) ? frX : frY, rf = (frX < frY) ? frY : frX;
            const double bX = m.max_dist * (rf * sqwX), bY = m.max_dist * (rf * sqwY);
            const double ub = (bX < bY) ? bY : bX;
            clsS = window_class(L, lf - ub, rf + ub);
        }
        ok &= !large | (clsS != WIN_OTHER);
    }
    const double iw = MGB_REC(q, R_IW), ie = MGB_REC(q, R_IE), io = MGB_REC(q, R_IO);
    double chiY, SY;
    bool weird;
    eta_chi_own<APPROX>(m, cls == WIN_OWN, iw, fabs(iwns), ie, io, frY - L.s_own, chiY, SY, weird);
    ok &= !weird;
    if (!__all_sync(0xffffffffu, ok | !part)) return false;
    if (!part) return true;
    const double dzX = MGB_REC(x, R_DZ);
    double dtau, Sx_large = 0.0, Sy = 0.0;
    if (large) {
        st.nlarge++;
        if (clsS == WIN_OWN) {
            const LineIntegral li = line_integrated_own3<APPROX>(m.min_opacity, L.s_own, frX, frY, dzX, MGB_REC(x, R_IW), iw, MGB_REC(x, R_IO), io,
                                                                 MGB_REC(x, R_SLINE), MGB_REC(q, R_SLINE));
            dtau = li.dtau; Sx_large = li.S_atX; Sy = li.S_atY;
        } else { /* no line in reach: sums stay 0, dtau bounded below by min_opacity * dz (solver.tpp:2095-2110) */
            double d = 0.0;
            const double bound_min_dtau = m.min_opacity * dzX;
            if (-bound_min_dtau < d) d = (bound_min_dtau < d) ? d : bound_min_dtau;
            dtau = d; Sx_large = zero_over(d); Sy = Sx_large;
        }
    } else {
        dtau = 0.5 * (st.chiX + chiY) * dzX;
    }
    /* interior row of X (solver.tpp:2264-2281; mirrored 2306-2311): its source is the line-integrated one of the segment in which X
       is `curr` (the near segment on side 0, the far one on side 1) */
    const double Sx      = side ? (st.large_far ? st.SatY_far : st.SX) : (large ? Sx_large : st.SX);
    const double avg     = 0.5 * (st.dtau_far + dtau);
    const double invFar  = avg * st.dtau_far;
    const double invNear = avg * dtau;
    const double P       = drcp_inl(invFar * (1.0 + st.F)); /* A * FI of the previous row */
    st.a = (st.a * P + Sx) * invNear;
    st.F = (st.F * P + 1.0) * invNear;
    st.dtau_far = dtau; st.SatY_far = Sy; st.chiX = chiY; st.SX = SY;
    st.large_far = large;
    return true;
}

#ifndef MGB_SOLVE3_WARPS
#define MGB_SOLVE3_WARPS 2
#endif
#ifndef MGB_SOLVE3_CTAS
#define MGB_SOLVE3_CTAS 8
#endif
constexpr int kSolve3Warps = MGB_SOLVE3_WARPS, kSolve3Ctas = MGB_SOLVE3_CTAS;

template <int APPROX>
__global__ void __launch_bounds__(32 * kSolve3Warps, kSolve3Ctas) k_solve3(const __grid_constant__ DevModel m, const __grid_constant__ PathView pv, ScratchView sc,
                                                                          const FreqChunk* __restrict__ chunks, uint32_t nchunks_f, uint32_t lpw) {
    const unsigned lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const size_t item   = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const size_t nitems = (size_t)pv.no * m.hnrays * nchunks_f;
    stage_exp_table();
    stage_erff_table();
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
    const uint32_t l   = f / m.nquads;       // line of this frequency bin (f = l*nquads + z, model.cpp:127-146)
    const uint32_t z   = f - l * m.nquads;
    const uint32_t ks  = l - ch.l0;          // record buffer of this lane's line
    const uint32_t ll  = (APPROX == APPROX_ONELINE) ? __ldg(&m.corr_line[f]) : l;
    const uint32_t ar  = __ldg(&m.antipod[rr]);
    const double freq  = __ldg(&m.nu[(size_t)o * m.nfreqs + f]);
    const size_t out   = ((size_t)rr * pv.no + ol) * m.nfreqs + f;

    const uint32_t len_b = __ldg(&pv.len[(size_t)o * m.nrays + rr]);
    const uint32_t len_f = __ldg(&pv.len[(size_t)o * m.nrays + ar]);
    if (len_b + len_f == 0) { // isolated origin (solver.tpp:510-518)
        if (active) { sc.u[out] = boundary_intensity(m, o, freq); sc.L[out] = 0.0; }
        const unsigned e1 = __popc(__ballot_sync(0xffffffffu, active && z == 0)); /* n_tot = 1 per line */
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

    unsigned nlarge = 0, nentries = 1; /* the origin's own entry */
    double SC = 0, F1 = 0, FI1 = 0, a1 = 0, dt1 = 0, Sc1 = 0;
    bool lg1 = false;
    {   /* side-0 results wait in shared memory while side 1 runs */
        double* s0 = ws.side0(lane);
        s0[0] = 0.0; s0[32] = 0.0; s0[64] = 0.0; s0[96] = 0.0; s0[128] = 0.0; s0[160] = 0.0;
    }
    const double* rb = ws.rec(ks) + 1; /* entry e of this lane's line */

    for (int side = 0; side < 2; side++) {
        const uint32_t r   = side ? ar : rr;
        const uint32_t len = side ? len_f : len_b;
        if (len == 0) continue;
        const size_t hr = (size_t)ol * m.nrays + r;
        const unsigned long long off = __ldg(&pv.poff[hr]);
        const double shift0 = __ldg(&pv.shift0[hr]);
        Lane3 st;
        st.haveX = false; st.X_is_end = false; st.large_far = false; st.nlarge = 0;
        st.chiX = st.SX = st.F = st.a = st.dtau_far = st.SatY_far = 0.0;
        __syncwarp();
        if (lane < ch.nl) { *ws.isub(lane) = 0; *ws.last(lane) = 0; *dsm_last(ws, lane) = 0.0; }
        __syncwarp();
        uint32_t j = len; /* geometric steps not yet expanded (warp-uniform) */
        for (;;) {
            const bool centre = (j == 0);
            carry_last_records(ws, ch.nl, lane);
            if (!centre) produce_window3(m, pv, ws, ch.l0, ch.nl, o, off, shift0, j, lane);
            else produce_centre3(m, ws, ch.l0, ch.nl, o, lane);
            __syncwarp();
            const uint32_t total = *ws.last(ks);
            if (!centre) nentries += total;
            const uint32_t tmax = __reduce_max_sync(0xffffffffu, total);
            for (uint32_t e = 0; e < tmax; e++) {
                const bool part = e < total;
                const double* q = rb + (part ? e : 0u);
                if (!step3_common<APPROX>(m, L, st, freq, side, q, part)) {
                    if (part) st = step3_general<APPROX>(m, L, st, freq, side, q, ll);
                }
            }
            __syncwarp();
            if (centre) break;
        }
        SC = st.SX; /* source function at the origin */
        const double FI = drcp(1.0 + st.F); /* FI and Su of the row next to the origin, from the carried b = Su * (1 + F) */
        const double aa = st.a * FI;
        nlarge += st.nlarge;
        if (side == 0) {
            double* s0 = ws.side0(lane);
            s0[0] = st.F; s0[32] = FI; s0[64] = aa; s0[96] = st.dtau_far; s0[128] = st.SatY_far; s0[160] = st.large_far ? 1.0 : 0.0;
        } else { F1 = st.F; FI1 = FI; a1 = aa; dt1 = st.dtau_far; lg1 = st.large_far; Sc1 = st.SatY_far; }
    }

    // ---- centre row; the Lambda increment's bin- and ray-independent factors are applied by k_reduce_lambda2 ----
    const double* s0 = ws.side0(lane);
    const CentreOut c = centre_row(m, o, freq, len_b != 0, len_f != 0, SC, s0[0], s0[32], s0[64], s0[96], s0[128], s0[160] != 0.0, F1, FI1, a1, dt1, Sc1, lg1);
    if (active) { sc.u[out] = c.u; sc.L[out] = c.Ldiag; }

    // work counters: Sigma N counts one ray per (o, rr, line); large-shift evaluations per bin
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

} // namespace mgb
