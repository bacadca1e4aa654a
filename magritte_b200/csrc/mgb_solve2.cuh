// mgb_solve2.cuh -- k_solve2: the hot kernel of the path for the diagonal Lambda operator (n_off_diag = 0), second generation.
//
// Same algorithm and the SAME ARITHMETIC as k_solve (mgb_solve.cuh: phase A records per (entry, line), phase B per frequency bin,
// two-sided Rybicki-Hummer elimination meeting at the origin; reference solver.tpp:2139-2186, 2193-2356) -- every result bit is
// k_solve's, which tests/test_gpu_parity.py pins to the oracle's two-sided restatement. What changed is the shape of the work,
// each item from an ncu finding on the round-1 kernel (profiles/ncu_k_solve_r1g_c3.txt: 414 thread-instructions per entry x bin
// of which 83 FP64, issue-active 45 %, 24.2 of 32 lanes, 212/316 B of spills at the 128-register cap):
//
//   * TWO frequency bins per lane, both of the SAME line (bins z and z + ceil(nquads/2)): the record loads, the loop control and
//     the bin-independent tests are shared, and the two dependent chains (expf -> division -> reciprocal -> recurrence) interleave
//     in one basic block. A warp owns whole lines (5 CO lines = 55 of 64 bin slots): no line is expanded by two warps any more
//     (phase A ran 54 line expansions for 40 lines).
//   * The common step is straight-line code: Gaussian profile (glibc expf inlined, branch-free over its under-flow cases), IEEE
//     division and reciprocal on the raw operands (no exponent re-scaling: one range test), no calls. Round 1 had 137 CALLs and
//     ~1000 IMAD.MOV register shuffles around them in 5800 SASS instructions.
//   * Everything that does not depend on the bin moved to phase A: the large-shift test of a segment (|dshift| > dshift_max,
//     compute_source_dtau solver.tpp:2150-2153) is a flag in the record (sign of the pre-scaled inverse width), inverse_width/sqrt(pi)
//     is stored next to inverse_width.
//   * The Lambda increment (update_Lambda, solver.tpp:1760-1806) needs eta/chi and the line profile AT THE ORIGIN, which depend on
//     (o, f) only: k_solve evaluated them per (o, rr, f), 96 times too often. k_solve2 writes L_diag; k_reduce_lambda2 forms
//     constante * nu * phi * L_diag / chi in the same operation order once per (o, f) and ray pair.
//   * Lanes walk their own line's entries linearly (no per-step max-reduction and offset loads); the centre-row divisions use the
//     Newton reciprocal instead of the compiler's slow path.
// The banded operator (n_off_diag > 0) stays with k_solve<APPROX, true>.
#pragma once
#include "mgb_solve.cuh"

namespace mgb {

// work item of one warp: lines [l0, l0+nl), of each the bins z in [z0, z0+nz); lane = k * lpl + j owns bins z0+j and z0+j+lpl of line l0+k
struct FreqChunk2 {
    uint32_t l0, nl, z0, nz, lpl;
};

constexpr int kRec2Fields = 11;
enum { R_IWN = 10 }; // inverse_width * (1/sqrt(pi)) with the large-shift flag of the segment (previous entry -> this entry) in its sign
constexpr int kLine2Doubles = kRec2Fields * kRecStride + kWordsPerLine / 2;
__host__ __device__ inline size_t solve2_smem_per_warp(uint32_t lpw) { return 8 * ((size_t)lpw * kLine2Doubles + 12 * 32 /* side-0 results of both bins */); }

struct WarpSmem2 {
    uint32_t base, lpw;
    __device__ __forceinline__ double* rec(uint32_t k) const { return mgb_smem + base + k * (kRec2Fields * kRecStride); }
    __device__ __forceinline__ uint32_t* words(uint32_t k) const {
        return reinterpret_cast<uint32_t*>(mgb_smem + base + lpw * (kRec2Fields * kRecStride)) + k * kWordsPerLine;
    }
    __device__ __forceinline__ uint32_t* ni(uint32_t k) const { return words(k); }
    __device__ __forceinline__ uint32_t* excl(uint32_t k) const { return words(k) + 32; }
    __device__ __forceinline__ uint32_t* mark(uint32_t k) const { return words(k) + 64; }
    __device__ __forceinline__ uint32_t* isub(uint32_t k) const { return words(k) + 100; }
    __device__ __forceinline__ uint32_t* last(uint32_t k) const { return words(k) + 101; }
    __device__ __forceinline__ double* side0(unsigned lane) const { return mgb_smem + base + lpw * kLine2Doubles + lane; } // [12][32]
};

// ---------------------------------------------------------------------------------------------------
// phase A (lanes = ray entries): the records of mgb_solve.cuh plus R_IWN and the large-shift flag
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void make_record2(const DevModel& m, double* r, uint32_t l, const StepData& sd, uint32_t i, uint32_t ni) {
    make_record(m, r, l, sd, i, ni);
    MGB_REC(r, R_IWN) = MGB_REC(r, R_IW) * kINV_SQRT_PI; /* gaussian(): inverse_width * INVERSE_SQRT_PI * expf(..), left to right */
}

// large-shift flag of the segments ending at this round's entries 0 .. total-1 of line k (slots 1 .. total; slot 0 = carried entry):
// compute_source_dtau's test |shift_next - shift_curr| > min(dshift_max[curr], dshift_max[next]) (solver.tpp:2150-2153)
__device__ __forceinline__ void flag_large(const WarpSmem2& ws, uint32_t k, uint32_t total, unsigned lane) {
    if (lane < total) {
        double* y = ws.rec(k) + 1 + lane;
        const double* x = y - 1;
        const double dshift_abs = fabs(MGB_REC(y, R_SHIFT) - MGB_REC(x, R_SHIFT));
        const double dX = MGB_REC(x, R_DSM), dY = MGB_REC(y, R_DSM);
        const double dshift_max = (dY < dX) ? dY : dX;
        if (dshift_abs > dshift_max) MGB_REC(y, R_IWN) = -MGB_REC(y, R_IWN);
    }
}

__device__ __forceinline__ void carry_last_records2(const WarpSmem2& ws, uint32_t nl, unsigned lane) {
    for (uint32_t k = 0; k < nl; k++) {
        const uint32_t last = *ws.last(k);
        if (last && lane < kRec2Fields) {
            double* row = ws.rec(k) + lane * kRecStride;
            row[0] = row[last];
        }
    }
    __syncwarp();
}

// produce_window of mgb_solve.cuh on the 11-field records (same expansion; all lines advance by the same geometric steps).
// Returns nothing: *ws.last(k) holds the number of entries of line k now in slots 1 .. last.
__device__ __forceinline__ void produce_window2(const DevModel& m, const PathView& pv, const WarpSmem2& ws, uint32_t l0, uint32_t nl,
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
            uint32_t* mark = ws.mark(k);
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
            if (lane < total) make_record2(m, ws.rec(k) + 1 + lane, l0 + k, se, ni_e - (lane - excl_e), ni_e);
            if (lane == 0) *ws.last(k) = total;
            __syncwarp();
            flag_large(ws, k, total, lane);
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
        if (lane < nemit) make_record2(m, ws.rec(k) + 1 + lane, l0 + k, s0, ni - (is + lane), ni);
        __syncwarp();
        if (lane == 0) { *ws.isub(k) = is + nemit; *ws.last(k) = nemit; }
        flag_large(ws, k, nemit, lane);
        all_done = all_done && (is + nemit == ni);
    }
    if (all_done) {
        __syncwarp();
        if (lane < nl) *ws.isub(lane) = 0;
        j -= 1;
    }
}

__device__ __forceinline__ void produce_centre2(const DevModel& m, const WarpSmem2& ws, uint32_t l0, uint32_t nl, uint32_t o, unsigned lane) {
    if (lane < nl) {
        StepData sd;
        sd.nxt = o; sd.crt = o; sd.dZ = 0.0; sd.sh_nxt = 1.0; sd.sh_crt = 1.0;
        sd.dsm = __ldg(&m.dshift_max[o]); sd.sqw = __ldg(&m.sqrtw_max[o]);
        double* y = ws.rec(lane) + 1;
        make_record2(m, y, l0 + lane, sd, 1, 1);
        *ws.last(lane) = 1;
        const double* x = y - 1; /* flag of the segment (last entry of the half-ray -> origin) */
        const double dshift_abs = fabs(MGB_REC(y, R_SHIFT) - MGB_REC(x, R_SHIFT));
        const double dX = MGB_REC(x, R_DSM), dY = MGB_REC(y, R_DSM);
        if (dshift_abs > ((dY < dX) ? dY : dX)) MGB_REC(y, R_IWN) = -MGB_REC(y, R_IWN);
    }
}

// ---------------------------------------------------------------------------------------------------
// phase B (lanes = pairs of frequency bins of one line)
// ---------------------------------------------------------------------------------------------------
struct BinState {
    double freq;
    double chiX, SX;   // eta/chi evaluation at the far entry X of the open segment
    double dtau_far;   // optical depth of X's far segment
    double SatY_far;   // line-integrated source of that segment at its near end (= at X)
    double F, a;       // Rybicki-Hummer F (FF or GG), eliminated source carried as b = Su * (1 + F)
};
struct LaneState {
    bool haveX, X_is_end, large_far;
    unsigned nlarge;
};

// k_solve's sweep_step_general on one bin of the pair (boundary rows, windows holding other lines, ApproximationType None)
template <int APPROX>
__device__ __noinline__ BinState step2_general(const DevModel& m, LaneLine L, BinState b, LaneState ls, int side, const double* q, uint32_t ll,
                                               bool own) { /* by value in and out: the pair's state must stay in registers in the hot loop */
    SweepState st;
    st.chiX = b.chiX; st.SX = b.SX; st.dtau_far = b.dtau_far; st.SatY_far = b.SatY_far; st.F = b.F; st.FI = 0.0; st.a = b.a;
    st.haveX = ls.haveX; st.X_is_end = ls.X_is_end; st.large_far = ls.large_far; st.nlarge = 0;
    L.own = own;
    unsigned nl = 0;
    sweep_step<APPROX>(m, L, st, side, q, ll, b.freq, nl);
    b.chiX = st.chiX; b.SX = st.SX; b.dtau_far = st.dtau_far; b.SatY_far = st.SatY_far; b.F = st.F; b.a = st.a;
    return b;
}

// chi and S = eta/chi at the entry Y for one bin whose window holds the own line (cls == WIN_OWN) or nothing (WIN_EMPTY)
template <int APPROX>
__device__ __forceinline__ void eta_chi_own(const DevModel& m, bool own_line, double iw, double iwn, double ie, double io, double diff,
                                            double& chi, double& S, bool& weird) {
    const double sqrt_exp = iw * diff;
    const double prof = iwn * gauss_expf(-sqrt_exp * sqrt_exp, weird);
    /* the reference accumulates into eta = 0, chi = min_opacity (CloseLines) or adds min_opacity last (OneLine); an empty window
       leaves eta = 0, chi = min_opacity, and S = 0 / chi is what the division returns for a zero numerator */
    const double eta = own_line ? 0.0 + prof * ie : 0.0;
    if (APPROX == APPROX_ONELINE) chi = own_line ? (0.0 + prof * io) + m.min_opacity : m.min_opacity;
    else chi = own_line ? m.min_opacity + prof * io : m.min_opacity;
    S = ddiv_raw(eta, chi);
}

// The common step for both bins of the lane: an interior row whose windows hold the own line or no line. Returns false (nothing
// done) when a participating lane of the warp needs the general step.
template <int APPROX>
__device__ __forceinline__ bool step2_common(const DevModel& m, const LaneLine& L, BinState& A, BinState& B, LaneState& ls, bool ownA, bool ownB,
                                             int side, const double* q, bool part) {
    if (APPROX == APPROX_NONE) return false;
    const double* x  = q - 1;
    const double shY = MGB_REC(q, R_SHIFT), sqwY = MGB_REC(q, R_SQW);
    const double iwns = MGB_REC(q, R_IWN);
    const bool large  = __double2hiint(iwns) < 0;
    const double frYA = A.freq * shY, frYB = B.freq * shY;
    int clsA, clsB;
    if (APPROX == APPROX_CLOSELINES) {
        const double ubA = m.max_dist * (frYA * sqwY), ubB = m.max_dist * (frYB * sqwY);
        clsA = window_class(L, frYA - ubA, frYA + ubA);
        clsB = window_class(L, frYB - ubB, frYB + ubB);
    } else {
        clsA = ownA ? WIN_OWN : WIN_OTHER;
        clsB = ownB ? WIN_OWN : WIN_OTHER;
    }
    bool ok = ls.haveX & !ls.X_is_end & (clsA != WIN_OTHER) & (clsB != WIN_OTHER);
    int clsSA = clsA, clsSB = clsB;
    double frXA = 0.0, frXB = 0.0;
    if (__any_sync(0xffffffffu, large & part)) { /* the segment's own window matters to large shifts only */
        const double shX = MGB_REC(x, R_SHIFT);
        frXA = A.freq * shX; frXB = B.freq * shX;
        if (APPROX == APPROX_CLOSELINES) {
            const double sqwX = MGB_REC(x, R_SQW);
            {
                const double lf = (frXA < frYA) ? frXA : frYA, rf = (frXA < frYA) ? frYA : frXA;
                const double bX = m.max_dist * (rf * sqwX), bY = m.max_dist * (rf * sqwY);
                const double ub = (bX < bY) ? bY : bX;
                clsSA = window_class(L, lf - ub, rf + ub);
            }
            {
                const double lf = (frXB < frYB) ? frXB : frYB, rf = (frXB < frYB) ? frYB : frXB;
                const double bX = m.max_dist * (rf * sqwX), bY = m.max_dist * (rf * sqwY);
                const double ub = (bX < bY) ? bY : bX;
                clsSB = window_class(L, lf - ub, rf + ub);
            }
        }
        ok &= !large | ((clsSA != WIN_OTHER) & (clsSB != WIN_OTHER));
    }
    const double iw = MGB_REC(q, R_IW), ie = MGB_REC(q, R_IE), io = MGB_REC(q, R_IO);
    const double iwn = fabs(iwns);
    double chiA, SA, chiB, SB;
    bool wA, wB;
    eta_chi_own<APPROX>(m, clsA == WIN_OWN, iw, iwn, ie, io, frYA - L.s_own, chiA, SA, wA);
    eta_chi_own<APPROX>(m, clsB == WIN_OWN, iw, iwn, ie, io, frYB - L.s_own, chiB, SB, wB);
    ok &= !(wA | wB);
    if (!__all_sync(0xffffffffu, ok | !part)) return false;
    if (!part) return true;
    const double dzX = MGB_REC(x, R_DZ);
    double dtauA, dtauB, SxA_large = 0.0, SxB_large = 0.0, SyA = 0.0, SyB = 0.0;
    if (large) {
        ls.nlarge++;
        if (clsSA == WIN_OWN) {
            const LineIntegral li = line_integrated_own<APPROX>(m.min_opacity, L.s_own, frXA, frYA, dzX, MGB_REC(x, R_IW), iw, MGB_REC(x, R_IO), io,
                                                                MGB_REC(x, R_SLINE), MGB_REC(q, R_SLINE));
            dtauA = li.dtau; SxA_large = li.S_atX; SyA = li.S_atY;
        } else { /* no line in reach: sums stay 0, dtau bounded below by min_opacity * dz (solver.tpp:2095-2110) */
            double d = 0.0;
            const double bound_min_dtau = m.min_opacity * dzX;
            if (-bound_min_dtau < d) d = (bound_min_dtau < d) ? d : bound_min_dtau;
            dtauA = d; SxA_large = zero_over(d); SyA = SxA_large;
        }
        if (clsSB == WIN_OWN) {
            const LineIntegral li = line_integrated_own<APPROX>(m.min_opacity, L.s_own, frXB, frYB, dzX, MGB_REC(x, R_IW), iw, MGB_REC(x, R_IO), io,
                                                                MGB_REC(x, R_SLINE), MGB_REC(q, R_SLINE));
            dtauB = li.dtau; SxB_large = li.S_atX; SyB = li.S_atY;
        } else {
            double d = 0.0;
            const double bound_min_dtau = m.min_opacity * dzX;
            if (-bound_min_dtau < d) d = (bound_min_dtau < d) ? d : bound_min_dtau;
            dtauB = d; SxB_large = zero_over(d); SyB = SxB_large;
        }
    } else {
        dtauA = 0.5 * (A.chiX + chiA) * dzX;
        dtauB = 0.5 * (B.chiX + chiB) * dzX;
    }
    /* interior row of X (solver.tpp:2264-2281; mirrored 2306-2311): its source is the line-integrated one of the segment in which X
       is `curr` (the near segment on side 0, the far one on side 1) */
    {
        const double Sx      = side ? (ls.large_far ? A.SatY_far : A.SX) : (large ? SxA_large : A.SX);
        const double avg     = 0.5 * (A.dtau_far + dtauA);
        const double invFar  = avg * A.dtau_far;
        const double invNear = avg * dtauA;
        const double P       = drcp_inl(invFar * (1.0 + A.F)); /* A * FI of the previous row */
        A.a = (A.a * P + Sx) * invNear;
        A.F = (A.F * P + 1.0) * invNear;
        A.dtau_far = dtauA; A.SatY_far = SyA; A.chiX = chiA; A.SX = SA;
    }
    {
        const double Sx      = side ? (ls.large_far ? B.SatY_far : B.SX) : (large ? SxB_large : B.SX);
        const double avg     = 0.5 * (B.dtau_far + dtauB);
        const double invFar  = avg * B.dtau_far;
        const double invNear = avg * dtauB;
        const double P       = drcp_inl(invFar * (1.0 + B.F));
        B.a = (B.a * P + Sx) * invNear;
        B.F = (B.F * P + 1.0) * invNear;
        B.dtau_far = dtauB; B.SatY_far = SyB; B.chiX = chiB; B.SX = SB;
    }
    ls.large_far = large;
    return true;
}

// The centre row and L_diag of one bin (k_solve's tail; reference solver.tpp:2240-2262, 2283-2317) with the Newton reciprocal /
// division in place of the compiler's (same bits).
struct CentreOut { double u, Ldiag; };
__device__ __forceinline__ CentreOut centre_row(const DevModel& m, uint32_t o, double freq, bool has_b, bool has_f, double SC, double F0,
                                                double FI0, double a0, double dt0, double Sn0, bool lg0, double F1, double FI1, double a1,
                                                double dt1, double Sc1, bool lg1) {
    CentreOut r;
    if (has_b && has_f) {
        const double S_c  = lg1 ? Sc1 : SC;
        const double avg  = 0.5 * (dt0 + dt1);
        const double invA = avg * dt0, invC = avg * dt1;
        const double A = drcp(invA), Cc = drcp(invC);
        const double FFc = (A * F0 * FI0 + 1.0) * invC;
        const double GP  = F1 * FI1;
        r.Ldiag = ddiv(invC, FFc + GP);
        r.u     = (S_c + A * a0 + Cc * a1) * r.Ldiag;
    } else if (has_f) { /* centre == first */
        const double S_c = lg1 ? Sc1 : SC;
        const double inv = drcp(dt1);
        const double Cf  = 2.0 * inv * inv;
        const double Bm  = 1.0 + 2.0 * inv;
        const double Ib  = boundary_intensity(m, o, freq);
        const double FFf = 0.5 * Bm * dt1 * dt1;
        const double GP  = F1 * FI1;
        r.Ldiag = ddiv(drcp(Cf), FFf + GP);
        r.u     = (S_c + 2.0 * Ib * inv + Cf * a1) * r.Ldiag;
    } else { /* centre == last */
        const double S_c = lg0 ? Sn0 : SC;
        const double inv = drcp(dt0);
        const double Al  = 2.0 * inv * inv;
        const double Bm  = 1.0 + 2.0 * inv;
        const double Bl  = Bm + Al;
        const double Ib  = boundary_intensity(m, o, freq);
        r.Ldiag = ddiv(1.0 + F0, Bm + Bl * F0);
        r.u     = (Al * a0 + (S_c + 2.0 * Ib * inv)) * r.Ldiag;
    }
    return r;
}

#ifndef MGB_SOLVE2_WARPS
#define MGB_SOLVE2_WARPS 2
#endif
#ifndef MGB_SOLVE2_CTAS
#define MGB_SOLVE2_CTAS 5
#endif
constexpr int kSolve2Warps = MGB_SOLVE2_WARPS, kSolve2Ctas = MGB_SOLVE2_CTAS;

template <int APPROX>
__global__ void __launch_bounds__(32 * kSolve2Warps, kSolve2Ctas) k_solve2(const __grid_constant__ DevModel m, const __grid_constant__ PathView pv, ScratchView sc,
                                                                          const FreqChunk2* __restrict__ chunks, uint32_t nchunks_f, uint32_t lpw) {
    const unsigned lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const size_t item   = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const size_t nitems = (size_t)pv.no * m.hnrays * nchunks_f;
    stage_exp_table();
    if (item >= nitems) return; /* whole warps only */
    WarpSmem2 ws;
    ws.lpw  = lpw;
    ws.base = wib * (uint32_t)(solve2_smem_per_warp(lpw) / 8);
    const uint32_t cf = (uint32_t)(item % nchunks_f);
    const size_t t    = item / nchunks_f;
    const uint32_t rr = (uint32_t)(t % m.hnrays);
    const uint32_t ol = (uint32_t)(t / m.hnrays);
    const uint32_t o  = pv.o0 + ol;
    const FreqChunk2 ch = chunks[cf];
    // lane -> (line slot ks, bin pair): idle lanes shadow the last line's first pair (computed, never written)
    uint32_t ks = lane / ch.lpl, jp = lane - ks * ch.lpl;
    const bool active = ks < ch.nl;
    if (!active) { ks = ch.nl - 1; jp = 0; }
    const uint32_t zA = ch.z0 + jp;
    const bool hasB   = active && (jp + ch.lpl < ch.nz);
    const uint32_t zB = (jp + ch.lpl < ch.nz) ? zA + ch.lpl : zA;
    const uint32_t l  = ch.l0 + ks;
    const uint32_t fA = l * m.nquads + zA, fB = l * m.nquads + zB; // f = l*nquads + z (model.cpp:127-146)
    const uint32_t llA = (APPROX == APPROX_ONELINE) ? __ldg(&m.corr_line[fA]) : l;
    const uint32_t llB = (APPROX == APPROX_ONELINE) ? __ldg(&m.corr_line[fB]) : l;
    const uint32_t ar  = __ldg(&m.antipod[rr]);
    BinState A, B;
    A.freq = __ldg(&m.nu[(size_t)o * m.nfreqs + fA]);
    B.freq = __ldg(&m.nu[(size_t)o * m.nfreqs + fB]);
    const size_t outA = ((size_t)rr * pv.no + ol) * m.nfreqs + fA, outB = ((size_t)rr * pv.no + ol) * m.nfreqs + fB;

    const uint32_t len_b = __ldg(&pv.len[(size_t)o * m.nrays + rr]);
    const uint32_t len_f = __ldg(&pv.len[(size_t)o * m.nrays + ar]);
    if (len_b + len_f == 0) { // isolated origin (solver.tpp:510-518)
        if (active) { sc.u[outA] = boundary_intensity(m, o, A.freq); sc.L[outA] = 0.0; }
        if (hasB) { sc.u[outB] = boundary_intensity(m, o, B.freq); sc.L[outB] = 0.0; }
        const unsigned e1 = __popc(__ballot_sync(0xffffffffu, active && jp == 0 && ch.z0 == 0)); /* n_tot = 1 per line */
        if (lane == 0 && e1) atomicAdd(&sc.counters[0], (unsigned long long)e1);
        return;
    }

    LaneLine L;
    {
        const uint32_t pos = __ldg(&m.sorted_pos[l]);
        L.s_own  = __ldg(&m.line[l]);
        L.s_prev = pos ? __ldg(&m.sorted_line[pos - 1]) : -CUDART_INF;
        L.s_next = (pos + 1 < m.nlines) ? __ldg(&m.sorted_line[pos + 1]) : CUDART_INF;
        L.own    = (APPROX == APPROX_CLOSELINES);
    }
    const bool ownA = (APPROX == APPROX_CLOSELINES) || (APPROX == APPROX_ONELINE && llA == l);
    const bool ownB = (APPROX == APPROX_CLOSELINES) || (APPROX == APPROX_ONELINE && llB == l);

    unsigned nlarge = 0, nentries = 1; /* the origin's own entry */
    double SCA = 0, SCB = 0;
    double F1A = 0, FI1A = 0, a1A = 0, dt1A = 0, Sc1A = 0, F1B = 0, FI1B = 0, a1B = 0, dt1B = 0, Sc1B = 0;
    bool lg1 = false;
    {
        double* s0 = ws.side0(lane);
#pragma unroll
        for (int i = 0; i < 12; i++) s0[32 * i] = 0.0;
    }
    const double* rb = ws.rec(ks) + 1; /* entry e of this lane's line */

    for (int side = 0; side < 2; side++) {
        const uint32_t r   = side ? ar : rr;
        const uint32_t len = side ? len_f : len_b;
        if (len == 0) continue;
        const size_t hr = (size_t)ol * m.nrays + r;
        const unsigned long long off = __ldg(&pv.poff[hr]);
        const double shift0 = __ldg(&pv.shift0[hr]);
        LaneState ls;
        ls.haveX = false; ls.X_is_end = false; ls.large_far = false; ls.nlarge = 0;
        A.chiX = A.SX = A.F = A.a = A.dtau_far = A.SatY_far = 0.0;
        B.chiX = B.SX = B.F = B.a = B.dtau_far = B.SatY_far = 0.0;
        __syncwarp();
        if (lane < ch.nl) { *ws.isub(lane) = 0; *ws.last(lane) = 0; }
        __syncwarp();
        uint32_t j = len; /* geometric steps not yet expanded (warp-uniform) */
        for (;;) {
            const bool centre = (j == 0);
            carry_last_records2(ws, ch.nl, lane);
            if (!centre) produce_window2(m, pv, ws, ch.l0, ch.nl, o, off, shift0, j, lane);
            else produce_centre2(m, ws, ch.l0, ch.nl, o, lane);
            __syncwarp();
            const uint32_t total = *ws.last(ks);
            if (!centre) nentries += total;
            const uint32_t tmax = __reduce_max_sync(0xffffffffu, total);
            for (uint32_t e = 0; e < tmax; e++) {
                const bool part = e < total;
                const double* q = rb + (part ? e : 0u);
                if (!step2_common<APPROX>(m, L, A, B, ls, ownA, ownB, side, q, part)) {
                    if (part) {
                        /* the general step evaluates the segment's shift test itself from dshift_max (the flag is not read there) */
                        A = step2_general<APPROX>(m, L, A, ls, side, q, llA, ownA);
                        B = step2_general<APPROX>(m, L, B, ls, side, q, llB, ownB);
                        // lane flags after the general step (k_solve's sweep_step): first entry opens the segment, the next one closes
                        // the boundary row; `large_far` is the step's own shift test
                        if (ls.haveX) {
                            const double* x = q - 1;
                            const double dshift_abs = fabs(MGB_REC(q, R_SHIFT) - MGB_REC(x, R_SHIFT));
                            const double dX = MGB_REC(x, R_DSM), dY = MGB_REC(q, R_DSM);
                            const bool large = dshift_abs > ((dY < dX) ? dY : dX);
                            if (large) ls.nlarge++;
                            ls.large_far = large;
                            ls.X_is_end  = false;
                        } else {
                            ls.haveX = true; ls.X_is_end = true;
                        }
                    }
                }
            }
            __syncwarp();
            if (centre) break;
        }
        SCA = A.SX; SCB = B.SX; /* source function at the origin */
        const double FIA = drcp(1.0 + A.F), FIB = drcp(1.0 + B.F); /* FI and Su of the row next to the origin, from b = Su * (1 + F) */
        const double aA = A.a * FIA, aB = B.a * FIB;
        nlarge += ls.nlarge;
        if (side == 0) {
            double* s0 = ws.side0(lane);
            s0[0] = A.F; s0[32] = FIA; s0[64] = aA; s0[96] = A.dtau_far; s0[128] = A.SatY_far; s0[160] = ls.large_far ? 1.0 : 0.0;
            s0[192] = B.F; s0[224] = FIB; s0[256] = aB; s0[288] = B.dtau_far; s0[320] = B.SatY_far;
        } else {
            F1A = A.F; FI1A = FIA; a1A = aA; dt1A = A.dtau_far; Sc1A = A.SatY_far;
            F1B = B.F; FI1B = FIB; a1B = aB; dt1B = B.dtau_far; Sc1B = B.SatY_far;
            lg1 = ls.large_far;
        }
    }

    // ---- centre row; the Lambda increment's bin- and ray-independent factors are applied by k_reduce_lambda2 ----
    const double* s0 = ws.side0(lane);
    const bool lg0 = s0[160] != 0.0;
    const CentreOut cA = centre_row(m, o, A.freq, len_b != 0, len_f != 0, SCA, s0[0], s0[32], s0[64], s0[96], s0[128], lg0, F1A, FI1A, a1A, dt1A, Sc1A, lg1);
    const CentreOut cB = centre_row(m, o, B.freq, len_b != 0, len_f != 0, SCB, s0[192], s0[224], s0[256], s0[288], s0[320], lg0, F1B, FI1B, a1B, dt1B, Sc1B, lg1);
    if (active) { sc.u[outA] = cA.u; sc.L[outA] = cA.Ldiag; }
    if (hasB) { sc.u[outB] = cB.u; sc.L[outB] = cB.Ldiag; }

    // work counters: Sigma N counts one ray per (o, rr, line); large-shift evaluations per bin
    unsigned e = (active && jp == 0 && ch.z0 == 0) ? nentries : 0u;
    unsigned g = active ? nlarge * (hasB ? 2u : 1u) : 0u;
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

// Lambda(o, lid) from the L_diag values k_solve2 left in the scratch: update_Lambda's (solver.tpp:1760-1806)
//   L = constante * frq * phi * L_diag * inverse_chi,  constante = A_k * w_z * 2 w(rr),  frq = nu(o,f) * shift[centre] (= 1),
//   phi = thermodynamics.profile(inverse_mass, o, freq_line, frq),  inverse_chi = 1 / chi(o, freq_line)
// multiplied in that order, accumulated over rr then z as Lambda::add_element saw them (lambda.tpp:79-89). chi and phi depend on (o, f)
// only and are evaluated once here instead of once per ray pair inside the solve kernel. One thread per (o, line).
template <int APPROX>
__global__ void __launch_bounds__(128) k_reduce_lambda2(const __grid_constant__ DevModel m, uint32_t o0, uint32_t no, const uint32_t* __restrict__ lens,
                                                        const double* __restrict__ Ld, double* __restrict__ lam_val, uint32_t* __restrict__ lam_cnt,
                                                        double* __restrict__ fac /* [no][nfreqs] scratch: frq, then phi in fac2 */, double* __restrict__ fac2) {
    stage_exp_table();
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)no * m.nlines) return;
    const uint32_t ol = (uint32_t)(i / m.nlines), l = (uint32_t)(i % m.nlines);
    const uint32_t o  = o0 + ol;
    const double freq_line = __ldg(&m.line[l]);
    double eta_l, chi_l;
    // NOTE: the reference passes the within-species transition index k as OneLine's line index here (solver.tpp:1798)
    eta_chi_generic<APPROX>(m, o, o, 0.0, __ldg(&m.line_k[l]), freq_line, eta_l, chi_l);
    const double inverse_chi = 1.0 / chi_l;
    const double invmass = __ldg(&m.line_invmass[l]);
    for (uint32_t z = 0; z < m.nquads; z++) {
        const uint32_t f = l * m.nquads + z;
        const double frq = __ldg(&m.nu[(size_t)o * m.nfreqs + f]) * 1.0; /* nu(o,f) * shift[centre] */
        fac[(size_t)ol * m.nfreqs + f]  = frq;
        fac2[(size_t)ol * m.nfreqs + f] = thermo_profile(m, invmass, o, freq_line, frq);
    }
    const double A_l = __ldg(&m.line_A[l]);
    double acc = 0.0;
    uint32_t cnt = 0;
    for (uint32_t rr = 0; rr < m.hnrays; rr++) {
        const uint32_t ar = __ldg(&m.antipod[rr]);
        if (lens[(size_t)o * m.nrays + rr] + lens[(size_t)o * m.nrays + ar] == 0) continue; /* n_tot <= 1: no add_element */
        const double w_ang = 2.0 * ray_weight(m, o, rr);
        for (uint32_t z = 0; z < m.nquads; z++) {
            const uint32_t f = l * m.nquads + z;
            const double constante = A_l * __ldg(&m.line_quadw[(size_t)l * m.nquads + z]) * w_ang;
            const double v = constante * fac[(size_t)ol * m.nfreqs + f] * fac2[(size_t)ol * m.nfreqs + f] * Ld[((size_t)rr * no + ol) * m.nfreqs + f] * inverse_chi;
            if (cnt == 0) acc = v; else acc += v;
            cnt = 1;
        }
    }
    lam_val[(size_t)o * m.nlines + l] = acc;
    lam_cnt[(size_t)o * m.nlines + l] = cnt;
}

} // namespace mgb
