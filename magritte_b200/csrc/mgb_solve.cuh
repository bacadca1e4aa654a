// mgb_solve.cuh -- k_solve: eta/chi assembly + second-order Feautrier solve.
//
// Work item = one warp = (origin o, ray pair rr, one chunk of <= 32 consecutive frequency bins, which belong
// to <= LPW lines). The warp alternates between two phases over the cached geometric path of each half-ray,
// walking from the far (boundary) end towards the origin:
//
//   phase A  (lanes = ray ENTRIES of one line): expands the next 32 entries of that line's ray -- the reference's
//            sub-stepping Solver::set_data_for_line (solver.tpp:1057-1097) with n_interp from
//            InterpHelper::get_n_interp_for_line (solver_interp_helper.tpp:55-75) -- and evaluates everything that
//            does not depend on the frequency bin ONCE per (entry, line): Doppler shift, segment length, the
//            interpolated inverse width / emissivity / opacity of the line (interp_linear / interp_log with powf,
//            solver.tpp:1259-1302), the line source function of the line-integrated branch (:1888-1911) and the
//            per-point dshift_max / CloseLines window width. Records go to shared memory (SoA, 32 per line).
//   phase B  (lanes = FREQUENCY bins): every lane streams over its line's 32 records and does the per-frequency
//            work: CloseLines window test, Gaussian profile (bit-exact expf), eta, chi, S, dtau
//            (compute_source_dtau, solver.tpp:2139-2186, incl. the erff line-integrated branch) and one step of
//            the elimination, all in registers.
//
// The tridiagonal system is eliminated from both ends simultaneously:
//   first side  (entries first..centre-1):  the reference's own forward elimination (FF, Su; :2240-2297)
//   last side   (entries last..centre+1):   its mirror image (GG and the mirrored Su), which the reference
//                                           already uses for L_diag (:2302-2315)
//   centre row:  u_c = (S_c + A_c a_{c-1} + C_c a'_{c+1}) * L_diag,  L_diag = inverse_C/(FF_c + GP_{c+1})
// so no per-entry coefficient array is ever stored. This is algebraically the reference's forward-elimination/
// back-substitution (solver.tpp:2193-2356, n_off_diag = 0); only the rounding of the elimination differs
// (observed |dJ/J| < 1e-13, tests/test_oracle_golden.py::test_twosided_reordering_error).
//
// The common case -- the CloseLines window of a bin contains exactly the bin's own line -- is served from the
// phase-A record with four comparisons against the neighbouring sorted line centres (exactly the result of
// the reference's std::lower_bound/upper_bound, solver.tpp:1230-1233); anything else (overlapping lines,
// ApproximationType None, a OneLine bin whose corresponding_line is another line) takes the generic
// out-of-line path that loads and interpolates every line in the window like the reference does.
#pragma once
#include "mgb_kernels.cuh"

namespace mgb {

// chunk of consecutive frequency bins handled by one warp (built on the host, mgb_api.cu: build_freq_chunks)
struct FreqChunk {
    uint32_t f0, nf; // bins [f0, f0+nf), nf <= 32
    uint32_t l0, nl; // lines l0 .. l0+nl-1 own those bins (f = l*nquads + z), nl <= lpw
};

// Record buffer of one line: kRecFields rows of kRecStride doubles. Slot 0 carries the last record of the
// previous round (the far end X of the first segment of this round), slots 1..32 hold this round's entries.
constexpr int kRecStride = 33;       // 33 slots; odd stride, and consecutive lines start 20 banks apart
constexpr int kRecFields = 10;
enum { R_SHIFT = 0, R_DZ, R_IW, R_IE, R_IO, R_SLINE, R_DSM, R_SQW, R_FACTOR, R_PTS /* uint2 {nr, nr_interp} */ };
#define MGB_REC(q, fld) ((q)[(fld) * kRecStride])

constexpr int kWordsPerLine = 104;   // u32 per line: ni[32], excl[32], eoff[36], isub, last, pad
constexpr int kLineDoubles  = kRecFields * kRecStride + kWordsPerLine / 2;

// bytes of dynamic shared memory per warp
__host__ __device__ inline size_t solve_smem_per_warp(uint32_t lpw) { return 8 * ((size_t)lpw * kLineDoubles + 6 * 32 /* side-0 results */); }

// Per-warp shared-memory region, addressed with 32-bit indices into the block's dynamic shared array so that
// every record access is one LDS/STS with an immediate offset.
extern __shared__ __align__(16) double mgb_smem[];
struct WarpSmem {
    uint32_t base; // index (in doubles) of this warp's region
    uint32_t lpw;
    __device__ __forceinline__ double* rec(uint32_t k) const { return mgb_smem + base + k * (kRecFields * kRecStride); } // [kRecFields][kRecStride]
    __device__ __forceinline__ uint32_t* words(uint32_t k) const {
        return reinterpret_cast<uint32_t*>(mgb_smem + base + lpw * (kRecFields * kRecStride)) + k * kWordsPerLine;
    }
    __device__ __forceinline__ uint32_t* ni(uint32_t k) const { return words(k); }         // [32] sub-steps of step j-1-lane
    __device__ __forceinline__ uint32_t* excl(uint32_t k) const { return words(k) + 32; }  // [32] exclusive prefix of ni
    __device__ __forceinline__ uint32_t* eoff(uint32_t k) const { return words(k) + 64; }  // [S+1] entry offset of each step in the buffer
    __device__ __forceinline__ uint32_t* isub(uint32_t k) const { return words(k) + 100; } // sub-entries of step j-1 already emitted
    __device__ __forceinline__ uint32_t* last(uint32_t k) const { return words(k) + 101; } // entries of the previous round (slot to carry)
    __device__ __forceinline__ double* side0(unsigned lane) const { return mgb_smem + base + lpw * kLineDoubles + lane; } // [6][32]
};

// ---------------------------------------------------------------------------------------------------
// generic (out-of-line) paths: exactly the reference's loops over the lines in the window
// ---------------------------------------------------------------------------------------------------
template <int APPROX>
__device__ __noinline__ void eta_chi_generic(const DevModel& m, uint32_t p1, uint32_t p2, double factor, uint32_t ll, double freq,
                                             double& eta, double& chi) {
    Entry e;
    e.p1 = p1; e.p2 = p2; e.factor = factor; e.shift = 1.0;
    eta_chi_interp<APPROX>(m, e, ll, freq, eta, chi);
}

template <int APPROX>
__device__ __noinline__ void S_dtau_generic(const DevModel& m, uint32_t cp1, uint32_t cp2, double cfac, uint32_t np1, uint32_t np2,
                                            double nfac, uint32_t ll, double currfreq, double nextfreq, double dz, double& dtau,
                                            double& Scurr, double& Snext) {
    Entry c, n;
    c.p1 = cp1; c.p2 = cp2; c.factor = cfac; c.shift = 1.0;
    n.p1 = np1; n.p2 = np2; n.factor = nfac; n.shift = 1.0;
    S_dtau_line_integrated<APPROX>(m, c, n, ll, currfreq, nextfreq, dz, dtau, Scurr, Snext);
}

// per-lane constants of the bin's own line
struct LaneLine {
    double s_prev, s_own, s_next; // sorted line centres around the own line (-inf / +inf sentinels)
    bool own;                     // the record of line l may be used for this bin's eta/chi
};

// Classify the CloseLines window [left,right] against the sorted line list using only the own line's neighbours.
// With lo = lower_bound(left), hi = upper_bound(right) (solver.tpp:1230-1233) and pos = sorted position of the own line:
//   WIN_OWN   : lo == pos && hi == pos+1   (exactly the own line)
//   WIN_EMPTY : lo >= hi                   (no line: the bin is Doppler-shifted out of every line's reach)
//   WIN_OTHER : anything else (or NaN)  -> generic path
enum { WIN_OTHER = 0, WIN_OWN = 1, WIN_EMPTY = 2 };
__device__ __forceinline__ int window_class(const LaneLine& L, double left, double right) {
    const bool prev_below = L.s_prev < left, own_below = L.s_own < left;
    const bool own_above = right < L.s_own, next_above = right < L.s_next;
    const bool is_own   = L.own & prev_below & !own_below & !own_above & next_above;
    const bool is_empty = L.own & ((prev_below & own_above) | (own_below & next_above));
    return is_own ? WIN_OWN : (is_empty ? WIN_EMPTY : WIN_OTHER);
}

// compute_S_dtau_line_integrated for the segment between the records X (far) and Y (near) when the window holds
// exactly the own line (solver.tpp:1878-1912 OneLine, 2012-2111 CloseLines; compute_dtau_single_line :2374-2455).
// dtau is bit-symmetric under exchanging the two ends (differences and the float erf difference negate exactly,
// sums and min/max commute), so the roles are fixed here and the caller maps S_atX / S_atY onto the reference's
// Scurr / Snext according to the side. Out of line: it holds two inlined erff and is taken by a minority of segments.
struct LineIntegral { double dtau, S_atX, S_atY; }; /* returned by value: results stay in registers across the call */
template <int APPROX>
__device__ __noinline__ LineIntegral line_integrated_own(double min_opacity, double linefreq, double freqX, double freqY, double dz, double iwX,
                                                         double iwY, double ioX, double ioY, double slX, double slY) {
    double dtau, S_atX, S_atY;
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
        const float erfdiff   = erff_ool((float)posY) - erff_ool((float)posX); /* float subtraction */
        const double erfterm  = ddiv(avg_op, diff_pos) / 2.0 * (double)erfdiff;
        ld = dz * (avg_iw * erfterm + min_opacity);
    }
    if (APPROX == APPROX_ONELINE) { dtau = ld; S_atX = slX; S_atY = slY; return LineIntegral{dtau, S_atX, S_atY}; }
    const double sum_dtau = 0.0 + ld, sum_x = 0.0 + ld * slX, sum_y = 0.0 + ld * slY;
    double d = sum_dtau;
    const double bound_min_dtau = min_opacity * dz;
    if (-bound_min_dtau < d) d = (bound_min_dtau < d) ? d : bound_min_dtau;
    const Div2 dv = ddiv2(sum_x, sum_y, d); /* one reciprocal refinement for both ends */
    dtau = d; S_atX = dv.q1; S_atY = dv.q2;
    return LineIntegral{dtau, S_atX, S_atY};
}

struct SweepState {
    double chiX, SX;     // eta/chi evaluation at the far entry X of the open segment
    double dtau_far;     // optical depth of X's far segment
    double SatY_far;     // line-integrated source of that segment at its near end (= at X)
    double F, FI, a;     // Rybicki-Hummer F (FF or GG), 1/(1+F), eliminated source (Su)
    bool haveX, X_is_end, large_far;
    unsigned nlarge;     // segments of this side that took the line-integrated branch
};

// One entry of the sweep: evaluate eta, chi, S at the record Y = q[0] (get_eta_and_chi_interpolated, then S = eta/chi,
// solver.tpp:2166-2176), close the segment between the previous record X = q[-1] and Y (compute_source_dtau,
// :2139-2186) and eliminate X's row. side 0: reference orientation curr = X, next = Y.  side 1: curr = Y, next = X.
template <int APPROX>
__device__ __forceinline__ void sweep_step(const DevModel& m, const LaneLine& L, SweepState& st, int side, const double* q, uint32_t ll,
                                           double freq, unsigned& nlarge) {
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
            /* the reference accumulates into eta = 0, chi = min_opacity (CloseLines) or adds min_opacity last (OneLine) */
            const double eta = 0.0 + prof * MGB_REC(q, R_IE);
            if (APPROX == APPROX_ONELINE) chiY = (0.0 + prof * MGB_REC(q, R_IO)) + m.min_opacity;
            else chiY = m.min_opacity + prof * MGB_REC(q, R_IO);
            SY = ddiv(eta, chiY);
        } else if (cls == WIN_EMPTY) {
            chiY = m.min_opacity;
            SY   = zero_over(chiY); /* eta = 0 */
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
        // compute_source_dtau (solver.tpp:2139-2186)
        const double dshift_abs = fabs(shY - shX);
        const double dX = MGB_REC(x, R_DSM), dY = MGB_REC(q, R_DSM);
        const double dshift_max = (dY < dX) ? dY : dX;
        double dtau, S_atX = 0.0, S_atY = 0.0;
        const bool large = dshift_abs > dshift_max;
        if (large) {
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
                dtau = li.dtau; S_atX = li.S_atX; S_atY = li.S_atY;
            } else if (cls == WIN_EMPTY) {
                /* no line in reach: sums stay 0, dtau is bounded below by min_opacity*dz (solver.tpp:2095-2110) */
                double d = 0.0;
                const double bound_min_dtau = m.min_opacity * dzX;
                if (-bound_min_dtau < d) d = (bound_min_dtau < d) ? d : bound_min_dtau;
                dtau = d; S_atX = zero_over(d); S_atY = S_atX; /* (0 + 0) / d */
            } else {
                const uint2 pX = *reinterpret_cast<const uint2*>(&MGB_REC(x, R_PTS)), pY = *reinterpret_cast<const uint2*>(&MGB_REC(q, R_PTS));
                const double fX = MGB_REC(x, R_FACTOR), fY = MGB_REC(q, R_FACTOR);
                if (side == 0) S_dtau_generic<APPROX>(m, pX.x, pX.y, fX, pY.x, pY.y, fY, ll, frX, frY, dzX, dtau, S_atX, S_atY);
                else S_dtau_generic<APPROX>(m, pY.x, pY.y, fY, pX.x, pX.y, fX, ll, frY, frX, dzX, dtau, S_atY, S_atX);
            }
        } else {
            dtau = 0.5 * (st.chiX + chiY) * dzX;
        }
        if (st.X_is_end) {
            // boundary row (solver.tpp:2240-2262 for `first`, 2283-2297 mirrored for `last`)
            const double S0  = large ? S_atX : st.SX;
            const double inv = drcp(dtau);
            const double Cb  = 2.0 * inv * inv;
            const double Bm  = 1.0 + 2.0 * inv;
            const double B   = Bm + Cb;
            const uint2 pX   = *reinterpret_cast<const uint2*>(&MGB_REC(x, R_PTS));
            const double Ib  = boundary_intensity(m, pX.x, freq * shX);
            const double Su0 = ddiv(S0 + 2.0 * Ib * inv, B);
            st.F  = 0.5 * Bm * dtau * dtau;
            st.a  = Su0 * (1.0 + st.F); /* the sweep carries b = Su * (1 + F): one reciprocal per row instead of two */
        } else {
            // interior row (solver.tpp:2264-2281; mirrored 2306-2311): the row's source is the line-integrated one of the
            // segment in which X is `curr` (the near segment on side 0, the far one on side 1)
            const double Sx = side ? (st.large_far ? st.SatY_far : st.SX) : (large ? S_atX : st.SX);
            const double avg     = 0.5 * (st.dtau_far + dtau);
            const double invFar  = avg * st.dtau_far;
            const double invNear = avg * dtau;
            const double P       = drcp(invFar * (1.0 + st.F)); /* A * FI of the previous row */
            st.a  = (st.a * P + Sx) * invNear;
            st.F  = (st.F * P + 1.0) * invNear;
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
}

// Out-of-line instance of the general step (boundary rows, windows with other lines, ApproximationType None): taken by a
// small minority of the entries, kept out of the instruction stream of the common case.
template <int APPROX>
__device__ __noinline__ SweepState sweep_step_general(const DevModel& m, LaneLine L, SweepState st, int side, const double* q, uint32_t ll,
                                                      double freq) {
    sweep_step<APPROX>(m, L, st, side, q, ll, freq, st.nlarge);
    return st;
}

// The common case of sweep_step, arithmetic unchanged: an interior row whose windows (at Y, and over the segment when the
// shift is large) hold exactly the own line or no line at all. `part` = this lane has an entry in this iteration.
// Returns false (nothing done) when some participating lane of the warp needs the general step.
template <int APPROX>
__device__ __forceinline__ bool sweep_step_common(const DevModel& m, const LaneLine& L, SweepState& st, int side, const double* q, double freq,
                                                  bool part) {
    if (APPROX == APPROX_NONE) return false;
    const double* x  = q - 1;
    const double shY = MGB_REC(q, R_SHIFT), shX = MGB_REC(x, R_SHIFT);
    const double frY = freq * shY;
    int clsY;
    if (APPROX == APPROX_CLOSELINES) {
        const double ub = m.max_dist * (frY * MGB_REC(q, R_SQW));
        clsY = window_class(L, frY - ub, frY + ub);
    } else {
        clsY = L.own ? WIN_OWN : WIN_OTHER;
    }
    const double dshift_abs = fabs(shY - shX);
    const double dX = MGB_REC(x, R_DSM), dY = MGB_REC(q, R_DSM);
    const double dshift_max = (dY < dX) ? dY : dX;
    const bool large = dshift_abs > dshift_max;
    const double frX = freq * shX;
    int clsS = WIN_OWN;
    if (APPROX == APPROX_CLOSELINES && __any_sync(0xffffffffu, large)) { /* the segment's window matters to large shifts only */
        const double left_freq = (frX < frY) ? frX : frY, right_freq = (frX < frY) ? frY : frX;
        const double bX = m.max_dist * (right_freq * MGB_REC(x, R_SQW));
        const double bY = m.max_dist * (right_freq * MGB_REC(q, R_SQW));
        const double ub = (bX < bY) ? bY : bX;
        clsS = window_class(L, left_freq - ub, right_freq + ub);
    } else if (APPROX != APPROX_CLOSELINES) {
        clsS = clsY;
    }
    const bool ok = !part || (st.haveX & !st.X_is_end & (clsY != WIN_OTHER) & (!large | (clsS != WIN_OTHER)));
    if (!__all_sync(0xffffffffu, ok)) return false;
    if (!part) return true;
    const double dzX = MGB_REC(x, R_DZ);
    double chiY, SY;
    if (clsY == WIN_OWN) {
        const double prof = gaussian(MGB_REC(q, R_IW), frY - L.s_own);
        const double eta  = 0.0 + prof * MGB_REC(q, R_IE);
        if (APPROX == APPROX_ONELINE) chiY = (0.0 + prof * MGB_REC(q, R_IO)) + m.min_opacity;
        else chiY = m.min_opacity + prof * MGB_REC(q, R_IO);
        SY = ddiv(eta, chiY);
    } else {
        chiY = m.min_opacity;
        SY   = zero_over(chiY);
    }
    double dtau, S_atX = 0.0, S_atY = 0.0;
    if (large) {
        st.nlarge++;
        if (clsS == WIN_OWN) {
            const LineIntegral li = line_integrated_own<APPROX>(m.min_opacity, L.s_own, frX, frY, dzX, MGB_REC(x, R_IW), MGB_REC(q, R_IW),
                                                                MGB_REC(x, R_IO), MGB_REC(q, R_IO), MGB_REC(x, R_SLINE), MGB_REC(q, R_SLINE));
            dtau = li.dtau; S_atX = li.S_atX; S_atY = li.S_atY;
        } else {
            double d = 0.0;
            const double bound_min_dtau = m.min_opacity * dzX;
            if (-bound_min_dtau < d) d = (bound_min_dtau < d) ? d : bound_min_dtau;
            dtau = d; S_atX = zero_over(d); S_atY = S_atX; /* (0 + 0) / d */
        }
    } else {
        dtau = 0.5 * (st.chiX + chiY) * dzX;
    }
    const double Sx      = side ? (st.large_far ? st.SatY_far : st.SX) : (large ? S_atX : st.SX);
    const double avg     = 0.5 * (st.dtau_far + dtau);
    const double invFar  = avg * st.dtau_far;
    const double invNear = avg * dtau;
    const double P       = drcp_inl(invFar * (1.0 + st.F)); /* A * FI of the previous row */
    st.a  = (st.a * P + Sx) * invNear;
    st.F  = (st.F * P + 1.0) * invNear;
    st.dtau_far  = dtau;
    st.large_far = large;
    st.SatY_far  = S_atY;
    st.chiX = chiY; st.SX = SY;
    return true;
}

// ---------------------------------------------------------------------------------------------------
// phase A: one record = everything of a ray entry that does not depend on the frequency bin
// (Solver::set_data_for_line, solver.tpp:1057-1097, fused with the interpolation part of
// get_eta_and_chi_interpolated, :1259-1302, and the line source function of :1888-1911)
// ---------------------------------------------------------------------------------------------------
struct StepData { // one geometric step crt -> nxt of the cached path (independent of the line)
    uint32_t nxt, crt;
    double dZ, sh_nxt, sh_crt, dsm, sqw;
};

__device__ __forceinline__ void make_record(const DevModel& m, double* r /* record slot */, uint32_t l, const StepData& sd, uint32_t i,
                                            uint32_t ni) {
    double dz = sd.dZ, factor = 0.0, shift = sd.sh_nxt;
    if (ni > 1) {
        dz = sd.dZ / ni;
        const double dshi = (sd.sh_nxt - sd.sh_crt) / ni;
        factor = 1.0 - ((double)i) / ni;
        shift  = sd.sh_crt + i * dshi;
    }
    const LineRow a = load_row(m, sd.nxt, l);
    double iw = a.inverse_width, ie = a.emissivity, io = a.opacity;
    double sline = ddiv(a.emissivity, a.opacity);
    if (factor != 0.0) {
        const LineRow b = load_row(m, sd.crt, l);
        iw    = interp_linear(a.inverse_width, b.inverse_width, factor);
        ie    = interp_log(a.emissivity, b.emissivity, factor);
        io    = interp_log(a.opacity, b.opacity, factor);
        sline = interp_log(sline, ddiv(b.emissivity, b.opacity), factor);
    }
    MGB_REC(r, R_SHIFT)  = shift;
    MGB_REC(r, R_DZ)     = dz;
    MGB_REC(r, R_IW)     = iw;
    MGB_REC(r, R_IE)     = ie;
    MGB_REC(r, R_IO)     = io;
    MGB_REC(r, R_SLINE)  = sline;
    MGB_REC(r, R_DSM)    = sd.dsm;
    MGB_REC(r, R_SQW)    = sd.sqw;
    MGB_REC(r, R_FACTOR) = factor;
    *reinterpret_cast<uint2*>(&MGB_REC(r, R_PTS)) = make_uint2(sd.nxt, sd.crt);
}

// Before a round overwrites the buffers: keep each line's last record in slot 0 (it is the far end of the next segment).
__device__ __forceinline__ void carry_last_records(const WarpSmem& ws, uint32_t nl, unsigned lane) {
    for (uint32_t k = 0; k < nl; k++) {
        const uint32_t last = *ws.last(k);
        if (last && lane < kRecFields) {
            double* row = ws.rec(k) + lane * kRecStride;
            row[0] = row[last];
        }
    }
    __syncwarp();
}

// Expand the next geometric steps j-1, j-2, ... of the half-ray for all nl lines of the warp. Returns the number S of
// steps whose entries are now in the record buffers (eoff(k)[0..S] = entry offsets per step of line k; entry e sits in
// slot 1+e); all lines advance by the SAME steps so that phase B runs in lock-step over segments (uniform
// large/small-shift branch). S is limited by the 32 entry slots per line. A single step with more than 32 sub-steps on
// some line is emitted 32 entries at a time (isub cursors) and reported as S = 1 with j unchanged until every line is through.
__device__ __forceinline__ uint32_t produce_window(const DevModel& m, const PathView& pv, const WarpSmem& ws, uint32_t l0, uint32_t nl,
                                                   uint32_t o, unsigned long long off, double shift0, uint32_t& j, unsigned lane) {
    // ---- lanes = steps ----
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
            const uint32_t total = __shfl_sync(0xffffffffu, excl + ni, S - 1); /* <= 32 entries in this window */
            // ---- entry e -> the step that owns it: scatter step ids at their first entry, then a running maximum ----
            uint32_t* mark = ws.eoff(k); /* eoff doubles as the scatter array before it receives the offsets */
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
            // ---- lanes = entries ----
            if (lane < total) make_record(m, ws.rec(k) + 1 + lane, l0 + k, se, ni_e - (lane - excl_e), ni_e);
            if (lane < S) {
                ws.eoff(k)[lane] = excl;
                if (lane == S - 1) { ws.eoff(k)[S] = excl + ni; *ws.last(k) = excl + ni; }
            }
        }
        j -= S;
        return S;
    }
    // ---- a step with more than 32 sub-steps on some line: lanes = sub-entries of step j-1 ----
    StepData s0;
    s0.nxt = __shfl_sync(0xffffffffu, sd.nxt, 0); s0.crt = __shfl_sync(0xffffffffu, sd.crt, 0);
    s0.dZ = __shfl_sync(0xffffffffu, sd.dZ, 0); s0.sh_nxt = __shfl_sync(0xffffffffu, sd.sh_nxt, 0);
    s0.sh_crt = __shfl_sync(0xffffffffu, sd.sh_crt, 0); s0.dsm = __shfl_sync(0xffffffffu, sd.dsm, 0);
    s0.sqw = __shfl_sync(0xffffffffu, sd.sqw, 0);
    bool all_done = true;
    for (uint32_t k = 0; k < nl; k++) {
        const uint32_t ni = ws.ni(k)[0], is = *ws.isub(k);
        const uint32_t rem = ni - is, nemit = rem < 32u ? rem : 32u;
        if (lane < nemit) make_record(m, ws.rec(k) + 1 + lane, l0 + k, s0, ni - (is + lane), ni);
        __syncwarp();
        if (lane == 0) { ws.eoff(k)[0] = 0; ws.eoff(k)[1] = nemit; *ws.isub(k) = is + nemit; *ws.last(k) = nemit; }
        all_done = all_done && (is + nemit == ni);
    }
    if (all_done) {
        __syncwarp();
        if (lane < nl) *ws.isub(lane) = 0;
        j -= 1;
    }
    return 1;
}

// The origin's own entry (nr_[centre] = o, shift_[centre] = 1; solver.tpp:484-485) as the single record of the last round.
__device__ __forceinline__ void produce_centre(const DevModel& m, const WarpSmem& ws, uint32_t l0, uint32_t nl, uint32_t o, unsigned lane) {
    if (lane < nl) {
        StepData sd;
        sd.nxt = o; sd.crt = o; sd.dZ = 0.0; sd.sh_nxt = 1.0; sd.sh_crt = 1.0;
        sd.dsm = __ldg(&m.dshift_max[o]); sd.sqw = __ldg(&m.sqrtw_max[o]);
        make_record(m, ws.rec(lane) + 1, l0 + lane, sd, 1, 1);
        ws.eoff(lane)[0] = 0; ws.eoff(lane)[1] = 1; *ws.last(lane) = 1;
    }
}

// ---------------------------------------------------------------------------------------------------
// History of the rows eliminated last on one side (nearest the origin first when read back): what the banded Lambda operator
// needs of the entries within n_off_diag of the origin. Ring buffers in local memory; only the OFFD instantiation has them.
struct SideHistory {
    static constexpr int R = kMaxOffDiag + 2;
    double F[R], FI[R], D[R], sh[R]; // FF/FI (side 0) or GG/GI (side 1) of the row, dtau of its segment towards the origin, shift
    uint32_t nr[R];
    unsigned cnt;
    __device__ __forceinline__ void push(double f, double fi, double d, uint32_t p, double s) {
        const unsigned i = cnt % R;
        F[i] = f; FI[i] = fi; D[i] = d; nr[i] = p; sh[i] = s;
        cnt++;
    }
    __device__ __forceinline__ unsigned slot(unsigned j) const { return (cnt - 1u - j) % R; } // j-th nearest to the origin
};

// CTA shape of k_solve: kSolveWarps warps, kSolveCtas CTAs per SM (register cap 65536 / (32 * warps * ctas)). Work items differ a lot
// in length and a CTA keeps its shared memory until its slowest warp is done: 8 CTAs of 2 warps beat 4 of 4 by 1.6 % (16 warps per SM
// either way; 8 x (2 x 13760 B + 256 B table + 1 KB reserved) just fits the 228 KB of an SM, which is why the record stride is 33).
#ifndef MGB_SOLVE_WARPS
#define MGB_SOLVE_WARPS 2
#endif
#ifndef MGB_SOLVE_CTAS
#define MGB_SOLVE_CTAS 8
#endif
constexpr int kSolveWarps = MGB_SOLVE_WARPS, kSolveCtas = MGB_SOLVE_CTAS;

template <int APPROX, bool OFFD>
__global__ void __launch_bounds__(32 * kSolveWarps, OFFD ? kSolveCtas / 2 : kSolveCtas) k_solve(const __grid_constant__ DevModel m, const __grid_constant__ PathView pv, ScratchView sc,
                                                  const FreqChunk* __restrict__ chunks, uint32_t nchunks_f, uint32_t lpw) {
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
    const uint32_t l   = f / m.nquads;       // line of this frequency bin (f = l*nquads + z, model.cpp:127-146)
    const uint32_t z   = f - l * m.nquads;
    const uint32_t ks  = l - ch.l0;          // record buffer of this lane's line
    const uint32_t ll  = (APPROX == APPROX_ONELINE) ? __ldg(&m.corr_line[f]) : l;
    const uint32_t ar  = __ldg(&m.antipod[rr]);
    const double freq  = __ldg(&m.nu[(size_t)o * m.nfreqs + f]);
    const size_t out   = ((size_t)rr * pv.no + ol) * m.nfreqs + f;

    const uint32_t len_b = __ldg(&pv.len[(size_t)o * m.nrays + rr]);
    const uint32_t len_f = __ldg(&pv.len[(size_t)o * m.nrays + ar]);
    if (len_b + len_f == 0) {
        // isolated origin (solver.tpp:510-518)
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
    double F1 = 0, FI1 = 0, a1 = 0, dt1 = 0, Sc1 = 0, SC = 0;
    bool lg1 = false;
    {   /* side-0 results wait in shared memory while side 1 runs */
        double* s0 = ws.side0(lane);
        s0[0] = 0.0; s0[32] = 0.0; s0[64] = 0.0; s0[96] = 0.0; s0[128] = 0.0; s0[160] = 0.0;
    }
    const double* rb   = ws.rec(ks) + 1; /* entry e of this lane's line */
    const uint32_t* eo = ws.eoff(ks);

    SideHistory hist[OFFD ? 2 : 1];
    if (OFFD) { hist[0].cnt = 0; hist[1].cnt = 0; }
    for (int side = 0; side < 2; side++) {
        const uint32_t r   = side ? ar : rr;
        const uint32_t len = side ? len_f : len_b;
        if (len == 0) continue;
        const size_t hr = (size_t)ol * m.nrays + r;
        const unsigned long long off = __ldg(&pv.poff[hr]);
        const double shift0 = __ldg(&pv.shift0[hr]);
        SweepState st;
        st.haveX = false; st.X_is_end = false; st.large_far = false; st.nlarge = 0;
        st.chiX = st.SX = st.F = st.FI = st.a = st.dtau_far = st.SatY_far = 0.0;
        __syncwarp();
        if (lane < ch.nl) { *ws.isub(lane) = 0; *ws.last(lane) = 0; }
        __syncwarp();
        uint32_t j = len; /* geometric steps not yet expanded (warp-uniform) */
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
                const uint32_t nmax = __reduce_max_sync(0xffffffffu, ns);
                for (uint32_t it = 0; it < nmax; it++) {
                    const bool part = it < ns;
                    const double* q = rb + e0 + (part ? it : 0u);
                    const bool hadX = st.haveX;
                    if (!sweep_step_common<APPROX>(m, L, st, side, q, freq, part)) {
                        if (part) st = sweep_step_general<APPROX>(m, L, st, side, q, ll, freq);
                    }
                    if (OFFD && part && hadX) { /* the row of X = q[-1] has just been eliminated */
                        const uint2 pX = *reinterpret_cast<const uint2*>(&MGB_REC(q - 1, R_PTS));
                        hist[OFFD ? side : 0].push(st.F, drcp(1.0 + st.F), st.dtau_far, pX.x, MGB_REC(q - 1, R_SHIFT));
                    }
                }
            }
            __syncwarp();
            if (centre) break;
        }
        SC = st.SX; /* source function at the origin */
        st.FI = drcp(1.0 + st.F); /* FI and Su of the row next to the origin, from the carried b = Su * (1 + F) */
        st.a  = st.a * st.FI;
        nlarge += st.nlarge;
        if (side == 0) {
            double* s0 = ws.side0(lane);
            s0[0] = st.F; s0[32] = st.FI; s0[64] = st.a; s0[96] = st.dtau_far; s0[128] = st.SatY_far; s0[160] = st.large_far ? 1.0 : 0.0;
        } else { F1 = st.F; FI1 = st.FI; a1 = st.a; dt1 = st.dtau_far; lg1 = st.large_far; Sc1 = st.SatY_far; }
    }

    // ---- centre row ----
    double u, Ldiag;
    const double* s0 = ws.side0(lane);
    const double F0 = s0[0], FI0 = s0[32], a0 = s0[64], dt0 = s0[96], Sn0 = s0[128];
    const bool lg0 = s0[160] != 0.0;
    if (len_b && len_f) {
        const double S_c     = lg1 ? Sc1 : SC;
        const double avg     = 0.5 * (dt0 + dt1);
        const double invA    = avg * dt0, invC = avg * dt1;
        const double A = drcp(invA), Cc = drcp(invC); /* Newton reciprocal / division: the bits of 1.0 / x and a / b without nvcc's slow path */
        const double FFc = (A * F0 * FI0 + 1.0) * invC;
        const double GP  = F1 * FI1;
        Ldiag = ddiv(invC, FFc + GP);
        u     = (S_c + A * a0 + Cc * a1) * Ldiag;
    } else if (len_f) {
        // centre == first (solver.tpp:2240-2262 with the back-substitution folded in)
        const double S_c = lg1 ? Sc1 : SC;
        const double inv = drcp(dt1);
        const double Cf  = 2.0 * inv * inv;
        const double Bm  = 1.0 + 2.0 * inv;
        const double Ib  = boundary_intensity(m, o, freq);
        const double FFf = 0.5 * Bm * dt1 * dt1;
        const double GP  = F1 * FI1;
        Ldiag = ddiv(drcp(Cf), FFf + GP);
        u     = (S_c + 2.0 * Ib * inv + Cf * a1) * Ldiag;
    } else {
        // centre == last (solver.tpp:2283-2297, 2317)
        const double S_c = lg0 ? Sn0 : SC;
        const double inv = drcp(dt0);
        const double Al  = 2.0 * inv * inv;
        const double Bm  = 1.0 + 2.0 * inv;
        const double Bl  = Bm + Al;
        const double Ib  = boundary_intensity(m, o, freq);
        Ldiag = ddiv(1.0 + F0, Bm + Bl * F0);
        u     = (Al * a0 + (S_c + 2.0 * Ib * inv)) * Ldiag;
    }

    if (!OFFD) {
        /* diagonal operator: L_diag leaves the kernel; the Lambda increment's factors (eta/chi and the line profile AT THE ORIGIN,
           update_Lambda solver.tpp:1760-1806) depend on (o, f) only and are applied once per (o, f) by k_reduce_lambda2 instead of
           once per (o, rr, f) here -- they were 20 % of this kernel's instructions (profiles/ncu_k_solve_r1g_c3.txt: tail + device other) */
        if (active) { sc.u[out] = u; sc.L[out] = Ldiag; }
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
        return;
    }
    // ---- Lambda increment (update_Lambda, solver.tpp:1760-1806) of the banded operator ----
    const double w_ang     = 2.0 * ray_weight(m, o, rr); /* get_weight<adaptive>(nr[centre], rr), solver.tpp:1777 */
    const double freq_line = __ldg(&m.line[l]);
    const double constante = __ldg(&m.line_A[l]) * __ldg(&m.line_quadw[(size_t)l * m.nquads + z]) * w_ang;
    double eta_l, chi_l;
    // NOTE: the reference passes the within-species transition index k as OneLine's line index here (solver.tpp:1798)
    eta_chi_generic<APPROX>(m, o, o, 0.0, __ldg(&m.line_k[l]), freq_line, eta_l, chi_l);
    const double inverse_chi = 1.0 / chi_l;
    const double frq = freq * 1.0; /* nu(o,f) * shift[centre] */
    const double phi = thermo_profile(m, __ldg(&m.line_invmass[l]), o, freq_line, frq);
    const double Lam = constante * frq * phi * Ldiag * inverse_chi;

    if (active) { sc.u[out] = u; sc.L[out] = Lam; }

    if (OFFD) {
        // ---- off-diagonal elements of the banded Lambda operator (solver.tpp:2320-2355 and update_Lambda :1806-1850) ----
        // Index i counts entries from the origin: i < 0 towards `first`, i > 0 towards `last`; n0 / n1 entries on the two sides.
        // FF, FI of i < 0 and GG, GI of i > 0 are the rows the two sweeps eliminated last (hist); the forward recursion is continued
        // across the origin to i = K and the backward one to i = -K+1 with the reference's formulas, then
        //   L_lower(m, -m-1) = L_diag[-m-1] * GI[-m] * ... * GI[0],   L_upper(m, m+1) = L_diag[m+1] * FI[m] * ... * FI[0].
        constexpr int KM = kMaxOffDiag;
        const int K  = (int)sc.K;
        const int n0 = (int)hist[0].cnt, n1 = (int)hist[OFFD ? 1 : 0].cnt;
        const SideHistory& H0 = hist[0];
        const SideHistory& H1 = hist[OFFD ? 1 : 0];
        auto seg = [&](int i) -> double { return i < 0 ? H0.D[H0.slot((unsigned)(-i - 1))] : H1.D[H1.slot((unsigned)i)]; }; /* dtau(i, i+1) */
        double FFw[2 * KM + 3], FIw[2 * KM + 3], GGw[2 * KM + 3], GIw[2 * KM + 3]; /* index i + KM + 1 */
        for (int i = -1; i >= -K - 1 && i >= -n0; i--) { FFw[i + KM + 1] = H0.F[H0.slot((unsigned)(-i - 1))]; FIw[i + KM + 1] = H0.FI[H0.slot((unsigned)(-i - 1))]; }
        for (int i = 1; i <= K + 1 && i <= n1; i++) { GGw[i + KM + 1] = H1.F[H1.slot((unsigned)(i - 1))]; GIw[i + KM + 1] = H1.FI[H1.slot((unsigned)(i - 1))]; }
        for (int i = 0; i <= K && i < n1; i++) { /* FF exists for first <= n < last */
            double FFi;
            if (i == -n0) { /* the origin is `first` */
                const double dt = seg(0), inv = 1.0 / dt;
                const double Bm = 1.0 + 2.0 * inv;
                FFi = 0.5 * Bm * dt * dt;
            } else {
                const double avg = 0.5 * (seg(i - 1) + seg(i));
                const double invA = avg * seg(i - 1), invC = avg * seg(i);
                const double A = 1.0 / invA;
                FFi = (A * FFw[i - 1 + KM + 1] * FIw[i - 1 + KM + 1] + 1.0) * invC;
            }
            FFw[i + KM + 1] = FFi; FIw[i + KM + 1] = 1.0 / (1.0 + FFi);
        }
        for (int i = 0; i >= -K + 1 && i > -n0; i--) { /* GG exists for first < n <= last */
            double GGi;
            if (i == n1) { /* the origin is `last` */
                const double dt = seg(-1), inv = 1.0 / dt;
                const double Bm = 1.0 + 2.0 * inv;
                GGi = 0.5 * Bm * dt * dt;
            } else {
                const double avg = 0.5 * (seg(i - 1) + seg(i));
                const double invA = avg * seg(i - 1), invC = avg * seg(i);
                const double C = 1.0 / invC;
                GGi = (C * (GGw[i + 1 + KM + 1] * GIw[i + 1 + KM + 1]) + 1.0) * invA;
            }
            GGw[i + KM + 1] = GGi; GIw[i + KM + 1] = 1.0 / (1.0 + GGi);
        }
        auto Ldiag_at = [&](int n) -> double {
            if (n == -n0) { /* first (solver.tpp:2335) */
                const double inv = 1.0 / seg(n);
                const double Cf = 2.0 * inv * inv, Bm = 1.0 + 2.0 * inv, Bf = Bm + Cf;
                return (1.0 + GGw[n + 1 + KM + 1]) / (Bm + Bf * GGw[n + 1 + KM + 1]);
            }
            if (n == n1) { /* last (solver.tpp:2324) */
                const double inv = 1.0 / seg(n - 1);
                const double Al = 2.0 * inv * inv, Bm = 1.0 + 2.0 * inv, Bl = Bm + Al;
                return (1.0 + FFw[n - 1 + KM + 1]) / (Bm + Bl * FFw[n - 1 + KM + 1]);
            }
            const double avg = 0.5 * (seg(n - 1) + seg(n));
            const double invC = avg * seg(n);
            return invC / (FFw[n + KM + 1] + GGw[n + 1 + KM + 1] * GIw[n + 1 + KM + 1]);
        };
        auto element = [&](uint32_t p, double shift_n, double Lband) -> double { /* update_Lambda :1816-1826 / :1838-1845 */
            double eta_n, chi_n;
            eta_chi_generic<APPROX>(m, p, p, 0.0, __ldg(&m.line_k[l]), freq_line, eta_n, chi_n);
            const double ichi = 1.0 / chi_n;
            const double fq   = __ldg(&m.nu[(size_t)p * m.nfreqs + f]) * shift_n;
            const double ph   = thermo_profile(m, __ldg(&m.line_invmass[l]), p, freq_line, fq);
            return constante * fq * ph * Lband * ichi;
        };
        double lowprod = 1.0, upprod = 1.0;
        for (int mm = 0; mm < K; mm++) {
            double vl = 0.0, vu = 0.0;
            uint32_t pl = 0xffffffffu, pu = 0xffffffffu;
            if (n0 >= mm + 1) {
                const int n = -mm - 1;
                lowprod *= GIw[n + 1 + KM + 1]; /* GI[-m] joins the product */
                const unsigned sl = H0.slot((unsigned)mm);
                pl = H0.nr[sl];
                vl = element(pl, H0.sh[sl], Ldiag_at(n) * lowprod);
            }
            if (n1 >= mm + 1) {
                const int n = mm + 1;
                upprod *= FIw[n - 1 + KM + 1]; /* FI[m] joins the product */
                const unsigned su = H1.slot((unsigned)mm);
                pu = H1.nr[su];
                vu = element(pu, H1.sh[su], Ldiag_at(n) * upprod);
            }
            if (active) {
                sc.offV[(out * K + mm) * 2] = vl; sc.offN[(out * K + mm) * 2] = pl;
                sc.offV[(out * K + mm) * 2 + 1] = vu; sc.offN[(out * K + mm) * 2 + 1] = pu;
            }
        }
    }

    // work counters: Sigma N counts one ray per (o, rr, line); large-shift evaluations per lane
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
// J(o,f) = sum_rr (u * 2) * w[rr] in rr order (solver.tpp:504-506, 515-517); optional copy of u into the full tensor.
__global__ void k_reduce_J(DevModel m, uint32_t o0, uint32_t no, const double* __restrict__ u, double* __restrict__ J,
                           double* __restrict__ u_full, uint32_t u_o0, uint32_t u_no) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)no * m.nfreqs) return;
    const uint32_t ol = (uint32_t)(i / m.nfreqs), f = (uint32_t)(i % m.nfreqs);
    double acc = 0.0;
    for (uint32_t rr = 0; rr < m.hnrays; rr++) {
        const double v = u[((size_t)rr * no + ol) * m.nfreqs + f];
        acc += v * 2.0 * ray_weight(m, o0 + ol, rr);
        if (u_full) u_full[((size_t)rr * u_no + (o0 + ol - u_o0)) * m.nfreqs + f] = v;
    }
    J[(size_t)(o0 + ol) * m.nfreqs + f] = acc;
}

// The memory-sparse solver's accumulator (Solver::solve_feautrier_order_2_sparse, solver.tpp:297-363): J(o, line) integrated over the
// line's quadrature, lspec.J(o,k) += quadrature.weights[z] * wt * Su[centre] in the loop nest rr > k > z, wt = 2 * weight(rr); the
// boundary intensity stands in for an isolated origin (scratch u holds it).
__global__ void k_reduce_J_lines(DevModel m, uint32_t o0, uint32_t no, const double* __restrict__ u, double* __restrict__ J_lines) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)no * m.nlines) return;
    const uint32_t ol = (uint32_t)(i / m.nlines), l = (uint32_t)(i % m.nlines);
    double acc = 0.0;
    for (uint32_t rr = 0; rr < m.hnrays; rr++) {
        const double wt = ray_weight(m, o0 + ol, rr) * 2.0;
        for (uint32_t z = 0; z < m.nquads; z++)
            acc += __ldg(&m.line_quadw[(size_t)l * m.nquads + z]) * wt * u[((size_t)rr * no + ol) * m.nfreqs + (size_t)l * m.nquads + z];
    }
    J_lines[(size_t)(o0 + ol) * m.nlines + l] = acc;
}

// Solver::solve_feautrier_order_2_anis (solver.tpp:365-456): per (o, line) the isotropic J and the five elements of the anisotropic
// radiation-field tensor, du = w_z * weight(o, rr) * u, J += 2 du, J2_x += wt_x(n(o, rr)) * du, loop nest rr > line > z.
// out [6][npoints][nlines]: J, J2_0, J2_1_Re, J2_1_Im, J2_2_Re, J2_2_Im.
__global__ void k_reduce_anis(DevModel m, uint32_t o0, uint32_t no, const double* __restrict__ u, double* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)no * m.nlines) return;
    const uint32_t ol = (uint32_t)(i / m.nlines), l = (uint32_t)(i % m.nlines);
    const uint32_t o  = o0 + ol;
    const double sqrt3 = sqrt(3.0), inv_sqrt2 = 1.0 / sqrt(2.0); /* tools/types.hpp:46-48 */
    double J = 0.0, J20 = 0.0, J21r = 0.0, J21i = 0.0, J22r = 0.0, J22i = 0.0;
    for (uint32_t rr = 0; rr < m.hnrays; rr++) {
        const size_t ri  = ray_index(m, o, rr);
        const double wt  = __ldg(&m.weight[ri]);
        const double nx = __ldg(&m.dir[3 * ri]), ny = __ldg(&m.dir[3 * ri + 1]), nz = __ldg(&m.dir[3 * ri + 2]);
        const double wt_0    = inv_sqrt2 * (3.0 * nz * nz - 1.0);
        const double wt_1_Re = -sqrt3 * nx * nz;
        const double wt_1_Im = -sqrt3 * ny * nz;
        const double wt_2_Re = 0.5 * sqrt3 * (nx * nx - ny * ny);
        const double wt_2_Im = sqrt3 * nx * ny;
        for (uint32_t z = 0; z < m.nquads; z++) {
            const double du = __ldg(&m.line_quadw[(size_t)l * m.nquads + z]) * wt * u[((size_t)rr * no + ol) * m.nfreqs + (size_t)l * m.nquads + z];
            J += 2.0 * du; J20 += wt_0 * du; J21r += wt_1_Re * du; J21i += wt_1_Im * du; J22r += wt_2_Re * du; J22i += wt_2_Im * du;
        }
    }
    const size_t plane = (size_t)m.npoints * m.nlines, at = (size_t)o * m.nlines + l;
    out[at] = J; out[plane + at] = J20; out[2 * plane + at] = J21r; out[3 * plane + at] = J21i; out[4 * plane + at] = J22r; out[5 * plane + at] = J22i;
}

// Lambda(o, lid) = sum over rr, then z, of the increments, in the order Lambda::add_element saw them
// (solver.tpp:476-508 loop nest rr > k > z; lambda.tpp:79-89).
__global__ void k_reduce_lambda(DevModel m, uint32_t o0, uint32_t no, const uint32_t* __restrict__ lens, const double* __restrict__ L,
                                double* __restrict__ lam_val, uint32_t* __restrict__ lam_cnt) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)no * m.nlines) return;
    const uint32_t ol = (uint32_t)(i / m.nlines), l = (uint32_t)(i % m.nlines);
    const uint32_t o  = o0 + ol;
    double acc = 0.0;
    uint32_t cnt = 0;
    for (uint32_t rr = 0; rr < m.hnrays; rr++) {
        const uint32_t ar = __ldg(&m.antipod[rr]);
        if (lens[(size_t)o * m.nrays + rr] + lens[(size_t)o * m.nrays + ar] == 0) continue; /* n_tot <= 1: no add_element */
        for (uint32_t z = 0; z < m.nquads; z++) {
            const double v = L[((size_t)rr * no + ol) * m.nfreqs + (size_t)l * m.nquads + z];
            if (cnt == 0) acc = v; else acc += v;
            cnt = 1;
        }
    }
    lam_val[(size_t)o * m.nlines + l] = acc;
    lam_cnt[(size_t)o * m.nlines + l] = cnt;
}

// Banded operator: the (p = o, line) list of (nr, Ls) built exactly as Lambda::add_element does (lambda.tpp:79-89) while
// update_Lambda runs over rr > z > (diagonal, then m = 0..K-1: lower, upper) (solver.tpp:476-508, 1806-1850).
__global__ void k_reduce_lambda_banded(DevModel m, uint32_t o0, uint32_t no, uint32_t K, uint32_t cap, const uint32_t* __restrict__ lens,
                                       const double* __restrict__ L, const double* __restrict__ offV, const uint32_t* __restrict__ offN,
                                       double* __restrict__ list_val, uint32_t* __restrict__ list_nr, double* __restrict__ lam_val,
                                       uint32_t* __restrict__ lam_cnt, unsigned long long* __restrict__ overflow) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)no * m.nlines) return;
    const uint32_t ol = (uint32_t)(i / m.nlines), l = (uint32_t)(i % m.nlines);
    const uint32_t o  = o0 + ol;
    double* lv   = list_val + ((size_t)o * m.nlines + l) * cap;
    uint32_t* ln = list_nr + ((size_t)o * m.nlines + l) * cap;
    uint32_t cnt = 0;
    auto add = [&](uint32_t nr_new, double Ls_new) {
        for (uint32_t k = 0; k < cnt; k++)
            if (ln[k] == nr_new) { lv[k] += Ls_new; return; }
        if (cnt >= cap) { atomicAdd(overflow, 1ull); return; }
        lv[cnt] = Ls_new; ln[cnt] = nr_new; cnt++;
    };
    for (uint32_t rr = 0; rr < m.hnrays; rr++) {
        const uint32_t ar = __ldg(&m.antipod[rr]);
        if (lens[(size_t)o * m.nrays + rr] + lens[(size_t)o * m.nrays + ar] == 0) continue;
        for (uint32_t z = 0; z < m.nquads; z++) {
            const size_t idx = ((size_t)rr * no + ol) * m.nfreqs + (size_t)l * m.nquads + z;
            add(o, L[idx]);
            for (uint32_t mm = 0; mm < K; mm++)
                for (uint32_t s = 0; s < 2; s++) {
                    const uint32_t p = offN[(idx * K + mm) * 2 + s];
                    if (p != 0xffffffffu) add(p, offV[(idx * K + mm) * 2 + s]);
                }
        }
    }
    for (uint32_t k = cnt; k < cap; k++) { lv[k] = 0.0; ln[k] = 0xffffffffu; }
    lam_val[(size_t)o * m.nlines + l] = cnt ? lv[0] : 0.0; /* the diagonal element is always added first */
    lam_cnt[(size_t)o * m.nlines + l] = cnt;
}

} // namespace mgb
