// mgb_solve.cuh -- k_solve: eta/chi assembly + second-order Feautrier solve, lanes = frequency bins.
//
// One warp owns (origin o, ray pair rr, 32 consecutive frequency bins). Every lane walks the cached
// geometric path of both half-rays from the far (boundary) end towards the origin, expands each step
// into the reference's sub-steps for ITS line (Solver::set_data_for_line, solver.tpp:1057-1097),
// evaluates eta, chi, S and dtau exactly as compute_source_dtau does (solver.tpp:2139-2186) and
// eliminates the tridiagonal system from both ends simultaneously:
//
//   first side  (entries first..centre-1):  the reference's own forward elimination (FF, Su; :2240-2297)
//   last side   (entries last..centre+1):   its mirror image (GG, and the mirrored Su), which the reference
//                                           already uses for L_diag (:2302-2315)
//   centre row:  u_c = (S_c + A_c a_{c-1} + C_c a'_{c+1}) * L_diag,  L_diag = inverse_C/(FF_c + GP_{c+1})
//
// so no per-entry coefficient array is ever stored: the whole solve lives in registers. This is
// algebraically the reference's forward-elimination/back-substitution (solver.tpp:2193-2356, n_off_diag = 0);
// only the rounding of the elimination differs (observed |dJ/J| < 1e-13, tests/test_twosided_error.py).
#pragma once
#include "mgb_kernels.cuh"

namespace mgb {

struct SideState {
    Entry X;          // the far entry of the segment being closed
    double chiX, SX;  // eta/chi evaluation at X
    double dzX;       // length of the segment between X and its nearer neighbour
    double dtau_far;  // optical depth of X's far segment
    double Scurr_far, Snext_far; // line-integrated sources of that segment (reference orientation)
    double F, FI, a;  // Rybicki-Hummer F (FF or GG), 1/(1+F), eliminated source (Su)
    bool haveX, X_is_end, large_far;
};

template <int APPROX>
__device__ __forceinline__ void eval_entry(const DevModel& m, const Entry& e, uint32_t ll, double freq, double& chi, double& S) {
    double eta;
    eta_chi_interp<APPROX>(m, e, ll, freq * e.shift, eta, chi);
    S = eta / chi;
}

// Close the segment between the far entry st.X and the nearer entry Y, then advance.
// side 0: reference orientation curr = X, next = Y.  side 1: curr = Y, next = X.
template <int APPROX>
__device__ __forceinline__ void push_entry(const DevModel& m, SideState& st, int side, const Entry& Y, double chiY, double SY, double dzY,
                                           uint32_t ll, double freq, unsigned& nlarge) {
    if (st.haveX) {
        const Entry& c = side ? Y : st.X;
        const Entry& n = side ? st.X : Y;
        // compute_source_dtau (solver.tpp:2139-2186)
        const double dshift_abs = fabs(n.shift - c.shift);
        const double da = __ldg(&m.dshift_max[c.p1]), db = __ldg(&m.dshift_max[n.p1]);
        const double dshift_max = (db < da) ? db : da;
        double dtau, Scurr = 0.0, Snext = 0.0;
        const bool large = dshift_abs > dshift_max;
        if (large) {
            S_dtau_line_integrated<APPROX>(m, c, n, ll, freq * c.shift, freq * n.shift, st.dzX, dtau, Scurr, Snext);
            nlarge++;
        } else {
            const double chic = side ? chiY : st.chiX, chin = side ? st.chiX : chiY;
            dtau = 0.5 * (chic + chin) * st.dzX;
        }
        if (st.X_is_end) {
            // boundary row (solver.tpp:2240-2262 for `first`, 2283-2297 mirrored for `last`)
            const double S0 = large ? (side ? Snext : Scurr) : st.SX;
            const double inv = 1.0 / dtau;
            const double Cb  = 2.0 * inv * inv;
            const double Bm  = 1.0 + 2.0 * inv;
            const double B   = Bm + Cb;
            const double Ib  = boundary_intensity(m, st.X.p1, freq * st.X.shift);
            st.a  = (S0 + 2.0 * Ib * inv) / B;
            st.F  = 0.5 * Bm * dtau * dtau;
            st.FI = 1.0 / (1.0 + st.F);
        } else {
            // interior row (solver.tpp:2264-2281; mirrored 2306-2311)
            const double Sx = side ? (st.large_far ? st.Scurr_far : st.SX) : (large ? Scurr : st.SX);
            const double avg     = 0.5 * (st.dtau_far + dtau);
            const double invFar  = avg * st.dtau_far;
            const double invNear = avg * dtau;
            const double Afar    = 1.0 / invFar;
            st.F  = (Afar * st.F * st.FI + 1.0) * invNear;
            st.FI = 1.0 / (1.0 + st.F);
            st.a  = (Afar * st.a + Sx) * st.FI * invNear;
        }
        st.dtau_far  = dtau;
        st.large_far = large;
        st.Scurr_far = Scurr;
        st.Snext_far = Snext;
        st.X_is_end  = false;
    } else {
        st.haveX    = true;
        st.X_is_end = true;
    }
    st.X = Y; st.chiX = chiY; st.SX = SY; st.dzX = dzY;
}

template <int APPROX>
__global__ void __launch_bounds__(128) k_solve(DevModel m, PathView pv, ScratchView sc, uint32_t nchunks_f) {
    const unsigned lane = threadIdx.x & 31;
    const size_t item   = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const size_t nitems = (size_t)pv.no * m.hnrays * nchunks_f;
    if (item >= nitems) return;
    const uint32_t cf = (uint32_t)(item % nchunks_f);
    const size_t t    = item / nchunks_f;
    const uint32_t rr = (uint32_t)(t % m.hnrays);
    const uint32_t ol = (uint32_t)(t / m.hnrays);
    const uint32_t o  = pv.o0 + ol;
    const uint32_t f_raw = cf * 32 + lane;
    const bool active    = f_raw < m.nfreqs;
    const uint32_t f     = active ? f_raw : m.nfreqs - 1;
    const uint32_t l     = f / m.nquads;       // line of this frequency bin (f = l*nquads + z, model.cpp:127-146)
    const uint32_t z     = f - l * m.nquads;
    const uint32_t ll    = (APPROX == APPROX_ONELINE) ? __ldg(&m.corr_line[f]) : l;
    const uint32_t ar    = __ldg(&m.antipod[rr]);
    const double freq    = __ldg(&m.nu[(size_t)o * m.nfreqs + f]);
    const size_t out     = ((size_t)rr * pv.no + ol) * m.nfreqs + f;

    const uint32_t len_b = __ldg(&pv.len[(size_t)o * m.nrays + rr]);
    const uint32_t len_f = __ldg(&pv.len[(size_t)o * m.nrays + ar]);
    if (len_b + len_f == 0) {
        // isolated origin (solver.tpp:510-518)
        if (active) { sc.u[out] = boundary_intensity(m, o, freq); sc.L[out] = 0.0; }
        const unsigned e1 = __popc(__ballot_sync(0xffffffffu, active && z == 0)); /* n_tot = 1 per line */
        if (lane == 0 && e1) atomicAdd(&sc.counters[0], (unsigned long long)e1);
        return;
    }

    Entry C;
    C.p1 = o; C.p2 = o; C.factor = 0.0; C.shift = 1.0;
    double chiC, SC;
    eval_entry<APPROX>(m, C, ll, freq, chiC, SC);

    unsigned nlarge = 0, nentries = 1;
    // results of the two sides
    double F0 = 0, FI0 = 0, a0 = 0, dt0 = 0, Sn0 = 0, F1 = 0, FI1 = 0, a1 = 0, dt1 = 0, Sc1 = 0;
    bool lg0 = false, lg1 = false;

    for (int side = 0; side < 2; side++) {
        const uint32_t r   = side ? ar : rr;
        const uint32_t len = side ? len_f : len_b;
        if (len == 0) continue;
        const size_t hr = (size_t)ol * m.nrays + r;
        const unsigned long long off = __ldg(&pv.poff[hr]);
        const double shift0 = __ldg(&pv.shift0[hr]);
        SideState st;
        st.haveX = false; st.X_is_end = false; st.large_far = false;
        st.F = st.FI = st.a = st.dtau_far = st.Scurr_far = st.Snext_far = 0.0;

        uint32_t nxt    = __ldg(&pv.nr[off + len - 1]);
        double sh_nxt   = __ldg(&pv.shift[off + len - 1]);
        double crit_nxt = __ldg(&m.lrow[(size_t)nxt * m.nlines + l].criterion);
        for (uint32_t j = len; j-- > 0;) {
            const double dZj   = __ldg(&pv.dZ[off + j]);
            const uint32_t crt = j ? __ldg(&pv.nr[off + j - 1]) : o;
            const double sh_crt = j ? __ldg(&pv.shift[off + j - 1]) : shift0;
            const double crit_crt = __ldg(&m.lrow[(size_t)crt * m.nlines + l].criterion);
            const uint32_t ni = n_interp_from_criteria(m, crit_crt, crit_nxt);
            nentries += ni;
            double dz = dZj, dshi = 0.0;
            if (ni > 1) { dz = dZj / ni; dshi = (sh_nxt - sh_crt) / ni; }
            for (uint32_t i = ni; i >= 1; i--) {
                Entry Y;
                Y.p1 = nxt; Y.p2 = crt;
                Y.factor = (ni > 1) ? 1.0 - ((double)i) / ni : 0.0;
                Y.shift  = (ni > 1) ? sh_crt + i * dshi : sh_nxt;
                double chiY, SY;
                eval_entry<APPROX>(m, Y, ll, freq, chiY, SY);
                push_entry<APPROX>(m, st, side, Y, chiY, SY, dz, ll, freq, nlarge);
            }
            nxt = crt; sh_nxt = sh_crt; crit_nxt = crit_crt;
        }
        push_entry<APPROX>(m, st, side, C, chiC, SC, 0.0, ll, freq, nlarge);
        if (side == 0) { F0 = st.F; FI0 = st.FI; a0 = st.a; dt0 = st.dtau_far; lg0 = st.large_far; Sn0 = st.Snext_far; }
        else { F1 = st.F; FI1 = st.FI; a1 = st.a; dt1 = st.dtau_far; lg1 = st.large_far; Sc1 = st.Scurr_far; }
    }

    // ---- centre row ----
    double u, Ldiag;
    if (len_b && len_f) {
        const double S_c     = lg1 ? Sc1 : SC;
        const double avg     = 0.5 * (dt0 + dt1);
        const double invA    = avg * dt0, invC = avg * dt1;
        const double A = 1.0 / invA, Cc = 1.0 / invC;
        const double FFc = (A * F0 * FI0 + 1.0) * invC;
        const double GP  = F1 * FI1;
        Ldiag = invC / (FFc + GP);
        u     = (S_c + A * a0 + Cc * a1) * Ldiag;
    } else if (len_f) {
        // centre == first (solver.tpp:2240-2262 with the back-substitution folded in)
        const double S_c = lg1 ? Sc1 : SC;
        const double inv = 1.0 / dt1;
        const double Cf  = 2.0 * inv * inv;
        const double Bm  = 1.0 + 2.0 * inv;
        const double Ib  = boundary_intensity(m, o, freq);
        const double FFf = 0.5 * Bm * dt1 * dt1;
        const double GP  = F1 * FI1;
        Ldiag = (1.0 / Cf) / (FFf + GP);
        u     = (S_c + 2.0 * Ib * inv + Cf * a1) * Ldiag;
    } else {
        // centre == last (solver.tpp:2283-2297, 2317)
        const double S_c = lg0 ? Sn0 : SC;
        const double inv = 1.0 / dt0;
        const double Al  = 2.0 * inv * inv;
        const double Bm  = 1.0 + 2.0 * inv;
        const double Bl  = Bm + Al;
        const double Ib  = boundary_intensity(m, o, freq);
        Ldiag = (1.0 + F0) / (Bm + Bl * F0);
        u     = (Al * a0 + (S_c + 2.0 * Ib * inv)) * Ldiag;
    }

    // ---- Lambda increment (update_Lambda, solver.tpp:1760-1806; diagonal element only, n_off_diag = 0) ----
    const double w_ang     = 2.0 * __ldg(&m.weight[rr]);
    const double freq_line = __ldg(&m.line[l]);
    const double constante = __ldg(&m.line_A[l]) * __ldg(&m.line_quadw[(size_t)l * m.nquads + z]) * w_ang;
    double eta_l, chi_l;
    // NOTE: the reference passes the within-species transition index k as OneLine's line index here (solver.tpp:1798)
    eta_chi_point<APPROX>(m, o, __ldg(&m.line_k[l]), freq_line, eta_l, chi_l);
    const double inverse_chi = 1.0 / chi_l;
    const double frq = freq * 1.0; /* nu(o,f) * shift[centre] */
    const double phi = thermo_profile(m, __ldg(&m.line_invmass[l]), o, freq_line, frq);
    const double Lam = constante * frq * phi * Ldiag * inverse_chi;

    if (active) { sc.u[out] = u; sc.L[out] = Lam; }

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
        acc += v * 2.0 * __ldg(&m.weight[rr]);
        if (u_full) u_full[((size_t)rr * u_no + (o0 + ol - u_o0)) * m.nfreqs + f] = v;
    }
    J[(size_t)(o0 + ol) * m.nfreqs + f] = acc;
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

} // namespace mgb
