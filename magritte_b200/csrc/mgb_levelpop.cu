// mgb_levelpop.cu -- the level-population half of an ALI iteration on the device.
//
// What the reference does on the host between two radiation-field calls (Model::compute_level_populations, src/model/model.cpp:654-747):
//   Model::compute_Jeff                                  model.cpp:514-545 (and compute_Jeff_sparse :549-570)
//   LineProducingSpecies::update_using_statistical_equilibrium      lineProducingSpecies.tpp:377-581 (n_off_diag = 0: one nlev x nlev
//                                                        block per point) / _sparse :589-705 (the same blocks, solved one by one)
//   LineProducingSpecies::correct_negative_populations   lineProducingSpecies.tpp:711-802
//   LineProducingSpecies::check_for_convergence[_trial]  lineProducingSpecies.tpp:128-165, 171-201
//   LineProducingSpecies::update_using_acceleration[_trial]  (Ng)  lineProducingSpecies.tpp:266-371
//   Lines::set_emissivity_and_opacity, set_inverse_width lines.tpp:29-59, InterpHelper criterion solver_interp_helper.hpp:15-30
// Here all of it reads J / Lambda where k_solve's reductions left them and writes the LineRow table k_solve reads next, so an
// iteration moves nothing across PCIe.
//
// Number types. The reference keeps populations in 80-bit long double (Eigen VectorXld, lineProducingSpecies.hpp:46) and solves the
// rate equations with a long-double SparseLU. The GPU has no 80-bit type: populations are double-double pairs {hi, lo} (106-bit
// significand, a superset of the 64-bit one) and the per-point system is eliminated in double-double with partial pivoting. Matrix
// ENTRIES are computed in plain double exactly as the reference does (A + Bs*Jeff, -Ls*opacity/freq, Cd*abn are Real expressions there).
// Agreement with the reference is therefore bounded by the reference's own rounding (long double, conditioning of the rate
// matrix), not by this solver: tests/test_gpu_levelpop_device.py.
#include <algorithm>

#include "mgb_ctx.h"

using namespace mgb;

#define MGB_TRY(x)            \
    do {                      \
        int rc_ = (x);        \
        if (rc_ != MGB_OK) return rc_; \
    } while (0)

namespace {

constexpr double kFOUR_PI         = 4.0 * 3.14159265358979323846;
constexpr double kHH_OVER_FOUR_PI = kHH / kFOUR_PI;

// ---- double-double arithmetic (error-free transformations; the file is compiled with -fmad=false, so only the explicit fma fuses) ----
struct dd { double hi, lo; };
__host__ __device__ __forceinline__ dd dd_make(double a) { return dd{a, 0.0}; }
__host__ __device__ __forceinline__ dd two_sum(double a, double b) {
    const double s = a + b, bb = s - a;
    return dd{s, (a - (s - bb)) + (b - bb)};
}
__host__ __device__ __forceinline__ dd quick_two_sum(double a, double b) {
    const double s = a + b;
    return dd{s, b - (s - a)};
}
__device__ __forceinline__ dd two_prod(double a, double b) {
    const double p = a * b;
    return dd{p, __fma_rn(a, b, -p)};
}
__device__ __forceinline__ dd dd_add(dd a, dd b) {
    dd s = two_sum(a.hi, b.hi);
    const dd t = two_sum(a.lo, b.lo);
    s.lo += t.hi;
    s = quick_two_sum(s.hi, s.lo);
    s.lo += t.lo;
    return quick_two_sum(s.hi, s.lo);
}
__device__ __forceinline__ dd dd_neg(dd a) { return dd{-a.hi, -a.lo}; }
__device__ __forceinline__ dd dd_sub(dd a, dd b) { return dd_add(a, dd_neg(b)); }
__device__ __forceinline__ dd dd_add_d(dd a, double b) {
    dd s = two_sum(a.hi, b);
    s.lo += a.lo;
    return quick_two_sum(s.hi, s.lo);
}
__device__ __forceinline__ dd dd_mul(dd a, dd b) {
    dd p = two_prod(a.hi, b.hi);
    p.lo += a.hi * b.lo + a.lo * b.hi;
    return quick_two_sum(p.hi, p.lo);
}
__device__ __forceinline__ dd dd_mul_d(dd a, double b) {
    dd p = two_prod(a.hi, b);
    p.lo += a.lo * b;
    return quick_two_sum(p.hi, p.lo);
}
__device__ __forceinline__ dd dd_div(dd a, dd b) {
    const double q1 = a.hi / b.hi;
    dd r = dd_sub(a, dd_mul_d(b, q1));
    const double q2 = r.hi / b.hi;
    r = dd_sub(r, dd_mul_d(b, q2));
    const double q3 = r.hi / b.hi;
    dd q = quick_two_sum(q1, q2);
    return dd_add_d(q, q3);
}
__device__ __forceinline__ double dd_to_double(dd a) { return a.hi + a.lo; }
__device__ __forceinline__ dd ld_dd(const double2* p) { const double2 v = *p; return dd{v.x, v.y}; }
__device__ __forceinline__ void st_dd(double2* p, dd v) { *p = make_double2(v.hi, v.lo); }

struct ColParView {
    uint32_t num_col_partner, orth_or_para, ntmp, ncol;
    const uint32_t *icol, *jcol;
    const double *tmp, *Ce, *Cd;
};
struct SpeciesView {
    uint32_t npoints, nlev, nrad, num, line0, nlines, nquads, nfreqs, nspecs, ncolpar;
    const uint32_t *irad, *jrad;
    const double *energy, *weight, *frequency, *A, *Ba, *Bs;
    const ColParView* colpar;
    double2 *pop, *prev1, *trial, *trial_prev;
    double *Jlin, *Jeff, *Jdif;
    const double* pop_tot;
    const double* abundance;   // [npoints][nspecs]
    const double* temperature; // [npoints]
    const double* vturb2;
    double min_line_opacity, min_line_emissivity, population_inversion_fraction, min_rel_pop, inverse_mass;
};

// LineProducingSpecies::get_emissivity / get_opacity (lineProducingSpecies.tpp:23-44): Real * long double products, max with the
// long double floor, returned as Real.
__device__ __forceinline__ double lp_emissivity(const SpeciesView& s, uint32_t p, uint32_t k) {
    const dd pi = ld_dd(&s.pop[(size_t)p * s.nlev + s.irad[k]]);
    const double c = s.frequency[k] * kHH_OVER_FOUR_PI * s.A[k];
    const double v = dd_to_double(dd_mul_d(pi, c));
    return v > s.min_line_emissivity ? v : s.min_line_emissivity; /* std::max(a, b): b unless a > b ... a if equal; same value */
}
__device__ __forceinline__ double lp_opacity(const SpeciesView& s, uint32_t p, uint32_t k) {
    const dd pi = ld_dd(&s.pop[(size_t)p * s.nlev + s.irad[k]]), pj = ld_dd(&s.pop[(size_t)p * s.nlev + s.jrad[k]]);
    const double c = s.frequency[k] * kHH_OVER_FOUR_PI;
    const dd d = dd_sub(dd_mul_d(pj, s.Ba[k]), dd_mul_d(pi, s.Bs[k]));
    const double v = dd_to_double(dd_mul_d(d, c));
    return v > s.min_line_opacity ? v : s.min_line_opacity;
}

// Lines::set_emissivity_and_opacity + set_inverse_width + the InterpHelper criterion, straight into the LineRow table of k_solve.
__global__ void k_lp_line_state(SpeciesView s, LineRow* __restrict__ rows) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)s.npoints * s.nrad) return;
    const uint32_t p = (uint32_t)(i / s.nrad), k = (uint32_t)(i % s.nrad);
    LineRow r;
    const double width = sqrt(kTWO_KB_OVER_AMU_CC2 * s.inverse_mass * s.temperature[p] + s.vturb2[p]); /* thermodynamics.tpp:27-34 */
    r.inverse_width = 1.0 / (s.frequency[k] * width);                                                 /* lines.tpp:45-59 */
    r.emissivity    = lp_emissivity(s, p, k);
    r.opacity       = lp_opacity(s, p, k);
    r.criterion     = r.opacity / s.pop_tot[p];
    rows[(size_t)p * s.nlines + s.line0 + k] = r;
}

// Model::compute_Jeff (model.cpp:514-545) / compute_Jeff_sparse (:549-570). J is radiation.J [npoints][nfreqs] (sparse: lspec.J per
// line, [npoints][nlines]); lam_val / lam_cnt the diagonal Lambda element (nr == p) and the list length.
__global__ void k_lp_jeff(SpeciesView s, int sparse, const double* __restrict__ J, const double* __restrict__ J_lines,
                          const double* __restrict__ quadw, const double* __restrict__ lam_val, const uint32_t* __restrict__ lam_cnt) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)s.npoints * s.nrad) return;
    const uint32_t p = (uint32_t)(i / s.nrad), k = (uint32_t)(i % s.nrad);
    const uint32_t l = s.line0 + k;
    double jl;
    if (sparse) jl = J_lines[(size_t)p * s.nlines + l];
    else {
        jl = 0.0;
        for (uint32_t z = 0; z < s.nquads; z++) jl += quadw[(size_t)l * s.nquads + z] * J[(size_t)p * s.nfreqs + (size_t)l * s.nquads + z];
    }
    double diff = 0.0;
    const uint32_t cnt = lam_cnt[(size_t)p * s.nlines + l];
    for (uint32_t m = 0; m < cnt; m++) { /* n_off_diag = 0: at most the element nr == p */
        const dd pi = ld_dd(&s.pop[(size_t)p * s.nlev + s.irad[k]]);
        diff = dd_to_double(dd_add_d(dd_mul_d(pi, lam_val[(size_t)p * s.nlines + l]), diff)); /* diff += Ls * population (Real += long double) */
    }
    s.Jlin[i] = jl;
    s.Jeff[i] = jl - kHH_OVER_FOUR_PI * diff;
    s.Jdif[i] = kHH_OVER_FOUR_PI * diff;
}

// tools/interpolation.hpp search() as used by CollisionPartner::interpolate_collision_coefficients (collisionPartner.tpp:6-30):
// index of the first tabulated temperature >= value (clamped to the table).
__device__ __forceinline__ uint32_t lp_search(const double* x, uint32_t n, double value) {
    uint32_t start = 0, stop = n - 1;
    if (value >= x[stop]) return stop;
    if (value <= x[start]) return start;
    while (stop > start + 1) {
        const uint32_t middle = (stop + start) / 2;
        if (value > x[middle]) start = middle;
        else if (value < x[middle]) stop = middle;
        else return middle;
    }
    return stop;
}

// One block per point: the point's nlev x nlev rate matrix (transposed as in the reference) with the conservation row, augmented
// with the right-hand side, in shared memory as double-double; Gaussian elimination with partial pivoting; back-substitution.
// Thread c owns COLUMN c while the matrix is assembled (contributions are added in the reference's triplet order, so every entry
// is the same sum) and during the elimination.
__global__ void __launch_bounds__(64) k_lp_stateq(SpeciesView s, const double* __restrict__ lam_val, const uint32_t* __restrict__ lam_cnt,
                                                  const LineRow* __restrict__ rows, int* __restrict__ failed) {
    extern __shared__ double2 lp_sm[];
    const uint32_t p = blockIdx.x, n = s.nlev, W = n + 1, t = threadIdx.x;
    double2* M = lp_sm;                     // [n][W]
    __shared__ uint32_t s_piv;
    __shared__ double s_T;
    stage_exp_table();
    for (uint32_t i = t; i < n * W; i += blockDim.x) M[i] = make_double2(0.0, 0.0);
    if (t == 0) s_T = s.temperature[p];
    __syncthreads();
    const double T = s_T;
    for (uint32_t c = t; c < n; c += blockDim.x) {
        auto add = [&](uint32_t r, double v) { st_dd(&M[r * W + c], dd_add_d(ld_dd(&M[r * W + c]), v)); };
        // radiative transitions
        for (uint32_t k = 0; k < s.nrad; k++) {
            const uint32_t I = s.irad[k], J = s.jrad[k];
            if (I != c && J != c) continue;
            const double Je   = s.Jeff[(size_t)p * s.nrad + k];
            const double v_IJ = s.A[k] + s.Bs[k] * Je;
            const double v_JI = s.Ba[k] * Je;
            if (J != n - 1) { if (I == c) add(J, +v_IJ); if (J == c) add(J, -v_JI); }
            if (I != n - 1) { if (J == c) add(I, +v_JI); if (I == c) add(I, -v_IJ); }
        }
        // approximated Lambda operator (diagonal element, nr == p)
        for (uint32_t k = 0; k < s.nrad; k++) {
            const uint32_t I = s.irad[k], J = s.jrad[k];
            if (I != c) continue;
            const uint32_t l = s.line0 + k;
            const uint32_t cnt = lam_cnt[(size_t)p * s.nlines + l];
            for (uint32_t m = 0; m < cnt; m++) {
                const double v_IJ = -lam_val[(size_t)p * s.nlines + l] * rows[(size_t)p * s.nlines + l].opacity / s.frequency[k];
                if (J != n - 1) add(J, +v_IJ);
                if (I != n - 1) add(I, -v_IJ);
            }
        }
        // collisional transitions
        for (uint32_t q = 0; q < s.ncolpar; q++) {
            const ColParView cp = s.colpar[q];
            double abn = s.abundance[(size_t)p * s.nspecs + cp.num_col_partner];
            if (cp.orth_or_para) { /* CollisionPartner::adjust_abundance_for_ortho_or_para (collisionPartner.tpp:33-40) */
                const double frac_H2_para = 1.0 / (1.0 + 9.0 * (double)mgb_expf((float)(-170.5 / T)));
                abn *= (cp.orth_or_para == 1) ? (1.0 - frac_H2_para) : frac_H2_para;
            }
            const uint32_t ti = lp_search(cp.tmp, cp.ntmp, T);
            const bool lo_end = (ti == 0), hi_end = (ti == cp.ntmp - 1);
            const double step = (lo_end || hi_end) ? 0.0 : (T - cp.tmp[ti - 1]) / (cp.tmp[ti] - cp.tmp[ti - 1]);
            for (uint32_t k = 0; k < cp.ncol; k++) {
                const uint32_t I = cp.icol[k], J = cp.jcol[k];
                if (I != c && J != c) continue;
                double Cd, Ce;
                if (lo_end || hi_end) { Cd = cp.Cd[(size_t)ti * cp.ncol + k]; Ce = cp.Ce[(size_t)ti * cp.ncol + k]; }
                else {
                    const double Ce0 = cp.Ce[(size_t)(ti - 1) * cp.ncol + k], Cd0 = cp.Cd[(size_t)(ti - 1) * cp.ncol + k];
                    Ce = Ce0 + (cp.Ce[(size_t)ti * cp.ncol + k] - Ce0) * step;
                    Cd = Cd0 + (cp.Cd[(size_t)ti * cp.ncol + k] - Cd0) * step;
                }
                const double v_IJ = Cd * abn, v_JI = Ce * abn;
                if (J != n - 1) { if (I == c) add(J, +v_IJ); if (J == c) add(J, -v_JI); }
                if (I != n - 1) { if (J == c) add(I, +v_JI); if (I == c) add(I, -v_IJ); }
            }
        }
        add(n - 1, 1.0); /* conservation row */
    }
    if (t == 0) st_dd(&M[(n - 1) * W + n], dd_make(s.pop_tot[p])); /* y[nlev-1] = population_tot */
    __syncthreads();
    // ---- elimination with partial pivoting (first largest |pivot| as std::max_element would pick) ----
    for (uint32_t c = 0; c < n; c++) {
        if (t == 0) {
            uint32_t piv = c;
            double big = fabs(M[c * W + c].x);
            for (uint32_t r = c + 1; r < n; r++) { const double v = fabs(M[r * W + c].x); if (v > big) { big = v; piv = r; } }
            s_piv = (big == 0.0 || big != big) ? 0xffffffffu : piv;
        }
        __syncthreads();
        const uint32_t piv = s_piv;
        if (piv == 0xffffffffu) { if (t == 0) atomicExch(failed, 1); return; }
        if (piv != c) for (uint32_t j = t; j < W; j += blockDim.x) { const double2 a = M[c * W + j]; M[c * W + j] = M[piv * W + j]; M[piv * W + j] = a; }
        __syncthreads();
        const dd pv = ld_dd(&M[c * W + c]);
        for (uint32_t r = c + 1 + t; r < n; r += blockDim.x) st_dd(&M[r * W + c], dd_div(ld_dd(&M[r * W + c]), pv)); /* multipliers */
        __syncthreads();
        for (uint32_t j = c + 1 + t; j < W; j += blockDim.x) {
            const dd top = ld_dd(&M[c * W + j]);
            if (top.hi == 0.0 && top.lo == 0.0) continue;
            for (uint32_t r = c + 1; r < n; r++) {
                const dd f = ld_dd(&M[r * W + c]);
                if (f.hi == 0.0) continue;
                st_dd(&M[r * W + j], dd_sub(ld_dd(&M[r * W + j]), dd_mul(f, top)));
            }
        }
        __syncthreads();
    }
    // ---- back-substitution: lanes of warp 0 share each row's dot product ----
    if (t < 32) {
        for (uint32_t r = n; r-- > 0;) {
            dd acc = dd_make(0.0);
            for (uint32_t k = r + 1 + t; k < n; k += 32) acc = dd_add(acc, dd_mul(ld_dd(&M[r * W + k]), ld_dd(&M[k * W + n])));
#pragma unroll
            for (int off = 16; off; off >>= 1) {
                dd o;
                o.hi = __shfl_xor_sync(0xffffffffu, acc.hi, off);
                o.lo = __shfl_xor_sync(0xffffffffu, acc.lo, off);
                acc  = dd_add(acc, o);
            }
            if (t == 0) st_dd(&M[r * W + n], dd_div(dd_sub(ld_dd(&M[r * W + n]), acc), ld_dd(&M[r * W + r])));
            __syncwarp();
        }
    }
    __syncthreads();
    for (uint32_t i = t; i < n; i += blockDim.x) s.pop[(size_t)p * n + i] = M[i * W + n];
}

// LineProducingSpecies::correct_negative_populations (lineProducingSpecies.tpp:711-802), one thread per point.
__device__ __forceinline__ void lp_set_LTE(const SpeciesView& s, uint32_t p, uint32_t k) { /* lineProducingSpecies.tpp:75-92 */
    double2* P = s.pop + (size_t)p * s.nlev;
    const uint32_t i = s.irad[k], j = s.jrad[k];
    const double tot_line = dd_to_double(dd_add(ld_dd(&P[i]), ld_dd(&P[j])));
    const double T = s.temperature[p];
    const dd ni = dd_make(s.weight[i] * exp(-s.energy[i] / (kKB * T)));
    const dd nj = dd_make(s.weight[j] * exp(-s.energy[j] / (kKB * T)));
    const double part = dd_to_double(dd_add(ni, nj));
    st_dd(&P[i], dd_mul_d(ni, tot_line / part));
    st_dd(&P[j], dd_mul_d(nj, tot_line / part));
}
__global__ void k_lp_negative(SpeciesView s, unsigned char* __restrict__ level_mask /* [npoints][nlev] scratch */) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= s.npoints) return;
    double2* P = s.pop + (size_t)p * s.nlev;
    const double ptot = s.pop_tot[p];
    if (ptot > 0.0) {
        double total_positive = 0.0;
        for (uint32_t i = 0; i < s.nlev; i++) {
            const dd v = ld_dd(&P[i]);
            if (v.hi < 0.0 || (v.hi == 0.0 && v.lo < 0.0)) st_dd(&P[i], dd_make(0.0));
            else total_positive = dd_to_double(dd_add_d(v, total_positive)); /* Real += long double */
        }
        for (uint32_t i = 0; i < s.nlev; i++) st_dd(&P[i], dd_div(dd_mul_d(ld_dd(&P[i]), ptot), dd_make(total_positive)));
    }
    // masers: transitions whose inversion exceeds the allowed fraction are put in LTE, until the set of touched levels is stable
    unsigned char* mask = level_mask + (size_t)p * s.nlev;
    for (uint32_t i = 0; i < s.nlev; i++) mask[i] = 0;
    auto inverted = [&](uint32_t k) {
        const dd pi = ld_dd(&P[s.irad[k]]), pj = ld_dd(&P[s.jrad[k]]);
        const dd rhs = dd_mul_d(dd_mul_d(pi, s.Bs[k] / s.Ba[k]), s.population_inversion_fraction);
        const dd d = dd_sub(pj, rhs);
        return d.hi < 0.0 || (d.hi == 0.0 && d.lo < 0.0);
    };
    uint32_t next_n = 0, curr_n;
    do {
        curr_n = next_n;
        bool any = false;
        for (uint32_t k = s.nrad; k-- > 0;)
            if (inverted(k)) { lp_set_LTE(s, p, k); mask[s.irad[k]] = 1; mask[s.jrad[k]] = 1; any = true; }
        for (uint32_t k = 0; k < s.nrad; k++)
            if (inverted(k)) { lp_set_LTE(s, p, k); mask[s.irad[k]] = 1; mask[s.jrad[k]] = 1; any = true; }
        if (any) { /* set_LTE_specific_levels (lineProducingSpecies.tpp:98-121) */
            double tot_levels = 0.0, part = 0.0;
            const double T = s.temperature[p];
            for (uint32_t i = 0; i < s.nlev; i++)
                if (mask[i]) {
                    tot_levels = dd_to_double(dd_add_d(ld_dd(&P[i]), tot_levels));
                    const double v = s.weight[i] * exp(-s.energy[i] / (kKB * T));
                    st_dd(&P[i], dd_make(v));
                    part += v;
                }
            for (uint32_t i = 0; i < s.nlev; i++)
                if (mask[i]) st_dd(&P[i], dd_mul_d(ld_dd(&P[i]), tot_levels / part));
        }
        next_n = 0;
        for (uint32_t i = 0; i < s.nlev; i++) next_n += mask[i];
    } while (next_n != curr_n);
}

// LineProducingSpecies::check_for_convergence (lineProducingSpecies.tpp:128-165) and _trial (:171-201): per-block partial
// {count of non-converged levels, sum of relative changes, max relative change}; the host adds the partials in block order.
__global__ void __launch_bounds__(256) k_lp_convergence(SpeciesView s, int trial, double pop_prec, double* __restrict__ partials) {
    __shared__ double sh_sum[256], sh_max[256];
    __shared__ unsigned sh_cnt[256];
    const size_t n = (size_t)s.npoints * s.nlev;
    double sum = 0.0, mx = 0.0;
    unsigned cnt = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const dd a = ld_dd(trial ? &s.trial[i] : &s.pop[i]), b = ld_dd(trial ? &s.trial_prev[i] : &s.prev1[i]);
        bool counted = true;
        if (!trial) {
            const uint32_t p = (uint32_t)(i / s.nlev);
            const dd thr = dd_sub(a, dd_make(s.min_rel_pop * s.pop_tot[p]));
            counted = thr.hi > 0.0 || (thr.hi == 0.0 && thr.lo > 0.0);
        }
        if (counted) {
            const double rc = 2.0 * fabs(dd_to_double(dd_div(dd_sub(a, b), dd_add(a, b))));
            if (rc > pop_prec) cnt++;
            sum += rc; /* NaN (0/0 of the reference's uninitialised trial arrays) propagates like there */
            mx = (rc > mx) ? rc : mx; /* std::max(relative_change, max): keeps max when rc is NaN */
        }
    }
    sh_sum[threadIdx.x] = sum; sh_max[threadIdx.x] = mx; sh_cnt[threadIdx.x] = cnt;
    __syncthreads();
    for (int off = 128; off; off >>= 1) {
        if ((int)threadIdx.x < off) {
            sh_sum[threadIdx.x] += sh_sum[threadIdx.x + off];
            sh_cnt[threadIdx.x] += sh_cnt[threadIdx.x + off];
            const double o = sh_max[threadIdx.x + off];
            if (o > sh_max[threadIdx.x]) sh_max[threadIdx.x] = o;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) { partials[3 * blockIdx.x] = (double)sh_cnt[0]; partials[3 * blockIdx.x + 1] = sh_sum[0]; partials[3 * blockIdx.x + 2] = sh_max[0]; }
}

__global__ void k_lp_diff(size_t n, const double2* __restrict__ a, const double2* __restrict__ b, double2* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) st_dd(&out[i], dd_sub(ld_dd(&a[i]), ld_dd(&b[i])));
}
// dot product partials in double-double (the reference's residuals[i].dot(residuals[j]) in long double, lineProducingSpecies.tpp:293-297)
__global__ void __launch_bounds__(256) k_lp_dot(size_t n, const double2* __restrict__ a, const double2* __restrict__ b, double2* __restrict__ partials) {
    __shared__ double2 sh[256];
    dd acc = dd_make(0.0);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        acc = dd_add(acc, dd_mul(ld_dd(&a[i]), ld_dd(&b[i])));
    sh[threadIdx.x] = make_double2(acc.hi, acc.lo);
    __syncthreads();
    for (int off = 128; off; off >>= 1) {
        if ((int)threadIdx.x < off) st_dd(&sh[threadIdx.x], dd_add(ld_dd(&sh[threadIdx.x]), ld_dd(&sh[threadIdx.x + off])));
        __syncthreads();
    }
    if (threadIdx.x == 0) partials[blockIdx.x] = sh[0];
}
// population = sum_i populations[popsize-1-i] * coef[order-1-i] (lineProducingSpecies.tpp:305-309)
struct CombineArgs { const double2* P[16]; double2 c[16]; uint32_t order; };
__global__ void k_lp_combine(size_t n, CombineArgs a, double2* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    dd acc = dd_make(0.0);
    for (uint32_t k = 0; k < a.order; k++) acc = dd_add(acc, dd_mul(ld_dd(&a.P[k][i]), dd{a.c[k].x, a.c[k].y}));
    st_dd(&out[i], acc);
}

constexpr int kLpBlocks = 592; // 4 x 148 partial blocks for the reductions

int upload(mgb_ctx* ctx, DevBuf& b, const void* host, size_t bytes) {
    MGB_CUDA(ctx, b.ensure(bytes));
    if (bytes) MGB_CUDA(ctx, cudaMemcpyAsync(b.p, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return MGB_OK;
}

int species_of(mgb_ctx* ctx, uint32_t l, LpSpecies** out, const char* who) {
    if (l >= ctx->lp.size() || !ctx->lp[l] || !ctx->lp[l]->nlev) return ctx->fail(MGB_ESTATE, "%s: species %u was not set (mgb_lp_set_species)", who, l);
    *out = ctx->lp[l];
    return MGB_OK;
}

int view_of(mgb_ctx* ctx, LpSpecies& sp, SpeciesView& v) {
    const DevModel& m = ctx->m;
    if (!ctx->have_geometry || !ctx->have_thermo || !ctx->have_lines) return ctx->fail(MGB_ESTATE, "level populations: geometry / thermodynamics / lines not set");
    if (!sp.have_pop) return ctx->fail(MGB_ESTATE, "level populations: mgb_lp_set_populations was not called for every species");
    v.npoints = m.npoints; v.nlev = sp.nlev; v.nrad = sp.nrad; v.num = sp.num; v.line0 = sp.line0; v.nlines = m.nlines; v.nquads = m.nquads;
    v.nfreqs = m.nfreqs; v.nspecs = ctx->lp_nspecs; v.ncolpar = (uint32_t)sp.colpar.size();
    v.irad = sp.irad.as<uint32_t>(); v.jrad = sp.jrad.as<uint32_t>();
    v.energy = sp.energy.as<double>(); v.weight = sp.weight.as<double>(); v.frequency = sp.frequency.as<double>();
    v.A = sp.A.as<double>(); v.Ba = sp.Ba.as<double>(); v.Bs = sp.Bs.as<double>();
    v.colpar = sp.colpar_views.as<ColParView>();
    v.pop = sp.pop.as<double2>(); v.prev1 = sp.prev1.as<double2>(); v.trial = sp.trial.as<double2>(); v.trial_prev = sp.trial_prev.as<double2>();
    v.Jlin = sp.Jlin.as<double>(); v.Jeff = sp.Jeff.as<double>(); v.Jdif = sp.Jdif.as<double>();
    v.pop_tot = sp.pop_tot.as<double>();
    v.abundance = ctx->lp_abundance.as<double>(); v.temperature = m.temperature; v.vturb2 = m.vturb2;
    v.min_line_opacity = ctx->lp_min_line_opacity; v.min_line_emissivity = ctx->lp_min_line_emissivity;
    v.population_inversion_fraction = ctx->lp_population_inversion_fraction; v.min_rel_pop = ctx->lp_min_rel_pop_for_convergence;
    v.inverse_mass = 0.0; /* make_view: the species' entry of mgb_set_lines */
    return MGB_OK;
}

int all_species_ready(mgb_ctx* ctx, const char* who) {
    if (ctx->lp.size() != ctx->m.nlspecs || ctx->lp.empty()) return ctx->fail(MGB_ESTATE, "%s: %zu of %u species set (mgb_lp_set_species)", who, ctx->lp.size(), ctx->m.nlspecs);
    for (size_t l = 0; l < ctx->lp.size(); l++) {
        if (!ctx->lp[l] || !ctx->lp[l]->have_pop) return ctx->fail(MGB_ESTATE, "%s: species %zu has no populations (mgb_lp_set_populations)", who, l);
        if (ctx->lp[l]->nrad != ctx->nrad[l]) return ctx->fail(MGB_EINVAL, "%s: species %zu has %u transitions but mgb_set_lines declared %u", who, l, ctx->lp[l]->nrad, ctx->nrad[l]);
    }
    return MGB_OK;
}

int make_view(mgb_ctx* ctx, uint32_t l, SpeciesView& v) {
    MGB_TRY(view_of(ctx, *ctx->lp[l], v));
    v.inverse_mass = ctx->inverse_mass[l];
    v.line0 = ctx->nrad_cum[l];
    return MGB_OK;
}

int copy_dd(mgb_ctx* ctx, DevBuf& dst, const DevBuf& src, size_t n) {
    MGB_CUDA(ctx, dst.ensure(16 * n));
    MGB_CUDA(ctx, cudaMemcpyAsync(dst.p, src.p, 16 * n, cudaMemcpyDeviceToDevice, ctx->stream));
    return MGB_OK;
}

// residuals.push_back(population - populations.back()); populations.push_back(population);
int push_history(mgb_ctx* ctx, LpSpecies& sp) {
    const size_t n = (size_t)ctx->m.npoints * sp.nlev;
    if (sp.populations.empty()) return ctx->fail(MGB_ESTATE, "level populations: empty history (mgb_lp_set_populations seeds it)");
    DevBuf* r = new DevBuf();
    sp.residuals.push_back(r);
    MGB_CUDA(ctx, r->ensure(16 * n));
    k_lp_diff<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(n, sp.pop.as<double2>(), sp.populations.back()->as<double2>(), r->as<double2>());
    MGB_CUDA(ctx, cudaGetLastError());
    DevBuf* q = new DevBuf();
    sp.populations.push_back(q);
    return copy_dd(ctx, *q, sp.pop, n);
}

// population_prev3 = prev2; prev2 = prev1; prev1 = population
int shift_previous(mgb_ctx* ctx, LpSpecies& sp) {
    const size_t n = (size_t)ctx->m.npoints * sp.nlev;
    std::swap(sp.prev3.p, sp.prev2.p); std::swap(sp.prev3.bytes, sp.prev2.bytes); /* prev3 <- old prev2, prev2 <- old prev3 (scratch) */
    std::swap(sp.prev2.p, sp.prev1.p); std::swap(sp.prev2.bytes, sp.prev1.bytes); /* prev2 <- old prev1, prev1 <- scratch */
    return copy_dd(ctx, sp.prev1, sp.pop, n);
}

int line_state(mgb_ctx* ctx) {
    const DevModel& m = ctx->m;
    MGB_CUDA(ctx, ctx->lrow.ensure(sizeof(LineRow) * (size_t)m.npoints * m.nlines));
    for (uint32_t l = 0; l < ctx->lp.size(); l++) {
        SpeciesView v;
        MGB_TRY(make_view(ctx, l, v));
        const size_t n = (size_t)m.npoints * v.nrad;
        k_lp_line_state<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(v, ctx->lrow.as<LineRow>());
        MGB_CUDA(ctx, cudaGetLastError());
    }
    ctx->m.lrow     = ctx->lrow.as<LineRow>();
    ctx->have_state = true;
    return MGB_OK;
}

int convergence(mgb_ctx* ctx, int trial, double pop_prec, mgb_lp_convergence* out) {
    MGB_CUDA(ctx, ctx->lp_partials.ensure(16 * (size_t)kLpBlocks * 4));
    for (uint32_t l = 0; l < ctx->lp.size(); l++) {
        SpeciesView v;
        MGB_TRY(make_view(ctx, l, v));
        k_lp_convergence<<<kLpBlocks, 256, 0, ctx->stream>>>(v, trial, pop_prec, ctx->lp_partials.as<double>());
        MGB_CUDA(ctx, cudaGetLastError());
        std::vector<double> h(3 * kLpBlocks);
        MGB_CUDA(ctx, cudaMemcpyAsync(h.data(), ctx->lp_partials.p, 8 * h.size(), cudaMemcpyDeviceToHost, ctx->stream));
        MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        double cnt = 0.0, sum = 0.0, mx = 0.0;
        for (int b = 0; b < kLpBlocks; b++) { cnt += h[3 * b]; sum += h[3 * b + 1]; if (h[3 * b + 2] > mx) mx = h[3 * b + 2]; }
        const double weight = 1.0 / ((double)ctx->m.npoints * v.nlev); /* lineProducingSpecies.tpp:129 */
        if (out) { out[l].fraction_not_converged = cnt * weight; out[l].relative_change_mean = sum * weight; out[l].relative_change_max = mx; }
    }
    return MGB_OK;
}

} // namespace

// ---------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------
extern "C" int mgb_lp_set_parameters(mgb_ctx* ctx, double min_line_opacity, double min_line_emissivity, double population_inversion_fraction,
                                     double min_rel_pop_for_convergence) {
    if (!ctx) return MGB_EINVAL;
    ctx->lp_min_line_opacity = min_line_opacity; ctx->lp_min_line_emissivity = min_line_emissivity;
    ctx->lp_population_inversion_fraction = population_inversion_fraction; ctx->lp_min_rel_pop_for_convergence = min_rel_pop_for_convergence;
    return MGB_OK;
}

extern "C" int mgb_lp_set_species(mgb_ctx* ctx, uint32_t l, uint32_t species_index, uint32_t nlev, uint32_t nrad, const uint32_t* irad,
                                  const uint32_t* jrad, const double* energy, const double* weight, const double* frequency, const double* A,
                                  const double* Ba, const double* Bs) {
    if (!ctx) return MGB_EINVAL;
    if (!ctx->have_lines) return ctx->fail(MGB_ESTATE, "mgb_lp_set_species before mgb_set_lines");
    if (l >= ctx->m.nlspecs) return ctx->fail(MGB_EINVAL, "mgb_lp_set_species: species %u of %u", l, ctx->m.nlspecs);
    if (!nlev || !nrad || !irad || !jrad || !energy || !weight || !frequency || !A || !Ba || !Bs) return ctx->fail(MGB_EINVAL, "mgb_lp_set_species: null or zero argument");
    if (nrad != ctx->nrad[l]) return ctx->fail(MGB_EINVAL, "mgb_lp_set_species: nrad = %u but mgb_set_lines declared %u", nrad, ctx->nrad[l]);
    if ((size_t)nlev * (nlev + 1) * 16 > 200 * 1024) return ctx->fail(MGB_EUNSUPPORTED, "mgb_lp_set_species: nlev = %u: the per-point rate matrix must fit 200 KB of shared memory (nlev <= 112)", nlev);
    for (uint32_t k = 0; k < nrad; k++) if (irad[k] >= nlev || jrad[k] >= nlev) return ctx->fail(MGB_EINVAL, "mgb_lp_set_species: level index out of range");
    cudaSetDevice(ctx->device);
    if (ctx->lp.size() < ctx->m.nlspecs) ctx->lp.resize(ctx->m.nlspecs, nullptr);
    delete ctx->lp[l];
    LpSpecies* sp = ctx->lp[l] = new LpSpecies();
    sp->nlev = nlev; sp->nrad = nrad; sp->num = species_index; sp->line0 = ctx->nrad_cum[l];
    MGB_TRY(upload(ctx, sp->irad, irad, 4 * (size_t)nrad)); MGB_TRY(upload(ctx, sp->jrad, jrad, 4 * (size_t)nrad));
    MGB_TRY(upload(ctx, sp->energy, energy, 8 * (size_t)nlev)); MGB_TRY(upload(ctx, sp->weight, weight, 8 * (size_t)nlev));
    MGB_TRY(upload(ctx, sp->frequency, frequency, 8 * (size_t)nrad)); MGB_TRY(upload(ctx, sp->A, A, 8 * (size_t)nrad));
    MGB_TRY(upload(ctx, sp->Ba, Ba, 8 * (size_t)nrad)); MGB_TRY(upload(ctx, sp->Bs, Bs, 8 * (size_t)nrad));
    MGB_CUDA(ctx, sp->colpar_views.ensure(8));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return MGB_OK;
}

extern "C" int mgb_lp_add_collision_partner(mgb_ctx* ctx, uint32_t l, uint32_t num_col_partner, int orth_or_para, uint32_t ntmp, uint32_t ncol,
                                            const uint32_t* icol, const uint32_t* jcol, const double* tmp, const double* Ce, const double* Cd) {
    if (!ctx) return MGB_EINVAL;
    LpSpecies* sp = nullptr;
    MGB_TRY(species_of(ctx, l, &sp, "mgb_lp_add_collision_partner"));
    if (!ntmp || !ncol || !icol || !jcol || !tmp || !Ce || !Cd || orth_or_para < 0 || orth_or_para > 2) return ctx->fail(MGB_EINVAL, "mgb_lp_add_collision_partner: bad argument");
    for (uint32_t k = 0; k < ncol; k++) if (icol[k] >= sp->nlev || jcol[k] >= sp->nlev) return ctx->fail(MGB_EINVAL, "mgb_lp_add_collision_partner: level index out of range");
    cudaSetDevice(ctx->device);
    LpColPar* cp = new LpColPar();
    sp->colpar.push_back(cp);
    cp->num_col_partner = num_col_partner; cp->orth_or_para = (uint32_t)orth_or_para; cp->ntmp = ntmp; cp->ncol = ncol;
    MGB_TRY(upload(ctx, cp->icol, icol, 4 * (size_t)ncol)); MGB_TRY(upload(ctx, cp->jcol, jcol, 4 * (size_t)ncol));
    MGB_TRY(upload(ctx, cp->tmp, tmp, 8 * (size_t)ntmp));
    MGB_TRY(upload(ctx, cp->Ce, Ce, 8 * (size_t)ntmp * ncol)); MGB_TRY(upload(ctx, cp->Cd, Cd, 8 * (size_t)ntmp * ncol));
    std::vector<ColParView> views;
    for (auto* c : sp->colpar)
        views.push_back(ColParView{c->num_col_partner, c->orth_or_para, c->ntmp, c->ncol, c->icol.as<uint32_t>(), c->jcol.as<uint32_t>(),
                                   c->tmp.as<double>(), c->Ce.as<double>(), c->Cd.as<double>()});
    MGB_TRY(upload(ctx, sp->colpar_views, views.data(), sizeof(ColParView) * views.size()));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return MGB_OK;
}

extern "C" int mgb_lp_set_abundance(mgb_ctx* ctx, uint32_t nspecs, const double* abundance) {
    if (!ctx || !abundance || !nspecs) return MGB_EINVAL;
    if (!ctx->have_geometry) return ctx->fail(MGB_ESTATE, "mgb_lp_set_abundance before mgb_set_geometry");
    cudaSetDevice(ctx->device);
    MGB_TRY(upload(ctx, ctx->lp_abundance, abundance, 8 * (size_t)ctx->m.npoints * nspecs));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->lp_nspecs = nspecs;
    return MGB_OK;
}

extern "C" int mgb_lp_set_populations(mgb_ctx* ctx, uint32_t l, const double* population_hi, const double* population_lo, const double* population_tot) {
    if (!ctx || !population_hi || !population_tot) return MGB_EINVAL;
    LpSpecies* sp = nullptr;
    MGB_TRY(species_of(ctx, l, &sp, "mgb_lp_set_populations"));
    cudaSetDevice(ctx->device);
    const size_t N = ctx->m.npoints, n = N * sp->nlev;
    std::vector<double2> h(n);
    for (size_t i = 0; i < n; i++) h[i] = make_double2(population_hi[i], population_lo ? population_lo[i] : 0.0);
    MGB_TRY(upload(ctx, sp->pop, h.data(), 16 * n));
    MGB_TRY(upload(ctx, sp->pop_tot, population_tot, 8 * N));
    for (DevBuf* b : {&sp->prev1, &sp->prev2, &sp->prev3, &sp->trial, &sp->trial_prev}) {
        MGB_CUDA(ctx, b->ensure(16 * n));
        MGB_CUDA(ctx, cudaMemsetAsync(b->p, 0, 16 * n, ctx->stream));
    }
    for (DevBuf* b : {&sp->Jlin, &sp->Jeff, &sp->Jdif}) {
        MGB_CUDA(ctx, b->ensure(8 * N * sp->nrad));
        MGB_CUDA(ctx, cudaMemsetAsync(b->p, 0, 8 * N * sp->nrad, ctx->stream));
    }
    for (auto* b : sp->populations) delete b;
    for (auto* b : sp->residuals) delete b;
    sp->populations.clear(); sp->residuals.clear();
    DevBuf* first = new DevBuf(); /* update_using_LTE ends with populations.push_back(population) (lineProducingSpecies.tpp:68) */
    sp->populations.push_back(first);
    MGB_TRY(copy_dd(ctx, *first, sp->pop, n));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    sp->have_pop = true;
    return MGB_OK;
}

extern "C" int mgb_lp_get_populations(mgb_ctx* ctx, uint32_t l, double* population_hi, double* population_lo) {
    if (!ctx || !population_hi) return MGB_EINVAL;
    LpSpecies* sp = nullptr;
    MGB_TRY(species_of(ctx, l, &sp, "mgb_lp_get_populations"));
    if (!sp->have_pop) return ctx->fail(MGB_ESTATE, "mgb_lp_get_populations: no populations on the device");
    cudaSetDevice(ctx->device);
    const size_t n = (size_t)ctx->m.npoints * sp->nlev;
    std::vector<double2> h(n);
    MGB_CUDA(ctx, cudaMemcpy(h.data(), sp->pop.p, 16 * n, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; i++) { population_hi[i] = h[i].x; if (population_lo) population_lo[i] = h[i].y; }
    return MGB_OK;
}

extern "C" int mgb_lp_refresh_line_state(mgb_ctx* ctx) {
    if (!ctx) return MGB_EINVAL;
    MGB_TRY(all_species_ready(ctx, "mgb_lp_refresh_line_state"));
    cudaSetDevice(ctx->device);
    MGB_TRY(line_state(ctx));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return MGB_OK;
}

extern "C" int mgb_lp_get_line_state(mgb_ctx* ctx, double* emissivity, double* opacity, double* inverse_width) {
    if (!ctx) return MGB_EINVAL;
    if (!ctx->have_state) return ctx->fail(MGB_ESTATE, "mgb_lp_get_line_state: no line state on the device");
    cudaSetDevice(ctx->device);
    const size_t n = (size_t)ctx->m.npoints * ctx->m.nlines;
    std::vector<LineRow> h(n);
    MGB_CUDA(ctx, cudaMemcpy(h.data(), ctx->lrow.p, sizeof(LineRow) * n, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; i++) {
        if (emissivity) emissivity[i] = h[i].emissivity;
        if (opacity) opacity[i] = h[i].opacity;
        if (inverse_width) inverse_width[i] = h[i].inverse_width;
    }
    return MGB_OK;
}

extern "C" int mgb_lp_compute_Jeff(mgb_ctx* ctx, int sparse) {
    if (!ctx) return MGB_EINVAL;
    MGB_TRY(all_species_ready(ctx, "mgb_lp_compute_Jeff"));
    NvtxRange nvtx("mgb_lp_compute_Jeff");
    if (!ctx->res_valid || ctx->res_o0 != 0 || ctx->res_o1 != ctx->m.npoints)
        return ctx->fail(MGB_ESTATE, "mgb_lp_compute_Jeff: the radiation field is not complete on this device (rows [%u,%u) of %u; run all origins or "
                                     "mgb_allgather_results first)", ctx->res_o0, ctx->res_o1, ctx->m.npoints);
    if (ctx->res_K) return ctx->fail(MGB_EUNSUPPORTED, "mgb_lp_compute_Jeff: n_off_diag > 0 couples the points (one global sparse system); the device path solves the block-diagonal case");
    cudaSetDevice(ctx->device);
    for (uint32_t l = 0; l < ctx->lp.size(); l++) {
        SpeciesView v;
        MGB_TRY(make_view(ctx, l, v));
        const size_t n = (size_t)ctx->m.npoints * v.nrad;
        k_lp_jeff<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(v, sparse, ctx->J.as<double>(), ctx->J_lines.as<double>(), ctx->m.line_quadw,
                                                                        ctx->lam_val.as<double>(), ctx->lam_cnt.as<uint32_t>());
        MGB_CUDA(ctx, cudaGetLastError());
    }
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return MGB_OK;
}

extern "C" int mgb_lp_get_Jeff(mgb_ctx* ctx, uint32_t l, double* Jlin, double* Jeff) {
    if (!ctx) return MGB_EINVAL;
    LpSpecies* sp = nullptr;
    MGB_TRY(species_of(ctx, l, &sp, "mgb_lp_get_Jeff"));
    if (!sp->have_pop) return ctx->fail(MGB_ESTATE, "mgb_lp_get_Jeff: nothing computed");
    cudaSetDevice(ctx->device);
    const size_t n = (size_t)ctx->m.npoints * sp->nrad;
    if (Jlin) MGB_CUDA(ctx, cudaMemcpy(Jlin, sp->Jlin.p, 8 * n, cudaMemcpyDeviceToHost));
    if (Jeff) MGB_CUDA(ctx, cudaMemcpy(Jeff, sp->Jeff.p, 8 * n, cudaMemcpyDeviceToHost));
    return MGB_OK;
}

// correct_negative_populations + check_for_convergence + set_emissivity_and_opacity: the tail every kind of update shares
// (Lines::iteration_using_statistical_equilibrium / _using_Ng_acceleration, lines.cpp:106-137).
static int finish_update(mgb_ctx* ctx, double pop_prec, mgb_lp_convergence* conv) {
    for (uint32_t l = 0; l < ctx->lp.size(); l++) {
        SpeciesView v;
        MGB_TRY(make_view(ctx, l, v));
        MGB_CUDA(ctx, ctx->lp_flags.ensure((size_t)ctx->m.npoints * v.nlev));
        k_lp_negative<<<(ctx->m.npoints + 127) / 128, 128, 0, ctx->stream>>>(v, ctx->lp_flags.as<unsigned char>());
        MGB_CUDA(ctx, cudaGetLastError());
    }
    MGB_TRY(convergence(ctx, 0, pop_prec, conv));
    MGB_TRY(line_state(ctx));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return MGB_OK;
}

extern "C" int mgb_lp_update_statistical_equilibrium(mgb_ctx* ctx, int keep_history, double pop_prec, mgb_lp_convergence* conv, float* ms) {
    if (!ctx) return MGB_EINVAL;
    MGB_TRY(all_species_ready(ctx, "mgb_lp_update_statistical_equilibrium"));
    NvtxRange nvtx("mgb_lp_update_statistical_equilibrium");
    if (!ctx->have_state) return ctx->fail(MGB_ESTATE, "mgb_lp_update_statistical_equilibrium: no line state");
    cudaSetDevice(ctx->device);
    cudaEvent_t ea = nullptr, eb = nullptr;
    if (ms) { cudaEventCreate(&ea); cudaEventCreate(&eb); cudaEventRecord(ea, ctx->stream); }
    MGB_CUDA(ctx, ctx->counters.ensure(64));
    int* failed = reinterpret_cast<int*>(ctx->counters.as<unsigned long long>() + 6);
    MGB_CUDA(ctx, cudaMemsetAsync(failed, 0, 4, ctx->stream));
    for (uint32_t l = 0; l < ctx->lp.size(); l++) {
        LpSpecies& sp = *ctx->lp[l];
        MGB_TRY(shift_previous(ctx, sp));
        if (keep_history) MGB_TRY(push_history(ctx, sp));
        SpeciesView v;
        MGB_TRY(make_view(ctx, l, v));
        const size_t smem = 16 * (size_t)v.nlev * (v.nlev + 1);
        MGB_CUDA(ctx, cudaFuncSetAttribute(k_lp_stateq, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_lp_stateq<<<ctx->m.npoints, 64, smem, ctx->stream>>>(v, ctx->lam_val.as<double>(), ctx->lam_cnt.as<uint32_t>(), ctx->lrow.as<LineRow>(), failed);
        MGB_CUDA(ctx, cudaGetLastError());
    }
    int h_failed = 0;
    MGB_CUDA(ctx, cudaMemcpyAsync(&h_failed, failed, 4, cudaMemcpyDeviceToHost, ctx->stream));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (h_failed) return ctx->fail(MGB_EINVAL, "Eigen solver ERROR."); /* the text of lineProducingSpecies.tpp:551: a singular rate matrix */
    MGB_TRY(finish_update(ctx, pop_prec, conv));
    if (ms) { cudaEventRecord(eb, ctx->stream); cudaEventSynchronize(eb); cudaEventElapsedTime(ms, ea, eb); cudaEventDestroy(ea); cudaEventDestroy(eb); }
    return MGB_OK;
}

// Gram matrix of the last `order` residuals, double-double (hi, lo) row-major [order][order]. With begin != 0 the call first does what
// both update_using_acceleration and _trial start with: residuals.push_back(population - populations.back()); populations.push_back(population)
// (and, for the real step, the shift of population_prev1..3; for the trial, trial_population_prev = trial_population).
extern "C" int mgb_lp_ng_gram(mgb_ctx* ctx, uint32_t l, uint32_t order, int trial, double* gram_hi, double* gram_lo) {
    if (!ctx || !gram_hi || !gram_lo || !order || order > 16) return MGB_EINVAL;
    MGB_TRY(all_species_ready(ctx, "mgb_lp_ng_gram"));
    LpSpecies* sp = nullptr;
    MGB_TRY(species_of(ctx, l, &sp, "mgb_lp_ng_gram"));
    cudaSetDevice(ctx->device);
    const size_t n = (size_t)ctx->m.npoints * sp->nlev;
    if (!trial) MGB_TRY(shift_previous(ctx, *sp));
    else MGB_TRY(copy_dd(ctx, sp->trial_prev, sp->trial, n));
    MGB_TRY(push_history(ctx, *sp));
    if (sp->populations.size() < sp->residuals.size() + 1) return ctx->fail(MGB_ESTATE, "Error during ng acceleration; inconstent sizes.");
    if ((trial ? sp->populations.size() : sp->residuals.size()) < order) /* lineProducingSpecies.tpp:283 (real step) and :335 (trial) */
        return ctx->fail(MGB_ESTATE, "Error during ng acceleration; order of acceleration greather than data stored.");
    if (sp->residuals.size() < order) return ctx->fail(MGB_ESTATE, "mgb_lp_ng_gram: %zu residuals stored, order %u", sp->residuals.size(), order);
    MGB_CUDA(ctx, ctx->lp_partials.ensure(16 * (size_t)kLpBlocks * 4));
    std::vector<double2> h(kLpBlocks);
    const size_t rs = sp->residuals.size();
    for (uint32_t i = 0; i < order; i++)
        for (uint32_t j = i; j < order; j++) { /* the Gram matrix is symmetric; the reference computes both halves with the same products */
            k_lp_dot<<<kLpBlocks, 256, 0, ctx->stream>>>(n, sp->residuals[rs - order + i]->as<double2>(), sp->residuals[rs - order + j]->as<double2>(),
                                                         ctx->lp_partials.as<double2>());
            MGB_CUDA(ctx, cudaGetLastError());
            MGB_CUDA(ctx, cudaMemcpyAsync(h.data(), ctx->lp_partials.p, 16 * (size_t)kLpBlocks, cudaMemcpyDeviceToHost, ctx->stream));
            MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            long double acc = 0.0L;
            for (int b = 0; b < kLpBlocks; b++) acc += (long double)h[b].x + (long double)h[b].y;
            const double hi = (double)acc, lo = (double)(acc - (long double)hi);
            gram_hi[i * order + j] = gram_hi[j * order + i] = hi;
            gram_lo[i * order + j] = gram_lo[j * order + i] = lo;
        }
    return MGB_OK;
}

// population (or trial_population) = sum_i populations[size-1-i] * coef[order-1-i]; then the history is trimmed like the reference:
// real step: residuals.clear(), populations keeps its last entry (lineProducingSpecies.tpp:311-313); trial: both pop_back (:366-369).
extern "C" int mgb_lp_ng_combine(mgb_ctx* ctx, uint32_t l, uint32_t order, int trial, const double* coef_hi, const double* coef_lo) {
    if (!ctx || !coef_hi || !coef_lo || !order || order > 16) return MGB_EINVAL;
    LpSpecies* sp = nullptr;
    MGB_TRY(species_of(ctx, l, &sp, "mgb_lp_ng_combine"));
    if (sp->populations.size() < order) return ctx->fail(MGB_ESTATE, "mgb_lp_ng_combine: history shorter than the order");
    cudaSetDevice(ctx->device);
    const size_t n = (size_t)ctx->m.npoints * sp->nlev, ps = sp->populations.size();
    CombineArgs a;
    a.order = order;
    for (uint32_t i = 0; i < order; i++) {
        a.P[i] = sp->populations[ps - 1 - i]->as<double2>();
        a.c[i] = make_double2(coef_hi[order - 1 - i], coef_lo[order - 1 - i]);
    }
    k_lp_combine<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(n, a, trial ? sp->trial.as<double2>() : sp->pop.as<double2>());
    MGB_CUDA(ctx, cudaGetLastError());
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (trial) {
        delete sp->residuals.back(); sp->residuals.pop_back();
        delete sp->populations.back(); sp->populations.pop_back();
    } else {
        for (auto* b : sp->residuals) delete b;
        sp->residuals.clear();
        for (size_t i = 0; i + 1 < sp->populations.size(); i++) delete sp->populations[i];
        sp->populations.erase(sp->populations.begin(), sp->populations.end() - 1);
    }
    return MGB_OK;
}

extern "C" int mgb_lp_finish_update(mgb_ctx* ctx, double pop_prec, mgb_lp_convergence* conv) {
    if (!ctx) return MGB_EINVAL;
    MGB_TRY(all_species_ready(ctx, "mgb_lp_finish_update"));
    cudaSetDevice(ctx->device);
    return finish_update(ctx, pop_prec, conv);
}

extern "C" int mgb_lp_trial_convergence(mgb_ctx* ctx, double pop_prec, mgb_lp_convergence* conv) {
    if (!ctx || !conv) return MGB_EINVAL;
    MGB_TRY(all_species_ready(ctx, "mgb_lp_trial_convergence"));
    cudaSetDevice(ctx->device);
    return convergence(ctx, 1, pop_prec, conv);
}
