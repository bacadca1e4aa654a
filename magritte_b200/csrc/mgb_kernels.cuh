// mgb_kernels.cuh -- the CUDA kernels of the Feautrier order-2 path (sm_100a).
//
//   k_prepare        per-(point,line) packed LineRow + per-point derived tables
//   k_trace<FILL>    one warp per half-ray (o, r): count pass / fill pass of the geometric path
//   k_solve<APPROX>  one warp per (o, rr, 32 frequencies): lanes = frequency bins; eta/chi assembly and
//                    the two-sided Thomas sweep entirely in registers
//   k_column         one warp per half-ray: column density of Model::density (Solver::get_column)
//   k_reduce_J / k_reduce_lambda   J(o,f) and Lambda(o,k) accumulated in the reference's summation order
#pragma once
#include "mgb_device.cuh"

namespace mgb {

// Geometric path of one origin chunk. Half-ray (o, r) owns steps [poff[(o-o0)*nrays+r], +len).
struct PathView {
    uint32_t o0, no;
    const uint32_t* len;      // [npoints][nrays] geometric steps per half-ray (global, not chunk-local)
    const unsigned long long* poff; // [no*nrays] chunk-local exclusive prefix of len
    uint32_t* nr;             // [steps] next point of each step
    double* dZ;               // [steps]
    double* shift;            // [steps] Doppler shift of that next point
    double* shift0;           // [no*nrays] shift of the origin itself (get_shift(o, r, o, 0))
};

struct ScratchView {
    double* u;     // [hnrays][no][nfreqs]   Su[centre] (or the boundary intensity for an isolated origin)
    double* L;     // [hnrays][no][nfreqs]   the Lambda increment of (rr, o, f)
    unsigned long long* counters; // [0] entries (Sigma N), [1] large-shift segment evaluations
    // banded Lambda operator (n_off_diag = K > 0): per (rr, o, f) the 2K off-diagonal increments in the order update_Lambda adds
    // them (m = 0..K-1: lower, upper), and the point each belongs to (0xffffffff: no such element)
    double* offV;   // [hnrays][no][nfreqs][2K]
    uint32_t* offN; // [hnrays][no][nfreqs][2K]
    uint32_t K;
};
constexpr int kMaxOffDiag = 4; // largest supported n_off_diag

// ---------------------------------------------------------------------------------------------------
__global__ void k_prepare_rows(uint32_t npoints, uint32_t nlines, const double* __restrict__ emis, const double* __restrict__ opac,
                               const double* __restrict__ invw, const double* __restrict__ pop_tot, const uint32_t* __restrict__ line_spec,
                               LineRow* __restrict__ rows) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)npoints * nlines) return;
    const uint32_t p = (uint32_t)(i / nlines), l = (uint32_t)(i % nlines);
    LineRow r;
    r.inverse_width = invw[i];
    r.emissivity    = emis[i];
    r.opacity       = opac[i];
    // InterpHelper: interpolation_criterion = opacity / population_tot (solver_interp_helper.hpp:15-30)
    r.criterion = opac[i] / pop_tot[(size_t)line_spec[l] * npoints + p];
    rows[i]     = r;
}

// sqrtw_max (CloseLines window, thermodynamics.tpp:42-53) and Model::set_dshift_max (model.cpp:1158-1178)
__global__ void k_prepare_points(uint32_t npoints, uint32_t nlspecs, const double* __restrict__ T, const double* __restrict__ vturb2,
                                 const double* __restrict__ inverse_mass, double max_inverse_mass, double max_width_fraction,
                                 double* __restrict__ sqrtw_max, double* __restrict__ dshift_max) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npoints) return;
    sqrtw_max[p] = sqrt(kTWO_KB_OVER_AMU_CC2 * max_inverse_mass * T[p] + vturb2[p]);
    double d = DBL_MAX;
    for (uint32_t l = 0; l < nlspecs; l++) {
        const double nd = max_width_fraction * sqrt(kTWO_KB_OVER_AMU_CC2 * inverse_mass[l] * T[p] + vturb2[p]);
        if (d > nd) d = nd;
    }
    dshift_max[p] = d;
}

// ---------------------------------------------------------------------------------------------------
// Tracer. reference: Solver::trace_ray_for_line (solver.tpp:773-800) minus the per-line sub-stepping
// (which k_solve redoes per frequency lane), Solver::get_ray_length (:79-103).
// FILL = false: lens[o][r] = number of valid steps, plus Sigma S / Sigma D counters.
// FILL = true : write (nr, dZ, shift) of every step at the half-ray's offset.
// ---------------------------------------------------------------------------------------------------
template <bool FILL>
__global__ void __launch_bounds__(128) k_trace(DevModel m, uint32_t o0, uint32_t no, uint32_t* __restrict__ lens, PathView pv,
                                               unsigned long long* __restrict__ counters) {
    const unsigned lane = threadIdx.x & 31;
    const size_t w      = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= (size_t)no * m.nrays) return;
    const uint32_t ol = (uint32_t)(w / m.nrays), r = (uint32_t)(w % m.nrays);
    const uint32_t o  = o0 + ol;
    const RayFrame fr = make_frame(m, o, r);
    unsigned long long off = 0;
    if (FILL) {
        off = pv.poff[w];
        if (lane == 0) pv.shift0[w] = get_shift_comoving(m, fr, o, 0.0);
    }
    double Z = 0.0, dZ = 0.0;
    uint32_t len = 0, scanned = 0, ns;
    uint32_t nxt = warp_get_next(m, fr, o, Z, dZ, ns, lane);
    scanned += ns;
    if (nxt < m.npoints) {
        for (;;) {
            if (FILL && lane == 0) {
                pv.nr[off + len]    = nxt;
                pv.dZ[off + len]    = dZ;
                pv.shift[off + len] = get_shift_comoving(m, fr, nxt, Z);
            }
            len++;
            if (__ldg(&m.p2b[nxt]) != m.npoints) break; /* reached the boundary (geometry.tpp:233) */
            const uint32_t crt = nxt;
            nxt = warp_get_next(m, fr, crt, Z, dZ, ns, lane);
            scanned += ns;
            if (!(nxt < m.npoints)) break; /* dead end */
        }
    }
    if (!FILL && lane == 0) {
        lens[(size_t)o * m.nrays + r] = len;
        atomicAdd(&counters[0], (unsigned long long)len);
        atomicAdd(&counters[1], (unsigned long long)scanned);
        atomicMax((unsigned int*)&counters[2], len);
    }
}

// Column density. reference: Solver::get_column (solver.tpp:2968-2991), driven by Solver::set_column (:2953-2966).
// One warp per half-ray (o, r) on the same walker as k_trace; lane 0 keeps the running trapezoid sum in the reference's
// order (outward from the origin). column is [nrays][no] (Model::column rows of this origin range).
__global__ void __launch_bounds__(128) k_column(DevModel m, uint32_t o0, uint32_t no, const double* __restrict__ density,
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
        double rho_crt = __ldg(&density[o]);
        for (;;) {
            const double rho_nxt = __ldg(&density[nxt]);
            col = __dadd_rn(col, __dmul_rn(__dmul_rn(0.5, __dadd_rn(rho_crt, rho_nxt)), dZ));
            if (__ldg(&m.p2b[nxt]) != m.npoints) break; /* reached the boundary */
            rho_crt = rho_nxt;
            nxt     = warp_get_next(m, fr, nxt, Z, dZ, ns, lane);
            if (!(nxt < m.npoints)) break; /* dead end (the reference would index density[npoints]) */
        }
    }
    if (lane == 0) column[(size_t)r * no + ol] = col;
}

// per-origin total of lens over all rays (chunk planning)
__global__ void k_origin_totals(uint32_t o0, uint32_t no, uint32_t nrays, const uint32_t* __restrict__ lens,
                                unsigned long long* __restrict__ tot) {
    const uint32_t ol = blockIdx.x * blockDim.x + threadIdx.x;
    if (ol >= no) return;
    unsigned long long s = 0;
    for (uint32_t r = 0; r < nrays; r++) s += lens[(size_t)(o0 + ol) * nrays + r];
    tot[ol] = s;
}

} // namespace mgb
