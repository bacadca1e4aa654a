// mgb_solve4.cuh -- the Feautrier order-2 pass as a PIPELINE of small kernels that stream through HBM (n_off_diag = 0).
//
// k_solve / k_solve3 keep everything of a work item in one kernel: the ray records of phase A in shared memory, the per-bin sweep in
// registers, the line-integrated (erff) branch inline. ncu on those kernels (profiles/ncu_k_solve3_r2.txt) shows what that costs on
// B200: 128 registers and 13.7 KB of shared memory per warp (16 warps per SM), three code regions (phase A, common step, erff branch)
// that evict each other from the instruction cache (stall_no_instruction is the top stall, one sample group per 128-byte line of
// straight-line code), phase A at 14 of 32 lanes, the erff branch at 22 -- and HBM idle (0.1 % of its bandwidth). The pipeline trades
// that idle bandwidth for regular kernels; every per-entry number is computed by the same functions in the same operation order, so
// the results are bit-identical to k_solve (tests/test_gpu_parity.py runs every golden through all solve kernels).
//
//   k_ni / k_fill      one warp per half ray, lanes = (step, line) pairs: the sub-stepped ray entries of every line (Solver::
//                      set_data_for_line, solver.tpp:1057-1097; InterpHelper::get_n_interp_for_line, solver_interp_helper.tpp:55-75),
//                      written as 64-byte records [shift, dz, inverse_width, emissivity, opacity, sqrtw_max, line source, flag word]
//                      in walking order (far end -> origin, the origin's own entry last) + 16 bytes for the generic paths
//                      (interpolation factor, the two points). The flag word holds the large-shift test of the segment that ends at
//                      the entry (compute_source_dtau, solver.tpp:2150-2153) and, if set, the slot of that segment's line integrals.
//   k_large            one thread per (large-shift segment, frequency bin): compute_S_dtau_line_integrated (solver.tpp:1878-2111,
//                      2374-2455) -> (dtau, S at the far end, S at the near end), all windows (own line, empty, generic).
//   k_sweep            one warp per (origin, ray pair, <= 32 bins), lanes = bins: streams the records of its lines (prefetched one
//                      entry ahead), evaluates eta / chi / S (get_eta_and_chi_interpolated + the division, solver.tpp:2166-2176),
//                      reads the line integrals of flagged segments, and runs the two-sided elimination in registers. No shared
//                      memory beyond 1.5 KB per warp, no phase A, no erff code.
// Algorithmic HBM traffic per record: 80 B written + 64 B read per sweeping warp; per large-shift segment and bin 24 B written + read.
#pragma once
#include "mgb_solve3.cuh"

namespace mgb {

constexpr unsigned long long kNoSlot = ~0ull; // flag word of an entry whose segment takes the small-shift branch

struct RecView {
    uint32_t ola, nob;              // origins [o0 + ola, o0 + ola + nob) of the path chunk
    const unsigned long long* off;  // [nob * nrays * nlines + 1] exclusive prefix of the list lengths; list id = ((ol - ola) * nrays + r) * nlines + l
    double* rec;                    // [nrec][8]
    double2* pts;                   // [nrec] {interpolation factor, bits of (nr, nr_interp)}
    unsigned long long* lcount;     // slots handed out so far
    struct SlotInfo* slots;         // [nrec] (capacity: every record could be flagged) slot -> segment
    const uint8_t* side_of;         // [nrays] 0: the pair's ray rr, 1: its antipode
};

struct SlotInfo { unsigned long long recY; uint32_t o, lside; }; // lside = line | side << 31

// Everything of one ray entry that does not depend on the frequency bin: make_record of mgb_solve.cuh, in registers.
struct RecVals { double shift, dz, iw, ie, io, sline, factor; };
__device__ __forceinline__ RecVals compute_record(const DevModel& m, uint32_t l, const StepData& sd, uint32_t i, uint32_t ni) {
    RecVals v;
    v.dz = sd.dZ; v.factor = 0.0; v.shift = sd.sh_nxt;
    if (ni > 1) {
        v.dz = sd.dZ / ni;
        const double dshi = (sd.sh_nxt - sd.sh_crt) / ni;
        v.factor = 1.0 - ((double)i) / ni;
        v.shift  = sd.sh_crt + i * dshi;
    }
    const LineRow a = load_row(m, sd.nxt, l);
    v.iw = a.inverse_width; v.ie = a.emissivity; v.io = a.opacity;
    v.sline = ddiv(a.emissivity, a.opacity);
    if (v.factor != 0.0) {
        const LineRow b = load_row(m, sd.crt, l);
        v.iw    = interp_linear(a.inverse_width, b.inverse_width, v.factor);
        v.ie    = interp_log(a.emissivity, b.emissivity, v.factor);
        v.io    = interp_log(a.opacity, b.opacity, v.factor);
        v.sline = interp_log(v.sline, ddiv(b.emissivity, b.opacity), v.factor);
    }
    return v;
}

// ---------------------------------------------------------------------------------------------------
// Dense record production: lanes = (step, line) pairs of one half ray (consecutive lanes = consecutive lines of the same step, so
// the LineRow / criterion loads coalesce), instead of k_rec's lanes = steps or entries of ONE line (13-20 of 32 lanes on typical rays).
//   k_ecount : elements per half ray, (len + 1) * nlines (the + 1 is the origin's own entry), 0 for a half ray without steps
//   k_ni     : the sub-step count of every element, stored list-major: ni[eoff[hr] + l * (len + 1) + k], k = 0 the far end, k = len
//              the origin; an exclusive sum over that array is the record index of each element's first entry
//   k_off    : list offsets for k_slots / k_sweep
//   k_fill   : every thread writes the ni records of its element
// ---------------------------------------------------------------------------------------------------
__global__ void k_ecount(uint32_t nhr, uint32_t nrays, uint32_t nlines, uint32_t o_first, const uint32_t* __restrict__ len_all, uint32_t* __restrict__ ecnt) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > nhr) return;
    uint32_t c = 0;
    if (i < nhr) {
        const uint32_t len = len_all[(size_t)o_first * nrays + i];
        c = len ? (len + 1) * nlines : 0;
    }
    ecnt[i] = c; /* ecnt[nhr] = 0: the scan's last element is the total */
}

__device__ __forceinline__ void st_global_v4(double* p, double a, double b, double c, double d) {
    asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}

struct DenseView {
    const unsigned long long* eoff; // [nhr + 1]
    uint32_t* ni;                   // [E + 1]
    const unsigned long long* base; // [E + 1] exclusive sum of ni
};

__device__ __forceinline__ StepData load_step(const DevModel& m, const PathView& pv, uint32_t o, unsigned long long off, double shift0, uint32_t js) {
    StepData sd;
    sd.nxt    = __ldg(&pv.nr[off + js]);
    sd.crt    = js ? __ldg(&pv.nr[off + js - 1]) : o;
    sd.dZ     = __ldg(&pv.dZ[off + js]);
    sd.sh_nxt = __ldg(&pv.shift[off + js]);
    sd.sh_crt = js ? __ldg(&pv.shift[off + js - 1]) : shift0;
    sd.dsm    = __ldg(&m.dshift_max[sd.nxt]);
    sd.sqw    = __ldg(&m.sqrtw_max[sd.nxt]);
    return sd;
}

// one warp per half ray
__global__ void __launch_bounds__(128) k_ni(const __grid_constant__ DevModel m, const __grid_constant__ PathView pv, uint32_t ola, uint32_t nob, DenseView dv) {
    const unsigned lane = threadIdx.x & 31;
    const size_t hrl = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    stage_exp_table();
    if (hrl >= (size_t)nob * m.nrays) return;
    const uint32_t r = (uint32_t)(hrl % m.nrays), ol = ola + (uint32_t)(hrl / m.nrays), o = pv.o0 + ol;
    const uint32_t len = __ldg(&pv.len[(size_t)o * m.nrays + r]);
    if (len == 0) return;
    const unsigned long long off = __ldg(&pv.poff[(size_t)ol * m.nrays + r]);
    uint32_t* out = dv.ni + dv.eoff[hrl];
    const uint32_t npairs = len * m.nlines;
    for (uint32_t t = lane; t < npairs; t += 32) {
        const uint32_t k = t / m.nlines, l = t - k * m.nlines, js = len - 1 - k;
        const uint32_t nxt = __ldg(&pv.nr[off + js]);
        const uint32_t crt = js ? __ldg(&pv.nr[off + js - 1]) : o;
        const double crit_nxt = __ldg(&m.lrow[(size_t)nxt * m.nlines + l].criterion);
        const double crit_crt = __ldg(&m.lrow[(size_t)crt * m.nlines + l].criterion);
        out[l * (len + 1) + k] = n_interp_from_criteria(m, crit_crt, crit_nxt);
    }
    for (uint32_t l = lane; l < m.nlines; l += 32) out[l * (len + 1) + len] = 1; /* the origin's own entry */
}

__global__ void k_off(const __grid_constant__ DevModel m, const __grid_constant__ PathView pv, uint32_t ola, uint32_t nob, DenseView dv, unsigned long long nelem,
                      unsigned long long* __restrict__ off) {
    const size_t list = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t nlists = (size_t)nob * m.nrays * m.nlines;
    if (list > nlists) return;
    if (list == nlists) { off[list] = dv.base[nelem]; return; }
    const uint32_t l = (uint32_t)(list % m.nlines);
    const size_t hrl = list / m.nlines;
    const uint32_t r = (uint32_t)(hrl % m.nrays), o = pv.o0 + ola + (uint32_t)(hrl / m.nrays);
    const uint32_t len = __ldg(&pv.len[(size_t)o * m.nrays + r]);
    off[list] = dv.base[dv.eoff[hrl] + (len ? (unsigned long long)l * (len + 1) : 0ull)];
}

// one warp per half ray; a thread owns the entries of one (step, line) element
#ifndef MGB_FILL_CTAS
#define MGB_FILL_CTAS 5
#endif
__global__ void __launch_bounds__(128, MGB_FILL_CTAS) k_fill(const __grid_constant__ DevModel m, const __grid_constant__ PathView pv, RecView rv, DenseView dv) {
    const unsigned lane = threadIdx.x & 31;
    const size_t hrl = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    stage_exp_table();
    if (hrl >= (size_t)rv.nob * m.nrays) return;
    const uint32_t r = (uint32_t)(hrl % m.nrays), ol = rv.ola + (uint32_t)(hrl / m.nrays), o = pv.o0 + ol;
    const uint32_t len = __ldg(&pv.len[(size_t)o * m.nrays + r]);
    if (len == 0) return;
    const size_t hr = (size_t)ol * m.nrays + r;
    const unsigned long long off = __ldg(&pv.poff[hr]);
    const double shift0 = __ldg(&pv.shift0[hr]);
    const unsigned long long e0 = dv.eoff[hrl];
    const uint32_t nel = (len + 1) * m.nlines;
    for (uint32_t t0 = 0; t0 < nel; t0 += 32) {
        const uint32_t t = t0 + lane;
        const bool valid = t < nel;
        uint32_t k = 0, l = 0, ni = 0;
        unsigned long long out = 0;
        StepData sd;
        sd.nxt = o; sd.crt = o; sd.dZ = 0.0; sd.sh_nxt = 1.0; sd.sh_crt = 1.0; sd.dsm = 0.0; sd.sqw = 0.0;
        double prev_shift = 0.0, prev_dsm = 0.0;
        bool have_prev = false;
        if (valid) {
            k = t / m.nlines; l = t - k * m.nlines;
            const unsigned long long el = e0 + (unsigned long long)l * (len + 1) + k;
            ni  = dv.ni[el];
            out = dv.base[el];
            if (k < len) sd = load_step(m, pv, o, off, shift0, len - 1 - k);
            else { sd.dsm = __ldg(&m.dshift_max[o]); sd.sqw = __ldg(&m.sqrtw_max[o]); }
            if (k > 0) { /* the entry before this element's first: the last sub-entry (i = 1) of the previous step */
                const uint32_t jp = len - k; /* step index of element k - 1 */
                const uint32_t nip = dv.ni[el - 1];
                const double sh_n = __ldg(&pv.shift[off + jp]), sh_c = jp ? __ldg(&pv.shift[off + jp - 1]) : shift0;
                prev_shift = sh_n;
                if (nip > 1) { const double dshi = (sh_n - sh_c) / nip; prev_shift = sh_c + 1 * dshi; }
                prev_dsm  = __ldg(&m.dshift_max[__ldg(&pv.nr[off + jp])]);
                have_prev = true;
            }
        }
        const uint32_t side_bit = (uint32_t)__ldg(&rv.side_of[r]) << 31;
        // stores one record (and, for a flagged segment, its slot table entry); slots are handed out per warp
        auto emit = [&](bool act, const RecVals& v, bool large, unsigned long long at, uint32_t line, uint32_t crt, uint32_t nxt, double sqw) {
            unsigned long long word = kNoSlot;
            const unsigned lmask = __ballot_sync(0xffffffffu, large);
            if (lmask) {
                unsigned long long b = 0;
                if (lane == 0) b = atomicAdd(rv.lcount, (unsigned long long)__popc(lmask));
                b = __shfl_sync(0xffffffffu, b, 0);
                if (large) {
                    word = b + __popc(lmask & ((1u << lane) - 1u));
                    SlotInfo si;
                    si.recY = at; si.o = o; si.lside = line | side_bit;
                    rv.slots[word] = si;
                }
            }
            if (act) { /* two 32-byte stores: whole sectors, no read-for-ownership in L2 (sm_100 256-bit vector stores) */
                double* rp = rv.rec + at * 8;
                st_global_v4(rp, v.shift, v.dz, v.iw, v.ie);
                st_global_v4(rp + 4, v.io, sqw, v.sline, __longlong_as_double((long long)word));
                rv.pts[at] = make_double2(v.factor, __hiloint2double((int)crt, (int)nxt));
            }
        };
        // ---- the first entry of every element (sub-index i = ni: the far point of the step itself, no interpolation) ----
        {
            RecVals v;
            v.shift = 0.0; v.dz = 0.0; v.iw = 0.0; v.ie = 0.0; v.io = 0.0; v.sline = 0.0; v.factor = 0.0;
            bool large = false;
            if (valid) {
                v = compute_record(m, l, sd, ni, ni);
                if (have_prev) {
                    const double dshift_max = (sd.dsm < prev_dsm) ? sd.dsm : prev_dsm;
                    large = fabs(v.shift - prev_shift) > dshift_max;
                }
            }
            emit(valid, v, large, out, l, sd.crt, sd.nxt, sd.sqw);
        }
        // ---- the sub-step entries (i = ni - 1 .. 1) of all elements of the batch, compacted: lanes = entries. They are the expensive
        //      ones (three powf interpolations each) and only some elements have them: looping per element left 8 of 32 lanes busy ----
        const uint32_t extra = valid ? ni - 1 : 0u;
        uint32_t incl = extra;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t x = __shfl_up_sync(0xffffffffu, incl, d);
            if ((int)lane >= d) incl += x;
        }
        const uint32_t ex = incl - extra, T = __shfl_sync(0xffffffffu, incl, 31);
        for (uint32_t q0 = 0; q0 < T; q0 += 32) {
            const uint32_t q = q0 + lane;
            const bool act = q < T;
            int ow = 0; /* owner: the last lane whose exclusive prefix is <= q */
#pragma unroll
            for (int d = 16; d; d >>= 1) {
                const uint32_t e2 = __shfl_sync(0xffffffffu, ex, (ow + d) & 31);
                if (ow + d < 32 && e2 <= q) ow += d;
            }
            StepData so;
            so.nxt = __shfl_sync(0xffffffffu, sd.nxt, ow); so.crt = __shfl_sync(0xffffffffu, sd.crt, ow);
            so.dZ = __shfl_sync(0xffffffffu, sd.dZ, ow); so.sh_nxt = __shfl_sync(0xffffffffu, sd.sh_nxt, ow);
            so.sh_crt = __shfl_sync(0xffffffffu, sd.sh_crt, ow); so.dsm = __shfl_sync(0xffffffffu, sd.dsm, ow);
            so.sqw = __shfl_sync(0xffffffffu, sd.sqw, ow);
            const uint32_t ni_o = __shfl_sync(0xffffffffu, ni, ow), ex_o = __shfl_sync(0xffffffffu, ex, ow), l_o = __shfl_sync(0xffffffffu, l, ow);
            const unsigned long long out_o = __shfl_sync(0xffffffffu, out, ow);
            RecVals v;
            v.shift = 0.0; v.dz = 0.0; v.iw = 0.0; v.ie = 0.0; v.io = 0.0; v.sline = 0.0; v.factor = 0.0;
            bool large = false;
            const uint32_t it = 1 + (q - ex_o);
            if (act) {
                const uint32_t i = ni_o - it;
                v = compute_record(m, l_o, so, i, ni_o);
                /* the entry before it is sub-entry i + 1 of the same step: compute_record's shift expression for i + 1 */
                const double dshi = (so.sh_nxt - so.sh_crt) / ni_o;
                const double pshift = so.sh_crt + (i + 1) * dshi;
                large = fabs(v.shift - pshift) > so.dsm; /* min(dshift_max, dshift_max) of the one point */
            }
            emit(act, v, large, out_o + it, l_o, so.crt, so.nxt, so.sqw);
        }
    }
}

// (dtau, S at the far end X, S at the near end Y) of every large-shift segment and bin: compute_source_dtau's line-integrated branch
// (solver.tpp:2150-2153 -> compute_S_dtau_line_integrated :1878-2111) exactly as sweep_step of mgb_solve.cuh evaluates it.
// trip [slot][nquads][3]. One thread per (slot, bin of the slot's line).
#ifndef MGB_LARGE_CTAS
#define MGB_LARGE_CTAS 8
#endif
template <int APPROX>
__global__ void __launch_bounds__(128, MGB_LARGE_CTAS) k_large(const __grid_constant__ DevModel m, const double* __restrict__ rec, const double2* __restrict__ pts,
                                               const SlotInfo* __restrict__ slots, unsigned long long nslots, double* __restrict__ trip) {
    stage_exp_table();
    stage_erff_table();
    /* a block takes blockDim / nquads whole slots (no 64-bit division per thread); the remaining threads idle */
    const uint32_t spb = blockDim.x / m.nquads, sib = threadIdx.x / m.nquads;
    const unsigned long long slot = (unsigned long long)blockIdx.x * spb + sib;
    const uint32_t z = threadIdx.x - sib * m.nquads;
    if (sib >= spb || slot >= nslots) return;
    const SlotInfo s = slots[slot];
    const uint32_t l = s.lside & 0x7fffffffu, side = s.lside >> 31;
    const uint32_t f = l * m.nquads + z;
    const uint32_t ll = (APPROX == APPROX_ONELINE) ? __ldg(&m.corr_line[f]) : l;
    const double freq = __ldg(&m.nu[(size_t)s.o * m.nfreqs + f]);
    const double* y = rec + s.recY * 8;
    const double* x = y - 8;
    const double shX = x[0], dzX = x[1], shY = y[0];
    const double frX = freq * shX, frY = freq * shY;
    LaneLine L;
    {
        const uint32_t pos = __ldg(&m.sorted_pos[l]);
        L.s_own  = __ldg(&m.line[l]);
        L.s_prev = pos ? __ldg(&m.sorted_line[pos - 1]) : -CUDART_INF;
        L.s_next = (pos + 1 < m.nlines) ? __ldg(&m.sorted_line[pos + 1]) : CUDART_INF;
        L.own    = (APPROX == APPROX_CLOSELINES) || (APPROX == APPROX_ONELINE && ll == l);
    }
    int cls = WIN_OTHER;
    if (APPROX == APPROX_CLOSELINES) {
        const double left_freq = (frX < frY) ? frX : frY, right_freq = (frX < frY) ? frY : frX;
        const double bX = m.max_dist * (right_freq * x[5]);
        const double bY = m.max_dist * (right_freq * y[5]);
        const double ub = (bX < bY) ? bY : bX;
        cls = window_class(L, left_freq - ub, right_freq + ub);
    } else if (APPROX == APPROX_ONELINE) {
        cls = L.own ? WIN_OWN : WIN_OTHER;
    }
    double dtau, S_atX, S_atY;
    if (cls == WIN_OWN) {
        const LineIntegral li = line_integrated_own3<APPROX>(m.min_opacity, L.s_own, frX, frY, dzX, x[2], y[2], x[4], y[4], x[6], y[6]);
        dtau = li.dtau; S_atX = li.S_atX; S_atY = li.S_atY;
    } else if (cls == WIN_EMPTY) { /* no line in reach: sums stay 0, dtau is bounded below by min_opacity*dz (solver.tpp:2095-2110) */
        double d = 0.0;
        const double bound_min_dtau = m.min_opacity * dzX;
        if (-bound_min_dtau < d) d = (bound_min_dtau < d) ? d : bound_min_dtau;
        dtau = d; S_atX = zero_over(d); S_atY = S_atX;
    } else {
        const double2 qX = pts[s.recY - 1], qY = pts[s.recY];
        const uint32_t pXx = (uint32_t)__double2loint(qX.y), pXy = (uint32_t)__double2hiint(qX.y);
        const uint32_t pYx = (uint32_t)__double2loint(qY.y), pYy = (uint32_t)__double2hiint(qY.y);
        if (side == 0) S_dtau_generic<APPROX>(m, pXx, pXy, qX.x, pYx, pYy, qY.x, ll, frX, frY, dzX, dtau, S_atX, S_atY);
        else S_dtau_generic<APPROX>(m, pYx, pYy, qY.x, pXx, pXy, qX.x, ll, frY, frX, dzX, dtau, S_atY, S_atX);
    }
    double* t = trip + (slot * m.nquads + z) * 3;
    t[0] = dtau; t[1] = S_atX; t[2] = S_atY;
}

// ---------------------------------------------------------------------------------------------------
// the sweep
// ---------------------------------------------------------------------------------------------------
struct Rec4 { double shift, dz, iw, ie, io, sqw; unsigned long long word; };
__device__ __forceinline__ Rec4 load_rec(const double* __restrict__ rec, unsigned long long i) {
    const double2* p = reinterpret_cast<const double2*>(rec + i * 8);
    const double2 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    Rec4 r;
    r.shift = a.x; r.dz = a.y; r.iw = b.x; r.ie = b.y; r.io = c.x; r.sqw = c.y;
    r.word = (unsigned long long)__double_as_longlong(__ldg(rec + i * 8 + 7));
    return r;
}
struct Trip { double dtau, SX, SY; };

// The general step (first entry, boundary row, windows holding other lines, ApproximationType None): sweep_step of mgb_solve.cuh on
// streamed records. X's record index is i - 1. Out of line, state by value.
template <int APPROX>
__device__ __noinline__ Lane3 step4_general(const DevModel& m, LaneLine L, Lane3 st, double freq, int side, const double* __restrict__ rec, double dzX,
                                            Rec4 y, Trip tr, unsigned long long i, const double2* __restrict__ pts, uint32_t ll) {
    const double frY = freq * y.shift;
    double chiY, SY;
    {
        int cls = WIN_OTHER;
        if (APPROX == APPROX_CLOSELINES) {
            const double ub = m.max_dist * (frY * y.sqw);
            cls = window_class(L, frY - ub, frY + ub);
        } else if (APPROX == APPROX_ONELINE) {
            cls = L.own ? WIN_OWN : WIN_OTHER;
        }
        if (cls == WIN_OWN) {
            const double prof = gaussian(y.iw, frY - L.s_own);
            const double eta  = 0.0 + prof * y.ie;
            if (APPROX == APPROX_ONELINE) chiY = (0.0 + prof * y.io) + m.min_opacity;
            else chiY = m.min_opacity + prof * y.io;
            SY = ddiv(eta, chiY);
        } else if (cls == WIN_EMPTY) {
            chiY = m.min_opacity;
            SY   = zero_over(chiY);
        } else {
            const double2 qY = pts[i];
            double eta;
            eta_chi_generic<APPROX>(m, (uint32_t)__double2loint(qY.y), (uint32_t)__double2hiint(qY.y), qY.x, ll, frY, eta, chiY);
            SY = ddiv(eta, chiY);
        }
    }
    if (st.haveX) {
        const bool large = y.word != kNoSlot;
        double dtau, S_atX = 0.0, S_atY = 0.0;
        if (large) { st.nlarge++; dtau = tr.dtau; S_atX = tr.SX; S_atY = tr.SY; }
        else dtau = 0.5 * (st.chiX + chiY) * dzX;
        if (st.X_is_end) { // boundary row (solver.tpp:2240-2262 for `first`, 2283-2297 mirrored for `last`)
            const double S0  = large ? S_atX : st.SX;
            const double inv = drcp(dtau);
            const double Cb  = 2.0 * inv * inv;
            const double Bm  = 1.0 + 2.0 * inv;
            const double B   = Bm + Cb;
            const double2 qX = pts[i - 1];
            const double Ib  = boundary_intensity(m, (uint32_t)__double2loint(qX.y), freq * rec[(i - 1) * 8]); /* X's shift */
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

// 9 CTAs of 2 warps (96 registers, 12 B of spill outside the loop) measured best: 389 ms against 408 at 8 x 128 registers,
// 398 at 10, 447 at 11 (80 registers, spills in the loop), 393 at 18 CTAs of one warp.
#ifndef MGB_SWEEP_WARPS
#define MGB_SWEEP_WARPS 2
#endif
#ifndef MGB_SWEEP_CTAS
#define MGB_SWEEP_CTAS 9
#endif
constexpr int kSweepWarps = MGB_SWEEP_WARPS, kSweepCtas = MGB_SWEEP_CTAS;

#ifndef MGB_SWEEP_UNROLL
#define MGB_SWEEP_UNROLL 1
#endif
#ifndef MGB_RING_DEPTH
#define MGB_RING_DEPTH 4
#endif
constexpr int kRingDepth = MGB_RING_DEPTH, kSweepUnroll = MGB_SWEEP_UNROLL;
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int APPROX>
__global__ void __launch_bounds__(32 * kSweepWarps, kSweepCtas) k_sweep(const __grid_constant__ DevModel m, const __grid_constant__ PathView pv, ScratchView sc,
                                                                        const FreqChunk* __restrict__ chunks, uint32_t nchunks_f, RecView rv,
                                                                        const double* __restrict__ trip) {
    __shared__ double s_side0[kSweepWarps][6][32];
    __shared__ __align__(16) double s_ring[kSweepWarps][kRingDepth][8][8]; /* records of the warp's <= 8 lines, kRingDepth entries deep */
    const unsigned lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const size_t item   = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const size_t nitems = (size_t)rv.nob * m.hnrays * nchunks_f;
    stage_exp_table();
    if (item >= nitems) return; /* whole warps only */
    const uint32_t cf = (uint32_t)(item % nchunks_f);
    const size_t t    = item / nchunks_f;
    const uint32_t rr = (uint32_t)(t % m.hnrays);
    const uint32_t olr = (uint32_t)(t / m.hnrays);   /* origin within the record range */
    const uint32_t ol = rv.ola + olr;                /* origin within the path chunk */
    const uint32_t o  = pv.o0 + ol;
    const FreqChunk ch = chunks[cf];
    const bool active  = lane < ch.nf;
    const uint32_t f   = ch.f0 + (active ? lane : ch.nf - 1);
    const uint32_t l   = f / m.nquads;       // line of this frequency bin (f = l*nquads + z, model.cpp:127-146)
    const uint32_t z   = f - l * m.nquads;
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
    double* s0 = &s_side0[wib][0][lane]; /* side-0 results wait in shared memory while side 1 runs */
    s0[0] = 0.0; s0[32] = 0.0; s0[64] = 0.0; s0[96] = 0.0; s0[128] = 0.0; s0[160] = 0.0;

    for (int side = 0; side < 2; side++) {
        const uint32_t r   = side ? ar : rr;
        const uint32_t len = side ? len_f : len_b;
        if (len == 0) continue;
        const size_t list = ((size_t)olr * m.nrays + r) * m.nlines + l;
        const unsigned long long b = __ldg(&rv.off[list]);
        const uint32_t n = (uint32_t)(__ldg(&rv.off[list + 1]) - b); /* entries of this lane's line, the origin's included */
        nentries += n - 1;
        Lane3 st;
        st.haveX = false; st.X_is_end = false; st.large_far = false; st.nlarge = 0;
        st.chiX = st.SX = st.F = st.a = st.dtau_far = st.SatY_far = 0.0;
        const uint32_t nmax = __reduce_max_sync(0xffffffffu, n);
        /* the records of the warp's lines arrive through a shared-memory ring filled by cp.async, kRingDepth - 1 entries ahead: lanes
           (k, c) = (lane / 4, lane % 4) copy the c-th 16 bytes of the next record of line slot k (one instruction per warp and entry);
           the line integrals of the next entry are requested as soon as its record is readable, one entry ahead */
        const uint32_t ks = l - ch.l0;
        const uint32_t kc = lane >> 2, cc = lane & 3;
        const int jk = (int)((ch.l0 + kc) * m.nquads) - (int)ch.f0; /* first lane of line slot kc */
        const unsigned long long bk = __shfl_sync(0xffffffffu, b, jk > 0 ? jk : 0);
        const uint32_t nk = __shfl_sync(0xffffffffu, n, jk > 0 ? jk : 0);
        const bool copier = kc < ch.nl;
        __syncwarp();
        const double* src = rv.rec + bk * 8 + 2 * cc; /* this lane's 16 bytes of the next record to copy; stops at the list's last record */
        uint32_t left = copier ? nk - 1 : 0u;          /* records after *src */
        for (int d = 0; d < kRingDepth - 1; d++) {
            if (copier) cp_async16(&s_ring[wib][d][kc][2 * cc], src);
            cp_async_commit();
            if (left) { src += 8; left--; }
        }
        Trip tr; tr.dtau = tr.SX = tr.SY = 0.0;
        double shX = 0.0, dzX = 0.0;
#pragma unroll kSweepUnroll
        for (uint32_t e = 0; e < nmax; e++) {
            const bool part = e < n;
            if (copier) cp_async16(&s_ring[wib][(e + kRingDepth - 1) % kRingDepth][kc][2 * cc], src);
            cp_async_commit();
            if (left) { src += 8; left--; }
            cp_async_wait<kRingDepth - 2>(); /* entries e and e + 1 have landed */
            __syncwarp();
            const double* yr = &s_ring[wib][e % kRingDepth][ks][0];
            Rec4 y;
            y.shift = yr[0]; y.dz = yr[1]; y.iw = yr[2]; y.ie = yr[3]; y.io = yr[4]; y.sqw = yr[5];
            y.word = (unsigned long long)__double_as_longlong(yr[7]);
            const unsigned long long w1 = (unsigned long long)__double_as_longlong(s_ring[wib][(e + 1) % kRingDepth][ks][7]);
            Trip tn; tn.dtau = tn.SX = tn.SY = 0.0;
            if (w1 != kNoSlot && e + 1 < n) {
                const double* tp = trip + (w1 * m.nquads + z) * 3;
                tn.dtau = __ldg(tp); tn.SX = __ldg(tp + 1); tn.SY = __ldg(tp + 2);
            }
            // ---- the common step: an interior row whose window at Y holds the own line or no line ----
            const bool large = y.word != kNoSlot;
            const double frY = freq * y.shift;
            int cls;
            if (APPROX == APPROX_CLOSELINES) {
                const double ub = m.max_dist * (frY * y.sqw);
                cls = window_class(L, frY - ub, frY + ub);
            } else {
                cls = (APPROX == APPROX_ONELINE && L.own) ? WIN_OWN : WIN_OTHER;
            }
            double chiY, SY;
            bool weird;
            eta_chi_own<APPROX>(m, cls == WIN_OWN, y.iw, y.iw * kINV_SQRT_PI, y.ie, y.io, frY - L.s_own, chiY, SY, weird);
            const bool ok = st.haveX & !st.X_is_end & (cls != WIN_OTHER) & !weird;
            if (__all_sync(0xffffffffu, ok | !part)) {
                if (part) {
                    double dtau, Sx_large = 0.0, Sy = 0.0;
                    if (large) { st.nlarge++; dtau = tr.dtau; Sx_large = tr.SX; Sy = tr.SY; }
                    else dtau = 0.5 * (st.chiX + chiY) * dzX;
                    const double Sx      = side ? (st.large_far ? st.SatY_far : st.SX) : (large ? Sx_large : st.SX);
                    const double avg     = 0.5 * (st.dtau_far + dtau);
                    const double invFar  = avg * st.dtau_far;
                    const double invNear = avg * dtau;
                    const double P       = drcp_inl(invFar * (1.0 + st.F)); /* A * FI of the previous row */
                    st.a = (st.a * P + Sx) * invNear;
                    st.F = (st.F * P + 1.0) * invNear;
                    st.dtau_far = dtau; st.SatY_far = Sy; st.chiX = chiY; st.SX = SY;
                    st.large_far = large;
                }
            } else if (part) {
                st = step4_general<APPROX>(m, L, st, freq, side, rv.rec, dzX, y, tr, b + e, rv.pts, ll);
            }
            if (part) { dzX = y.dz; tr = tn; }
            __syncwarp(); /* the slot read here is overwritten by the next iteration's copy */
        }
        cp_async_wait<0>(); /* copies issued past the last entry must not land in the other side's ring */
        __syncwarp();
        SC = st.SX; /* source function at the origin */
        const double FI = drcp(1.0 + st.F); /* FI and Su of the row next to the origin, from the carried b = Su * (1 + F) */
        const double aa = st.a * FI;
        nlarge += st.nlarge;
        if (side == 0) {
            s0[0] = st.F; s0[32] = FI; s0[64] = aa; s0[96] = st.dtau_far; s0[128] = st.SatY_far; s0[160] = st.large_far ? 1.0 : 0.0;
        } else { F1 = st.F; FI1 = FI; a1 = aa; dt1 = st.dtau_far; lg1 = st.large_far; Sc1 = st.SatY_far; }
    }

    // ---- centre row; the Lambda increment's bin- and ray-independent factors are applied by k_reduce_lambda2 ----
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
