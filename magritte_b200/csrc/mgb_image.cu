// mgb_image.cu -- the new imager: Solver::image_feautrier_order_2_new_imager<approx> (reference solver.tpp:576-627).
//
//   host      Image::set_coordinates_projection_surface (image.cpp:221-368), Geometry::get_closest_bdy_point_in_custom_raydir
//             (geometry.tpp:346-364): a loop over the boundary points, done once per call on the host copies kept by mgb_set_geometry
//   k_image_trace<MODE>   one warp per pixel: pixel -> 3D origin (image.cpp:370-398), walk over boundary neighbours to the boundary
//             point closest to the ray (trace_ray_imaging_get_start, solver.tpp:808-825), Rest-frame imaging trace with the
//             all-lines sub-stepping of set_data (trace_ray_imaging :840-897, set_data :1003-1043, InterpHelper::get_n_interp).
//             MODE 2 (the usual case) counts AND writes the entries (nr, nr_interp, interp_factor, shift, dZ) in one pass, at
//             pixel * cap with cap taken from the previous image's longest walk; when a walk does not fit (or the budget is small)
//             MODE 0 counts and MODE 1 writes at the scanned offsets (two passes). 40 registers, 48 warps per SM: a pointer chase.
//   k_image_records       one thread per entry: everything the bins of a pixel share (interpolated line data of every line, line
//             source function, sqrtw_max / dshift_max of the entry's point), computed once instead of once per bin
//   k_image_solve_rec<APPROX, TAU>  lanes = bins (3 full warps per pixel for 100 bins, the 4 left-over bins of 8 pixels share one
//             warp): image_feautrier_order_2 (:2462-2574), i.e. the reference's forward elimination from the far end of the ray to
//             the start point, in the reference's own operation order (no re-ordering: Image.I is bit-identical to the CPU
//             reference). Launched per slice of pixels; a finished slice goes to pinned staging on a second stream while the next is
//             solved, host threads move it on into the caller's buffer.
//   k_image_solve<APPROX, TAU>      the first version of the solve (MGB_IMAGE_KERNEL=1): no records, every lane calls the generic
//             eta_chi_interp / S_dtau_line_integrated of mgb_device.cuh on (nr, nr_interp, factor). Kept as the in-tree cross-check
//             (tests/test_gpu_image.py::test_image_kernels_bit_identical).
// Frequencies of an image are not tied to the lines (compute_spectral_discretisation(nu_min, nu_max, n)), so eta/chi use the
// generic window search (lower / upper bound on the sorted line frequencies), exactly the reference's loops.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

#include <cub/device/device_reduce.cuh>
#include <cub/device/device_scan.cuh>

#include "mgb_ctx.h"

using namespace mgb;

namespace {

struct ImageGeom {
    double ray[3];    // normed direction
    double scp[3];    // Image::surface_center_point
    double center[3]; // Points::center
    uint32_t start_bdy;
};

struct ImagePaths { // entries of every pixel of a chunk, in emission order (array index centre-1-e of the reference's scratch)
    const unsigned long long* poff; // [npix] exclusive prefix of len
    uint32_t* nr;
    uint32_t* nri;
    double* fac;
    double* sh;
    double* dz;
    uint32_t cap; // != 0: the entries of pixel w start at w * cap (single-pass trace) instead of poff[w]
};
__device__ __forceinline__ unsigned long long entry_offset(const ImagePaths& ip, size_t w) {
    return ip.cap ? (unsigned long long)w * ip.cap : ip.poff[w];
}
// Records of a pixel chunk: what the solve kernel reads instead of (nr, nr_interp, interp_factor) and the line table. The list of pixel
// w starts at poff[w] + w: its start point first (position 0), then entry e at position e + 1.
struct ImageRecs {
    double4* hdr; // per position {shift, dZ, sqrtw_max[nr], dshift_max[nr]}
    double4* rec; // per (position, line slot) {inverse_width, emissivity, opacity, S_line = emissivity / opacity}, interpolated (nr, nr_interp, factor)
    // CloseLines: the slots are the sorted line indices [slo, slo + nsel) that some bin of this image can have in its window (bounds of
    // frequency x shift over all pixels, widened by the largest window); ApproximationType None sums every line: slo = 0, nsel = nlines,
    // natural line order
    uint32_t slo, nsel;
};

// image.cpp:370-398
__device__ __forceinline__ void surface_coords_to_3D(const ImageGeom& g, double x, double y, double& ox, double& oy, double& oz) {
    const double rx = g.ray[0], ry = g.ray[1], rz = g.ray[2];
    const double denominator = sqrt(rx * rx + ry * ry);
    const double inverse_denominator = 1.0 / denominator;
    const double ix = ry * inverse_denominator, iy = -rx * inverse_denominator;
    const double jx = rx * rz * inverse_denominator, jy = ry * rz * inverse_denominator, jz = -denominator;
    if (denominator >= 1.0e-9) {
        ox = (x * ix + y * jx) + g.scp[0]; oy = (x * iy + y * jy) + g.scp[1]; oz = (y * jz) + g.scp[2];
    } else {
        ox = x + g.scp[0]; oy = y + g.scp[1]; oz = 0.0 + g.scp[2];
    }
}

// Geometry::get_boundary_point_closer_to_custom_ray (geometry.tpp:310-338): lanes scan the neighbour list; the winner is the
// first boundary neighbour (list order) with the smallest distance strictly below the current point's.
__device__ __forceinline__ uint32_t warp_bdy_closer(const DevModel& m, const RayFrame& fr, uint32_t crt, unsigned lane) {
    const double4 Pc = ldg4(&m.pos[crt]);
    const double Rcx = Pc.x - fr.ox, Rcy = Pc.y - fr.oy, Rcz = Pc.z - fr.oz;
    const double Zc  = dot3(Rcx, Rcy, Rcz, fr.dx, fr.dy, fr.dz);
    const double d0  = dot3(Rcx, Rcy, Rcz, Rcx, Rcy, Rcz) - Zc * Zc;
    const uint2 c = __ldg(&m.csr[crt]);
    double best = d0;
    uint32_t besti = 0xffffffffu, bestn = crt;
    for (uint32_t i = lane; i < c.y; i += 32) {
        const uint32_t n = __ldg(&m.nbs[c.x + i]);
        if (__ldg(&m.p2b[n]) != m.npoints) {
            const double4 P = ldg4(&m.pos[n]);
            const double Rx = P.x - fr.ox, Ry = P.y - fr.oy, Rz = P.z - fr.oz;
            const double Zn = dot3(Rx, Ry, Rz, fr.dx, fr.dy, fr.dz);
            const double d2 = dot3(Rx, Ry, Rz, Rx, Ry, Rz) - Zn * Zn;
            if (d2 < best) { best = d2; besti = i; bestn = n; }
        }
    }
    double mn = best;
#pragma unroll
    for (int off = 16; off; off >>= 1) {
        const double o = __shfl_xor_sync(0xffffffffu, mn, off);
        mn = (o < mn) ? o : mn;
    }
    const uint32_t key = (besti != 0xffffffffu && best == mn) ? besti : 0xffffffffu;
    const uint32_t win = __reduce_min_sync(0xffffffffu, key);
    if (win == 0xffffffffu) return crt;
    return __shfl_sync(0xffffffffu, bestn, (int)(win & 31u));
}

// Geometry::get_distance_origin_to_boundary (geometry.tpp:167-191)
__device__ __forceinline__ double distance_origin_to_boundary(const DevModel& m, const RayFrame& fr, uint32_t bdy) {
    const double4 P = ldg4(&m.pos[bdy]);
    double Z = dot3(fr.dx, fr.dy, fr.dz, P.x - fr.ox, P.y - fr.oy, P.z - fr.oz);
    if (m.spherical) {
        const double Rcos  = dot3(fr.ox, fr.oy, fr.oz, fr.dx, fr.dy, fr.dz);
        const double R2    = dot3(fr.ox, fr.oy, fr.oz, fr.ox, fr.oy, fr.oz);
        const double Rsin2 = R2 - Rcos * Rcos;
        const double horz_dist2 = dot3(P.x, P.y, P.z, P.x, P.y, P.z) - Rsin2;
        if (horz_dist2 > 0) Z = -sqrt(horz_dist2) - Rcos; else Z = -Rcos;
    }
    return Z;
}

// get_shift<Rest>(origin, 0, raydir, c, Z, false) (geometry.tpp:218-221, 408-418, 452-465)
__device__ __forceinline__ double get_shift_rest(const DevModel& m, const RayFrame& fr, uint32_t c, double Z) {
    const double4 vc = ldg4(&m.vel[c]);
    if (m.spherical) {
        const double px = ldg4(&m.pos[c]).x;
        if (px == 0.0) return 1.0;
        const double Rcos_plus_Z = dot3(fr.ox, fr.oy, fr.oz, fr.dx, fr.dy, fr.dz) + Z;
        return 1.0 + vc.x * Rcos_plus_Z / px;
    }
    return 1.0 + dot3(vc.x, vc.y, vc.z, fr.dx, fr.dy, fr.dz);
}

// InterpHelper::get_n_interp (solver_interp_helper.tpp:10-46): the largest |logf(crit_next / crit_curr)| over ALL lines
// (lanes stride the lines; std::max_element keeps the largest value, NaNs are never selected unless first).
__device__ __forceinline__ uint32_t warp_n_interp_all(const DevModel& m, uint32_t curr, uint32_t next, unsigned lane) {
    float mx = -1.0f; /* |.| >= 0 */
    for (uint32_t l = lane; l < m.nlines; l += 32) {
        const double x = __ldg(&m.lrow[(size_t)next * m.nlines + l].criterion);
        const double y = __ldg(&m.lrow[(size_t)curr * m.nlines + l].criterion);
        const float d  = fabsf(mgb_logf((float)(x / y)));
        if (mx < d) mx = d;
    }
#pragma unroll
    for (int off = 16; off; off >>= 1) {
        const float o = __shfl_xor_sync(0xffffffffu, mx, off);
        mx = (mx < o) ? o : mx;
    }
    const uint32_t n = to_size(ceil((double)mx / m.log_max_interp_diff));
    return n > 1 ? n : 1;
}

// ---------------------------------------------------------------------------------------------------
// Tracer: one warp per pixel. reference: solver.tpp:591-607 (driver part), 808-825, 840-897, 1003-1043
// ---------------------------------------------------------------------------------------------------
// MODE 0 counts the entries, MODE 1 writes them at the pixel's scanned offset, MODE 2 does both in one pass: entries go to w * cap
// while they fit (the host falls back to MODE 0 + 1 when some walk is longer than cap).
#ifndef MGB_TRACE_CTAS
#define MGB_TRACE_CTAS 12
#endif
template <int MODE>
__global__ void __launch_bounds__(128, MGB_TRACE_CTAS) k_image_trace(DevModel m, ImageGeom g, unsigned long long pix0, uint32_t npix,
                                                     const double* __restrict__ ImX, const double* __restrict__ ImY,
                                                     uint32_t* __restrict__ len, uint32_t* __restrict__ start, double* __restrict__ shift0,
                                                     ImagePaths ip, unsigned long long* __restrict__ counters) {
    const unsigned lane = threadIdx.x & 31;
    const size_t w      = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= npix) return;
    constexpr bool FILL = MODE != 0;
    if (MODE == 1 && len[w] == 0) return;
    const uint32_t room = MODE == 2 ? ip.cap : 0xffffffffu;
    RayFrame fr;
    surface_coords_to_3D(g, ImX[pix0 + w], ImY[pix0 + w], fr.ox, fr.oy, fr.oz);
    fr.dx = g.ray[0]; fr.dy = g.ray[1]; fr.dz = g.ray[2];
    fr.vx = fr.vy = fr.vz = 0.0;
    // trace_ray_imaging_get_start
    uint32_t c = g.start_bdy;
    for (;;) {
        const uint32_t nx = warp_bdy_closer(m, fr, c, lane);
        if (nx == c) break;
        c = nx;
    }
    double Z = distance_origin_to_boundary(m, fr, c);
    const double sh0 = get_shift_rest(m, fr, c, Z);
    double sh_min = sh0, sh_max = sh0; /* bounds of the shifts along this pixel's ray (the sub-steps interpolate between them) */
    // trace_ray_imaging
    const unsigned long long off = FILL ? entry_offset(ip, w) : 0ull;
    uint32_t n = 0, ns;
    bool accessed_non_boundary = false;
    double dZ = 0.0;
    uint32_t crt = c;
    uint32_t nxt = warp_get_next(m, fr, c, Z, dZ, ns, lane);
    if (nxt < m.npoints) {
        double shift_crt = get_shift_rest(m, fr, crt, Z - dZ);
        double shift_nxt = get_shift_rest(m, fr, nxt, Z);
        sh_min = fmin(sh_min, shift_crt); sh_max = fmax(sh_max, shift_crt);
        for (;;) {
            sh_min = fmin(sh_min, shift_nxt); sh_max = fmax(sh_max, shift_nxt);
            // set_data (solver.tpp:1003-1043)
            const uint32_t ni = warp_n_interp_all(m, crt, nxt, lane);
            if (FILL) {
                if (ni > 1) {
                    const double dZ_interpl     = dZ / ni;
                    const double dshift_interpl = (shift_nxt - shift_crt) / ni;
                    for (uint32_t i = 1 + lane; i <= ni && n + (i - 1) < room; i += 32) {
                        const unsigned long long e = off + n + (i - 1);
                        ip.nr[e] = nxt; ip.nri[e] = crt;
                        ip.sh[e]  = shift_crt + i * dshift_interpl;
                        ip.fac[e] = 1.0 - ((double)i) / ni;
                        ip.dz[e]  = dZ_interpl;
                    }
                } else if (lane == 0 && n < room) {
                    const unsigned long long e = off + n;
                    ip.nr[e] = nxt; ip.nri[e] = crt; ip.sh[e] = shift_nxt; ip.fac[e] = 0.0; ip.dz[e] = dZ;
                }
            }
            n += ni;
            const double4 P = ldg4(&m.pos[nxt]);
            const bool interior = __ldg(&m.p2b[nxt]) == m.npoints;
            if (!(interior || dot3(P.x - g.center[0], P.y - g.center[1], P.z - g.center[2], fr.dx, fr.dy, fr.dz) < 0)) break;
            crt = nxt; shift_crt = shift_nxt;
            if (interior) accessed_non_boundary = true;
            nxt = warp_get_next(m, fr, crt, Z, dZ, ns, lane);
            if (!(nxt < m.npoints)) break;
            shift_nxt = get_shift_rest(m, fr, nxt, Z);
        }
    }
    if (MODE != 1 && lane == 0) {
        // a ray that only touched boundary points is dropped in general geometry (solver.tpp:892-896): n_tot = 1
        const uint32_t L = (accessed_non_boundary || m.spherical) ? n : 0u;
        len[w] = L; start[w] = c; shift0[w] = sh0;
        atomicAdd(&counters[0], (unsigned long long)L);
        atomicMax((unsigned int*)&counters[1], n); /* get_ray_lengths_max_new_imager counts every walk */
        /* positive doubles order like their bit patterns; a non-positive or NaN shift makes the host keep every line */
        if (sh_min > 0.0 && sh_max < CUDART_INF) {
            atomicMin(&counters[2], (unsigned long long)__double_as_longlong(sh_min));
            atomicMax(&counters[3], (unsigned long long)__double_as_longlong(sh_max));
        } else {
            atomicMin(&counters[2], 0ull);
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// Solve: one warp per (pixel, 32 frequency bins); lanes = bins. reference: solver.tpp:2139-2186 (compute_source_dtau),
// 2462-2574 (image_feautrier_order_2), 609-621 (Image.I)
// ---------------------------------------------------------------------------------------------------
template <int APPROX>
__device__ __forceinline__ void image_source_dtau(const DevModel& m, const Entry& c, const Entry& n, double curr_freq, double next_freq,
                                                  double dz, bool& compute_curr_opacity, double& dtau_n, double& chi_c, double& chi_n,
                                                  double& S_c, double& S_n) {
    const double dshift_abs = fabs(n.shift - c.shift);
    const double a = __ldg(&m.dshift_max[c.p1]), b = __ldg(&m.dshift_max[n.p1]);
    const double dshift_max = (b < a) ? b : a; /* std::min(a, b) */
    if (dshift_abs > dshift_max) {
        compute_curr_opacity = true;
        S_dtau_line_integrated<APPROX>(m, c, n, 0u, curr_freq, next_freq, dz, dtau_n, S_c, S_n);
    } else {
        if (compute_curr_opacity) {
            compute_curr_opacity = false;
            double eta_c;
            eta_chi_interp<APPROX>(m, c, 0u, curr_freq, eta_c, chi_c);
            S_c = ddiv(eta_c, chi_c);
        }
        double eta_n;
        eta_chi_interp<APPROX>(m, n, 0u, next_freq, eta_n, chi_n);
        S_n    = ddiv(eta_n, chi_n);
        dtau_n = 0.5 * (chi_c + chi_n) * dz;
    }
}

// TAU: Solver::image_optical_depth (solver.tpp:2718-2760) instead -- the same walk, Image.I = sum of the segments' dtau from the far end
// to the start point (image_optical_depth_new_imager, :688-733; 0 for a ray without interior points).
template <int APPROX, bool TAU>
__global__ void __launch_bounds__(128) k_image_solve(DevModel m, uint32_t npix, uint32_t nfreqs, uint32_t nchunks_f,
                                                     const double* __restrict__ nu, uint32_t nu_stride,
                                                     const uint32_t* __restrict__ len, const uint32_t* __restrict__ start,
                                                     const double* __restrict__ shift0, ImagePaths ip, double* __restrict__ I) {
    const unsigned lane = threadIdx.x & 31;
    const size_t item   = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    stage_exp_table();
    if (item >= (size_t)npix * nchunks_f) return;
    const uint32_t pix = (uint32_t)(item / nchunks_f), cf = (uint32_t)(item % nchunks_f);
    const uint32_t f0  = cf * 32 + lane;
    const bool active  = f0 < nfreqs;
    const uint32_t f   = active ? f0 : nfreqs - 1;
    const uint32_t c   = start[pix];
    const uint32_t L   = len[pix];
    const double freq  = __ldg(&nu[(size_t)c * nu_stride + f]);
    double out;
    if (L == 0) {
        out = TAU ? 0.0 : boundary_intensity(m, c, freq); /* n_tot = 1 (solver.tpp:616-620; :726-730) */
    } else if (TAU) {
        const unsigned long long off = entry_offset(ip, pix);
        Entry cur, nxt;
        cur.p1 = ip.nr[off + L - 1]; cur.p2 = ip.nri[off + L - 1]; cur.factor = ip.fac[off + L - 1]; cur.shift = ip.sh[off + L - 1];
        bool compute_curr_opacity = true;
        double chi_c = 0.0, chi_n = 0.0, term_c = 0.0, term_n = 0.0, dtau_n = 0.0, tau = 0.0;
        for (uint32_t e = L; e-- > 0;) {
            if (e > 0) {
                nxt.p1 = ip.nr[off + e - 1]; nxt.p2 = ip.nri[off + e - 1]; nxt.factor = ip.fac[off + e - 1]; nxt.shift = ip.sh[off + e - 1];
            } else {
                nxt.p1 = c; nxt.p2 = 0u; nxt.factor = 0.0; nxt.shift = shift0[pix];
            }
            if (e != L - 1) { term_c = term_n; chi_c = chi_n; }
            image_source_dtau<APPROX>(m, cur, nxt, freq * cur.shift, freq * nxt.shift, ip.dz[off + e], compute_curr_opacity, dtau_n, chi_c, chi_n,
                                      term_c, term_n);
            tau += dtau_n;
            cur = nxt;
        }
        out = tau;
    } else {
        const unsigned long long off = entry_offset(ip, pix);
        Entry cur, nxt;
        cur.p1 = ip.nr[off + L - 1]; cur.p2 = ip.nri[off + L - 1]; cur.factor = ip.fac[off + L - 1]; cur.shift = ip.sh[off + L - 1];
        bool compute_curr_opacity = true;
        double chi_c = 0.0, chi_n = 0.0, term_c = 0.0, term_n = 0.0, dtau_n = 0.0, dtau_c;
        double Su = 0.0, FF = 0.0, FI = 0.0;
        for (uint32_t e = L; e-- > 0;) {
            if (e > 0) {
                nxt.p1 = ip.nr[off + e - 1]; nxt.p2 = ip.nri[off + e - 1]; nxt.factor = ip.fac[off + e - 1]; nxt.shift = ip.sh[off + e - 1];
            } else { /* the start point: nr_[centre], shift_[centre]; nr_interp_/interp_factor_ stay value-initialised (0) */
                nxt.p1 = c; nxt.p2 = 0u; nxt.factor = 0.0; nxt.shift = shift0[pix];
            }
            const double dz = ip.dz[off + e];
            if (e == L - 1) {
                image_source_dtau<APPROX>(m, cur, nxt, freq * cur.shift, freq * nxt.shift, dz, compute_curr_opacity, dtau_n, chi_c, chi_n,
                                          term_c, term_n);
                const double inverse_dtau_f = drcp(dtau_n);
                const double Cf        = 2.0 * inverse_dtau_f * inverse_dtau_f;
                const double Bf_min_Cf = 1.0 + 2.0 * inverse_dtau_f;
                const double Bf        = Bf_min_Cf + Cf;
                const double I_bdy_f   = boundary_intensity(m, cur.p1, freq * cur.shift);
                Su = term_c + 2.0 * I_bdy_f * inverse_dtau_f;
                Su = ddiv(Su, Bf);
                FF = 0.5 * Bf_min_Cf * dtau_n * dtau_n;
                FI = drcp(1.0 + FF);
            } else {
                term_c = term_n; dtau_c = dtau_n; chi_c = chi_n;
                image_source_dtau<APPROX>(m, cur, nxt, freq * cur.shift, freq * nxt.shift, dz, compute_curr_opacity, dtau_n, chi_c, chi_n,
                                          term_c, term_n);
                const double dtau_avg  = 0.5 * (dtau_c + dtau_n);
                const double inverse_A = dtau_avg * dtau_c;
                const double inverse_C = dtau_avg * dtau_n;
                const double A         = drcp(inverse_A);
                const double FFn       = (A * FF * FI + 1.0) * inverse_C;
                const double FIn       = drcp(1.0 + FFn);
                Su = (A * Su + term_c) * FIn * inverse_C;
                FF = FFn; FI = FIn;
            }
            cur = nxt;
        }
        const double inverse_dtau_l = drcp(dtau_n);
        const double Al          = 2.0 * inverse_dtau_l * inverse_dtau_l;
        const double Bl_min_Al   = 1.0 + 2.0 * inverse_dtau_l;
        const double Bl          = Bl_min_Al + Al;
        const double denominator = drcp(Bl * FF + Bl_min_Al);
        const double I_bdy_l     = boundary_intensity(m, cur.p1, freq * cur.shift);
        double Sl = term_n + 2.0 * I_bdy_l * inverse_dtau_l;
        Sl = (Al * Su + Sl) * (1.0 + FF) * denominator;
        out = 2.0 * Sl - boundary_intensity(m, cur.p1, freq);
    }
    if (active) I[(size_t)pix * nfreqs + f] = out;
}

// ---------------------------------------------------------------------------------------------------
// Records: what every frequency bin of a pixel needs from an entry does not depend on the bin -- the interpolated line data of
// get_eta_and_chi_interpolated (solver.tpp:1259-1302: interp_linear / interp_log of the two points' inverse width, emissivity,
// opacity), the line source function of compute_dtau_single_line (:2374-2455), sqrtw_max and dshift_max of the entry's point. The
// generic kernel above recomputes them in each of the nfreqs lanes (two powf and a division per line and bin); here one thread per
// entry computes them once, with the same functions in the same order, and k_image_solve_rec reads them back (warp-uniform loads).
// ---------------------------------------------------------------------------------------------------
template <bool CLOSE>
__global__ void __launch_bounds__(128) k_image_records(DevModel m, uint32_t npix, const uint32_t* __restrict__ len,
                                                       const uint32_t* __restrict__ start, const double* __restrict__ shift0, ImagePaths ip,
                                                       ImageRecs ir) {
    const unsigned lane = threadIdx.x & 31;
    const size_t w      = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    stage_exp_table(); /* interp_log's powf reads it */
    if (w >= npix) return;
    const uint32_t L = len[w];
    if (L == 0) return;
    const unsigned long long off = entry_offset(ip, w), roff = ip.poff[w] + w;
    for (uint32_t j = lane; j <= L; j += 32) {
        Entry e;
        double dz = 0.0;
        if (j == 0) { /* the start point: nr_[centre], shift_[centre]; nr_interp_ / interp_factor_ value-initialised */
            e.p1 = start[w]; e.p2 = 0u; e.factor = 0.0; e.shift = shift0[w];
        } else {
            const unsigned long long i = off + j - 1;
            e.p1 = ip.nr[i]; e.p2 = ip.nri[i]; e.factor = ip.fac[i]; e.shift = ip.sh[i]; dz = ip.dz[i];
        }
        ir.hdr[roff + j] = make_double4(e.shift, dz, __ldg(&m.sqrtw_max[e.p1]), __ldg(&m.dshift_max[e.p1]));
        double4* R = ir.rec + (roff + j) * ir.nsel;
        for (uint32_t k = 0; k < ir.nsel; k++) {
            const uint32_t l = CLOSE ? __ldg(&m.sorted_map[ir.slo + k]) : k;
            const LineRow a = load_row(m, e.p1, l);
            double iw = a.inverse_width, ie = a.emissivity, io = a.opacity, Sl = a.emissivity / a.opacity;
            if (e.factor != 0.0) {
                const LineRow b = load_row(m, e.p2, l);
                iw = interp_linear(a.inverse_width, b.inverse_width, e.factor);
                ie = interp_log(a.emissivity, b.emissivity, e.factor);
                io = interp_log(a.opacity, b.opacity, e.factor);
                Sl = interp_log(Sl, b.emissivity / b.opacity, e.factor);
            }
            R[k] = make_double4(iw, ie, io, Sl);
        }
    }
}

// get_eta_and_chi_interpolated<approx> (solver.tpp:1178-1202, 1259-1302) on the records of one position
template <int APPROX>
__device__ __forceinline__ void eta_chi_rec(const DevModel& m, const double4* __restrict__ R, uint32_t slo, uint32_t nsel, double sqw, double freq, double& eta, double& chi) {
    eta = 0.0;
    chi = m.min_opacity;
    if (APPROX == APPROX_NONE) {
        for (uint32_t l = 0; l < m.nlines; l++) {
            const double4 v   = ldg4(&R[l]);
            const double prof = gaussian(v.x, freq - __ldg(&m.line[l]));
            eta += prof * v.y; chi += prof * v.z;
        }
        return;
    }
    const double ub   = m.max_dist * (freq * sqw);
    const double left = freq - ub, right = freq + ub;
    /* every line a bin of this image can reach is in [slo, slo + nsel): the search over that range finds the full search's bounds */
    const uint32_t lo = slo + lower_bound_d(m.sorted_line + slo, nsel, left);
    const uint32_t hi = slo + upper_bound_d(m.sorted_line + slo, nsel, right);
    for (uint32_t sl = lo; sl < hi; sl++) {
        const double4 v   = ldg4(&R[sl - slo]); /* the records are in sorted-line order from slo on */
        const double prof = gaussian(v.x, freq - __ldg(&m.sorted_line[sl]));
        eta += prof * v.y; chi += prof * v.z;
    }
}

// compute_dtau_single_line (solver.tpp:2374-2455) on the records of the two positions
__device__ __forceinline__ double dtau_single_line_rec(const DevModel& m, const double4 vc, const double4 vn, double linefreq, double curr_freq,
                                                       double next_freq, double dz, double& Sc, double& Sn) {
    Sc = vc.w; Sn = vn.w;
    const double avg_iw   = (vc.x + vn.x) / 2.0;
    const double curr_op = vc.z, next_op = vn.z;
    if (curr_freq == next_freq) {
        const double prof   = gaussian(avg_iw, curr_freq - linefreq);
        const double avg_op = (curr_op + next_op) / 2.0;
        return dz * (prof * avg_op + m.min_opacity);
    }
    const double next_pos = (linefreq - next_freq) * avg_iw;
    const double curr_pos = (linefreq - curr_freq) * avg_iw;
    const double diff_pos = next_pos - curr_pos;
    const double avg_op   = (next_op + curr_op) / 2.0;
    const float erfdiff   = erff_ool((float)next_pos) - erff_ool((float)curr_pos); /* float subtraction */
    const double erfterm  = avg_op / diff_pos / 2.0 * (double)erfdiff;
    return dz * (avg_iw * erfterm + m.min_opacity);
}

// compute_S_dtau_line_integrated<approx> (solver.tpp:1940-1983, 2012-2111) on the records of the two positions
template <int APPROX>
__device__ __forceinline__ void S_dtau_rec(const DevModel& m, const double4* __restrict__ Rc, const double4* __restrict__ Rn, uint32_t slo, uint32_t nsel, double sqw_c,
                                           double sqw_n, double currfreq, double nextfreq, double dz, double& dtau, double& Scurr,
                                           double& Snext) {
    double sum_dtau = 0.0, sum_c = 0.0, sum_n = 0.0;
    if (APPROX == APPROX_NONE) {
        for (uint32_t l = 0; l < m.nlines; l++) {
            double Sc, Sn;
            const double ld = dtau_single_line_rec(m, ldg4(&Rc[l]), ldg4(&Rn[l]), __ldg(&m.line[l]), currfreq, nextfreq, dz, Sc, Sn);
            sum_dtau += ld; sum_c += ld * Sc; sum_n += ld * Sn;
        }
        dtau = sum_dtau; Scurr = sum_c / sum_dtau; Snext = sum_n / sum_dtau;
        return;
    }
    double left_freq, right_freq;
    if (currfreq < nextfreq) { left_freq = currfreq; right_freq = nextfreq; }
    else { right_freq = currfreq; left_freq = nextfreq; }
    const double cb = m.max_dist * (right_freq * sqw_c);
    const double nb = m.max_dist * (right_freq * sqw_n);
    const double ub = (cb < nb) ? nb : cb;
    const double left = left_freq - ub, right = right_freq + ub;
    const uint32_t lo = slo + lower_bound_d(m.sorted_line + slo, nsel, left);
    const uint32_t hi = slo + upper_bound_d(m.sorted_line + slo, nsel, right);
    for (uint32_t sl = lo; sl < hi; sl++) {
        double Sc, Sn;
        const double ld = dtau_single_line_rec(m, ldg4(&Rc[sl - slo]), ldg4(&Rn[sl - slo]), __ldg(&m.sorted_line[sl]), currfreq, nextfreq, dz, Sc, Sn);
        sum_dtau += ld; sum_c += ld * Sc; sum_n += ld * Sn;
    }
    double d = sum_dtau;
    const double bound_min_dtau = m.min_opacity * dz;
    if (-bound_min_dtau < d) d = (bound_min_dtau < d) ? d : bound_min_dtau;
    dtau = d; Scurr = sum_c / d; Snext = sum_n / d;
}

// compute_source_dtau (solver.tpp:2139-2186) between the positions with headers hc (current) and hn (next)
template <int APPROX>
__device__ __forceinline__ void image_source_dtau_rec(const DevModel& m, const double4 hc, const double4 hn, const double4* __restrict__ Rc,
                                                      const double4* __restrict__ Rn, uint32_t slo, uint32_t nsel, double freq, bool& compute_curr_opacity, double& dtau_n,
                                                      double& chi_c, double& chi_n, double& S_c, double& S_n) {
    const double dshift_abs = fabs(hn.x - hc.x);
    const double dshift_max = (hn.w < hc.w) ? hn.w : hc.w; /* std::min(curr, next) */
    const double curr_freq = freq * hc.x, next_freq = freq * hn.x, dz = hc.y;
    if (dshift_abs > dshift_max) {
        compute_curr_opacity = true;
        S_dtau_rec<APPROX>(m, Rc, Rn, slo, nsel, hc.z, hn.z, curr_freq, next_freq, dz, dtau_n, S_c, S_n);
    } else {
        if (compute_curr_opacity) {
            compute_curr_opacity = false;
            double eta_c;
            eta_chi_rec<APPROX>(m, Rc, slo, nsel, hc.z, curr_freq, eta_c, chi_c);
            S_c = ddiv(eta_c, chi_c);
        }
        double eta_n;
        eta_chi_rec<APPROX>(m, Rn, slo, nsel, hn.z, next_freq, eta_n, chi_n);
        S_n    = ddiv(eta_n, chi_n);
        dtau_n = 0.5 * (chi_c + chi_n) * dz;
    }
}

// k_image_solve on the records: the same walk from the far end (position L) to the start point (position 0), the same operations.
// Lanes = bins. A pixel takes nfreqs / 32 full warps; the nfreqs % 32 bins left over are packed with those of the next pixels into
// one warp (100 bins: 3 full warps per pixel + one warp per 8 pixels instead of 4 warps per pixel, a fifth of them at 4 of 32 lanes).
// Nothing in the walk is warp-collective, so lanes on different pixels simply run lists of different length.
struct ImageItems {
    uint32_t nfull, tail, group; // full warps per pixel, left-over bins, pixels per tail warp (0 without a tail)
    __host__ __device__ size_t count(uint32_t npix) const { return (size_t)npix * nfull + (group ? ((size_t)npix + group - 1) / group : 0); }
};
#ifndef MGB_IMAGE_CTAS
#define MGB_IMAGE_CTAS 8
#endif
#ifndef MGB_IMAGE_PREFETCH
#define MGB_IMAGE_PREFETCH 1
#endif
template <int APPROX, bool TAU>
__global__ void __launch_bounds__(128, MGB_IMAGE_CTAS) k_image_solve_rec(DevModel m, uint32_t pix0, uint32_t npix, uint32_t nfreqs, ImageItems it,
                                                                         const double* __restrict__ nu, uint32_t nu_stride,
                                                                         const uint32_t* __restrict__ len, const uint32_t* __restrict__ start,
                                                                         ImagePaths ip, const double4* __restrict__ hdr,
                                                                         const double4* __restrict__ rec, uint32_t slo, uint32_t nsel,
                                                                         double* __restrict__ I) {
    const unsigned lane = threadIdx.x & 31;
    const size_t item   = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    stage_exp_table();
    if (item >= it.count(npix)) return;
    const size_t nfull_items = (size_t)npix * it.nfull;
    size_t pix_l;
    uint32_t f;
    bool active = true;
    if (item < nfull_items) {
        pix_l = item / it.nfull;
        f     = (uint32_t)(item % it.nfull) * 32 + lane;
    } else {
        pix_l  = (item - nfull_items) * it.group + lane / it.tail;
        f      = it.nfull * 32 + lane % it.tail;
        active = lane < it.group * it.tail && pix_l < npix;
        if (!active) { pix_l = (item - nfull_items) * it.group; f = nfreqs - 1; } /* idle lanes shadow the warp's first pixel */
    }
    const uint32_t pix = pix0 + (uint32_t)pix_l; /* this launch images the pixels [pix0, pix0 + npix) of the chunk */
    const uint32_t c   = start[pix];
    const uint32_t L   = len[pix];
    const double freq  = __ldg(&nu[(size_t)c * nu_stride + f]);
    double out;
    if (L == 0) {
        out = TAU ? 0.0 : boundary_intensity(m, c, freq); /* n_tot = 1 (solver.tpp:616-620; :726-730) */
    } else {
        const unsigned long long roff = ip.poff[pix] + pix;
        const uint32_t far_point = ip.nr[entry_offset(ip, pix) + L - 1];
        const double4* Rc = rec + (roff + L) * nsel;
        double4 hc = ldg4(&hdr[roff + L]), hn = ldg4(&hdr[roff + L - 1]);
        bool compute_curr_opacity = true;
        double chi_c = 0.0, chi_n = 0.0, term_c = 0.0, term_n = 0.0, dtau_n = 0.0, dtau_c, tau = 0.0;
        double Su = 0.0, FF = 0.0, FI = 0.0;
        for (uint32_t e = L; e-- > 0;) { /* current = position e + 1, next = position e */
            const double4* Rn = Rc - nsel;
#if MGB_IMAGE_PREFETCH
            const double4 hnn = ldg4(&hdr[roff + (e ? e - 1 : 0)]); /* the header after next, one step ahead of its use */
#endif
            if (TAU) {
                if (e != L - 1) { term_c = term_n; chi_c = chi_n; }
                image_source_dtau_rec<APPROX>(m, hc, hn, Rc, Rn, slo, nsel, freq, compute_curr_opacity, dtau_n, chi_c, chi_n, term_c, term_n);
                tau += dtau_n;
            } else if (e == L - 1) {
                image_source_dtau_rec<APPROX>(m, hc, hn, Rc, Rn, slo, nsel, freq, compute_curr_opacity, dtau_n, chi_c, chi_n, term_c, term_n);
                const double inverse_dtau_f = drcp(dtau_n);
                const double Cf        = 2.0 * inverse_dtau_f * inverse_dtau_f;
                const double Bf_min_Cf = 1.0 + 2.0 * inverse_dtau_f;
                const double Bf        = Bf_min_Cf + Cf;
                const double I_bdy_f   = boundary_intensity(m, far_point, freq * hc.x);
                Su = term_c + 2.0 * I_bdy_f * inverse_dtau_f;
                Su = ddiv(Su, Bf);
                FF = 0.5 * Bf_min_Cf * dtau_n * dtau_n;
                FI = drcp(1.0 + FF);
            } else {
                term_c = term_n; dtau_c = dtau_n; chi_c = chi_n;
                image_source_dtau_rec<APPROX>(m, hc, hn, Rc, Rn, slo, nsel, freq, compute_curr_opacity, dtau_n, chi_c, chi_n, term_c, term_n);
                const double dtau_avg  = 0.5 * (dtau_c + dtau_n);
                const double inverse_A = dtau_avg * dtau_c;
                const double inverse_C = dtau_avg * dtau_n;
                const double A         = drcp(inverse_A);
                const double FFn       = (A * FF * FI + 1.0) * inverse_C;
                const double FIn       = drcp(1.0 + FFn);
                Su = (A * Su + term_c) * FIn * inverse_C;
                FF = FFn; FI = FIn;
            }
#if MGB_IMAGE_PREFETCH
            hc = hn; hn = hnn; Rc = Rn;
#else
            hc = hn; Rc = Rn;
            if (e) hn = ldg4(&hdr[roff + e - 1]);
#endif
        }
        if (TAU) {
            out = tau;
        } else { /* hc is the start point's header now */
            const double inverse_dtau_l = drcp(dtau_n);
            const double Al          = 2.0 * inverse_dtau_l * inverse_dtau_l;
            const double Bl_min_Al   = 1.0 + 2.0 * inverse_dtau_l;
            const double Bl          = Bl_min_Al + Al;
            const double denominator = drcp(Bl * FF + Bl_min_Al);
            const double I_bdy_l     = boundary_intensity(m, c, freq * hc.x);
            double Sl = term_n + 2.0 * I_bdy_l * inverse_dtau_l;
            Sl = (Al * Su + Sl) * (1.0 + FF) * denominator;
            out = 2.0 * Sl - boundary_intensity(m, c, freq);
        }
    }
    if (active) I[(size_t)pix * nfreqs + f] = out;
}

} // namespace

namespace {
struct Stopwatch {
    cudaEvent_t a = nullptr, b = nullptr;
    cudaStream_t s;
    explicit Stopwatch(cudaStream_t st) : s(st) { cudaEventCreate(&a); cudaEventCreate(&b); }
    ~Stopwatch() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
    void start() { cudaEventRecord(a, s); }
    float stop() { cudaEventRecord(b, s); cudaEventSynchronize(b); float ms = 0.f; cudaEventElapsedTime(&ms, a, b); return ms; }
};

// Image::set_coordinates_projection_surface (image.cpp:221-368) on the host copies of the boundary points.
void projection_surface(const mgb_ctx* ctx, int dimension, ImageGeom& g, uint32_t nx, uint32_t ny, std::vector<double>& ImX,
                        std::vector<double>& ImY) {
    const uint32_t nb = (uint32_t)ctx->h_b2p.size();
    const double* bp  = ctx->h_bpos.data();
    double max_x = -DBL_MAX, max_y = -DBL_MAX, min_x = DBL_MAX, min_y = DBL_MAX;
    if (dimension == 1) {
        for (uint32_t b = 0; b < nb; b++) { max_x = std::max(max_x, bp[3 * b]); min_x = std::min(min_x, bp[3 * b]); }
        max_x = std::max(std::abs(min_x), max_x);
        min_x = -max_x; min_y = min_x; max_y = max_x;
    }
    if (dimension == 3) {
        const double rx = g.ray[0], ry = g.ray[1], rz = g.ray[2];
        const double denominator = sqrt(rx * rx + ry * ry);
        const double inverse_denominator = 1.0 / denominator;
        const double ix = ry * inverse_denominator, iy = -rx * inverse_denominator;
        const double jx = rx * rz * inverse_denominator, jy = ry * rz * inverse_denominator, jz = -denominator;
        for (uint32_t b = 0; b < nb; b++) {
            const double* P = bp + 3 * (size_t)b;
            double X, Y;
            if (denominator >= 1.0e-9) { X = ix * P[0] + iy * P[1]; Y = jx * P[0] + jy * P[1] + jz * P[2]; }
            else { X = P[0]; Y = P[1]; }
            min_x = std::min(min_x, X); max_x = std::max(max_x, X); min_y = std::min(min_y, Y); max_y = std::max(max_y, Y);
        }
    }
    // Geometry::get_closest_bdy_point_in_custom_raydir (geometry.tpp:346-364). The reference's first guess is
    // boundary.point2boundary[0] used as a POINT index (sic); when point 0 is interior that is npoints and the reference reads one
    // element past the position array: such a first guess is dropped here. Boundary index 0 is never examined, as in the reference.
    const DevModel& m = ctx->m;
    double cp[3] = {0, 0, 0};
    if (m.spherical) {
        g.start_bdy = m.npoints - 1;
        for (int k = 0; k < 3; k++) cp[k] = ctx->h_pos_last[k];
    } else {
        uint32_t closest = ctx->h_p2b0;
        double dmin = HUGE_VAL;
        if (closest < m.npoints) {
            for (int k = 0; k < 3; k++) cp[k] = ctx->h_pos_p2b0[k];
            dmin = g.ray[0] * cp[0] + g.ray[1] * cp[1] + g.ray[2] * cp[2];
        }
        for (uint32_t b = 1; b < nb; b++) {
            const double* P = bp + 3 * (size_t)b;
            const double d  = g.ray[0] * P[0] + g.ray[1] * P[1] + g.ray[2] * P[2];
            if (d < dmin) { closest = ctx->h_b2p[b]; dmin = d; for (int k = 0; k < 3; k++) cp[k] = P[k]; }
        }
        g.start_bdy = closest;
    }
    double distance_from_origin = cp[0] * g.ray[0] + cp[1] * g.ray[1] + cp[2] * g.ray[2];
    if (m.spherical) distance_from_origin = -cp[0];
    for (int k = 0; k < 3; k++) g.scp[k] = g.ray[k] * distance_from_origin;
    const double deltax = (max_x - min_x) / (nx), deltay = (max_y - min_y) / (ny);
    ImX.resize((size_t)nx * ny); ImY.resize((size_t)nx * ny);
    for (uint32_t xi = 0; xi < nx; xi++) {
        const double xloc = min_x + deltax / 2.0 + xi * deltax;
        for (uint32_t yi = 0; yi < ny; yi++) {
            ImX[yi + (size_t)ny * xi] = xloc;
            ImY[yi + (size_t)ny * xi] = min_y + deltay / 2.0 + yi * deltay;
        }
    }
}
} // namespace

constexpr size_t kStageMaxBytes = 1ull << 30; // pinned staging of the image result: chunks above this take a plain cudaMemcpy
#define MGB_TRY_(x) do { int rc_ = (x); if (rc_ != MGB_OK) return rc_; } while (0)

static int image_impl(mgb_ctx* ctx, bool optical_depth, int approx, int dimension, const double* ray_dir, uint32_t nx, uint32_t ny,
                      uint32_t nfreqs, const double* nu, int nu_per_point, uint64_t pix_begin, uint64_t pix_end, double* ImX_out,
                      double* ImY_out, double* I_out, mgb_image_stats* stats) {
    if (!ctx) return MGB_EINVAL;
    if (!(ctx->have_geometry && ctx->have_thermo && ctx->have_lines && ctx->have_state))
        return ctx->fail(MGB_ESTATE, "mgb_image_feautrier_order_2: model incomplete (geometry %d thermo %d lines %d state %d)",
                         ctx->have_geometry, ctx->have_thermo, ctx->have_lines, ctx->have_state);
    if (!ray_dir || !nu || !I_out || !nx || !ny || !nfreqs) return ctx->fail(MGB_EINVAL, "mgb_image_feautrier_order_2: null or zero argument");
    const uint64_t npix_all = (uint64_t)nx * ny;
    if (pix_begin >= pix_end || pix_end > npix_all) return ctx->fail(MGB_EINVAL, "pixel range [%llu,%llu) invalid for %llu pixels",
                                                                     (unsigned long long)pix_begin, (unsigned long long)pix_end, (unsigned long long)npix_all);
    if (pix_end - pix_begin > 0xffffffffull) return ctx->fail(MGB_EINVAL, "more than 2^32 pixels in one call");
    // Model::compute_image_new (model.cpp:1008-1035)
    const double n2 = ray_dir[0] * ray_dir[0] + ray_dir[1] * ray_dir[1] + ray_dir[2] * ray_dir[2];
    if (n2 == 0.0) return ctx->fail(MGB_EINVAL, "The given ray direction vector does not point in a direction. Please use a non-zero (normed) "
                                                "direction vector to generate an image.");
    if (approx == APPROX_ONELINE)
        return ctx->fail(MGB_EUNSUPPORTED, optical_depth ? "One line approximation is currently not supported for imaging."
                                                         : "One line approximation is not supported for imaging.");
    if (approx != APPROX_NONE && approx != APPROX_CLOSELINES) return ctx->fail(MGB_EINVAL, "approx must be 0 (None) or 2 (CloseLines)");
    if (dimension != 1 && dimension != 3) return ctx->fail(MGB_EINVAL, "dimension must be 1 or 3 (Image::set_coordinates_projection_surface)");
    ImageGeom g;
    const double inv_norm = 1 / std::sqrt(n2);
    for (int k = 0; k < 3; k++) g.ray[k] = ray_dir[k] * inv_norm;
    if (dimension == 1 && (g.ray[0] != 0.0 || g.ray[1] != 1.0 || g.ray[2] != 0.0))
        return ctx->fail(MGB_EINVAL, "In 1D, the image ray has to be (0,1,0)"); /* image.cpp:80-86 */
    for (int k = 0; k < 3; k++) g.center[k] = ctx->h_center[k];
    cudaSetDevice(ctx->device);
    const DevModel& m = ctx->m;
    std::vector<double> hX, hY;
    projection_surface(ctx, dimension, g, nx, ny, hX, hY);
    if (ImX_out) memcpy(ImX_out, hX.data(), 8 * hX.size());
    if (ImY_out) memcpy(ImY_out, hY.data(), 8 * hY.size());

    mgb_image_stats st;
    memset(&st, 0, sizeof st);
    Stopwatch total(ctx->stream);
    total.start();
    MGB_TRY_(mgb_internal_refresh_points(ctx)); /* Model::set_dshift_max (setup_new_imager, solver.tpp:26) and the CloseLines window width */
    const uint32_t npix = (uint32_t)(pix_end - pix_begin);
    const size_t nu_elems = nu_per_point ? (size_t)m.npoints * nfreqs : (size_t)nfreqs;
    MGB_CUDA(ctx, ctx->im_X.ensure(8 * hX.size())); MGB_CUDA(ctx, ctx->im_Y.ensure(8 * hY.size()));
    MGB_CUDA(ctx, ctx->im_nu.ensure(8 * nu_elems));
    MGB_CUDA(ctx, cudaMemcpyAsync(ctx->im_X.p, hX.data(), 8 * hX.size(), cudaMemcpyHostToDevice, ctx->stream));
    MGB_CUDA(ctx, cudaMemcpyAsync(ctx->im_Y.p, hY.data(), 8 * hY.size(), cudaMemcpyHostToDevice, ctx->stream));
    MGB_CUDA(ctx, cudaMemcpyAsync(ctx->im_nu.p, nu, 8 * nu_elems, cudaMemcpyHostToDevice, ctx->stream));
    MGB_CUDA(ctx, ctx->im_len.ensure(4 * (size_t)npix)); MGB_CUDA(ctx, ctx->im_start.ensure(4 * (size_t)npix));
    MGB_CUDA(ctx, ctx->im_shift0.ensure(8 * (size_t)npix)); MGB_CUDA(ctx, ctx->im_poff.ensure(8 * (size_t)npix));
    MGB_CUDA(ctx, ctx->counters.ensure(64));
    MGB_CUDA(ctx, cudaMemsetAsync(ctx->counters.p, 0, 64, ctx->stream));
    MGB_CUDA(ctx, cudaMemsetAsync(ctx->counters.as<unsigned long long>() + 2, 0xff, 8, ctx->stream)); /* the smallest shift */
    {   /* the largest sqrtw_max of the model (the window of a bin is at most max_dist * freq * that) -> counters[4] */
        size_t tmp = 0;
        cub::DeviceReduce::Max(nullptr, tmp, m.sqrtw_max, ctx->counters.as<double>() + 4, (int64_t)m.npoints, ctx->stream);
        MGB_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
        MGB_CUDA(ctx, cub::DeviceReduce::Max(ctx->cub_tmp.p, tmp, m.sqrtw_max, ctx->counters.as<double>() + 4, (int64_t)m.npoints, ctx->stream));
    }
    // ---- pixel chunks: bounded entry storage (32 B per entry, + 32 + 32 * selected lines with records) and bounded result rows ----
    const char* env_kernel = getenv("MGB_IMAGE_KERNEL"); /* 1: the generic solve kernel, no records, two trace passes */
    const bool generic = env_kernel && env_kernel[0] == '1';
    unsigned long long entry_budget = std::max<unsigned long long>(1ull << 20, ctx->params.path_cache_bytes / (generic ? 32ull : 128ull));
    const uint32_t pix_budget = (uint32_t)std::max<size_t>(1, ctx->params.scratch_bytes / (8 * (size_t)nfreqs));
    const uint32_t nchunks_f  = (nfreqs + 31) / 32;
    const unsigned trace_blocks = (unsigned)(((size_t)npix * 32 + 127) / 128);
    // per-entry buffers grow with a quarter of headroom: the next direction's rays are a little longer or shorter, and a cudaFree +
    // cudaMalloc of GBs on a full GPU took 0.3 s
    auto grow = [](mgb::DevBuf& b, size_t n) -> cudaError_t { return (b.p && n <= b.bytes) ? cudaSuccess : b.ensure(n + n / 4); };
    auto ensure_entries = [&](size_t ne) -> cudaError_t {
        cudaError_t e;
        if ((e = grow(ctx->im_nr, 4 * ne)) != cudaSuccess || (e = grow(ctx->im_nri, 4 * ne)) != cudaSuccess ||
            (e = grow(ctx->im_fac, 8 * ne)) != cudaSuccess || (e = grow(ctx->im_sh, 8 * ne)) != cudaSuccess ||
            (e = grow(ctx->im_dz, 8 * ne)) != cudaSuccess) return e;
        return cudaSuccess;
    };
    // ---- first pass over the whole pixel range: the entry counts -- and, when a per-pixel capacity derived from the previous image's
    //      longest walk (512 the first time) fits the budget, the entries themselves at pixel * cap (one trace instead of two) ----
    uint32_t cap = 0;
    if (!generic && npix <= pix_budget) {
        uint32_t want = ctx->im_last_max ? std::max<uint32_t>(128u, (ctx->im_last_max + ctx->im_last_max / 4 + 31u) & ~31u) : 512u;
        if (const char* v = getenv("MGB_IMAGE_CAP")) want = (uint32_t)std::max(1l, atol(v)); /* tests: force the fallback */
        if ((unsigned long long)want * npix <= entry_budget) cap = want;
    }
    Stopwatch tt(ctx->stream);
    tt.start();
    if (cap) {
        MGB_CUDA(ctx, ensure_entries((size_t)cap * npix));
        ImagePaths capped{nullptr, ctx->im_nr.as<uint32_t>(), ctx->im_nri.as<uint32_t>(), ctx->im_fac.as<double>(), ctx->im_sh.as<double>(),
                          ctx->im_dz.as<double>(), cap};
        k_image_trace<2><<<trace_blocks, 128, 0, ctx->stream>>>(m, g, pix_begin, npix, ctx->im_X.as<double>(), ctx->im_Y.as<double>(),
                                                                 ctx->im_len.as<uint32_t>(), ctx->im_start.as<uint32_t>(),
                                                                 ctx->im_shift0.as<double>(), capped, ctx->counters.as<unsigned long long>());
    } else {
        ImagePaths none{};
        k_image_trace<0><<<trace_blocks, 128, 0, ctx->stream>>>(m, g, pix_begin, npix, ctx->im_X.as<double>(), ctx->im_Y.as<double>(),
                                                                 ctx->im_len.as<uint32_t>(), ctx->im_start.as<uint32_t>(),
                                                                 ctx->im_shift0.as<double>(), none, ctx->counters.as<unsigned long long>());
    }
    MGB_CUDA(ctx, cudaGetLastError());
    st.n_launches++;
    st.ms_trace += tt.stop();
    unsigned long long cnt[5];
    MGB_CUDA(ctx, cudaMemcpy(cnt, ctx->counters.p, 40, cudaMemcpyDeviceToHost));
    st.n_entries = cnt[0]; st.max_ray_length = (uint32_t)cnt[1];
    // ---- the lines some bin of this image can have in its window: sorted indices [slo, slo + nsel). A bin at frequency nu sees
    //      nu * shift -+ max_dist * (nu * shift) * sqrtw_max (solver.tpp:1259-1302, 2012-2111); bounds over all bins, all shifts of the
    //      traced rays and the largest sqrtw_max, widened by 1e-9 relative (rounding is 1e-16), never miss a line the kernels' own
    //      searches would find. ApproximationType None sums every line; per-point frequency rows keep every line as well. ----
    uint32_t slo = 0, nsel = m.nlines;
    if (!generic && approx == APPROX_CLOSELINES && !nu_per_point && cnt[2] != 0 && cnt[2] != ~0ull && ctx->h_sorted_line.size() == m.nlines) {
        double sh_min, sh_max, sqw;
        memcpy(&sh_min, &cnt[2], 8); memcpy(&sh_max, &cnt[3], 8); memcpy(&sqw, &cnt[4], 8);
        const double nu_min = *std::min_element(nu, nu + nfreqs), nu_max = *std::max_element(nu, nu + nfreqs);
        if (nu_min > 0.0 && sqw >= 0.0 && std::isfinite(nu_max) && std::isfinite(sqw) && std::isfinite(ctx->m.max_dist) && ctx->m.max_dist >= 0.0) {
            const double wide = 1.0 + 1.0e-9;
            const double hi_f = nu_max * sh_max * wide, lo_f = nu_min * sh_min / wide;
            const double ub = ctx->m.max_dist * hi_f * sqw * wide;
            const auto& sl = ctx->h_sorted_line;
            slo  = (uint32_t)(std::lower_bound(sl.begin(), sl.end(), lo_f - ub) - sl.begin());
            nsel = (uint32_t)(std::upper_bound(sl.begin(), sl.end(), hi_f + ub) - sl.begin()) - slo;
        }
    }
    if (!generic) entry_budget = std::max<unsigned long long>(1ull << 20, ctx->params.path_cache_bytes / (64ull + 32ull * nsel));
    ctx->im_last_max = (uint32_t)cnt[1];
    if (cap && (cnt[1] > cap || cnt[0] + npix > entry_budget)) cap = 0; /* some walk did not fit: count is known, fill in a second pass */
    std::vector<uint32_t> hlen;
    if (!cap) {
        hlen.resize(npix);
        MGB_CUDA(ctx, cudaMemcpy(hlen.data(), ctx->im_len.p, 4 * (size_t)npix, cudaMemcpyDeviceToHost));
    }
    for (uint32_t p0 = 0; p0 < npix;) {
        uint32_t p1 = p0;
        unsigned long long ents = 0;
        if (cap) { p1 = npix; ents = cnt[0]; }
        else
            while (p1 < npix && (p1 - p0) < pix_budget) {
                if (p1 > p0 && ents + hlen[p1] + (p1 - p0) > entry_budget) break;
                ents += hlen[p1]; p1++;
            }
        const uint32_t np = p1 - p0;
        const size_t ne   = (size_t)std::max<unsigned long long>(ents, 1);
        if (!cap) MGB_CUDA(ctx, ensure_entries(ne));
        MGB_CUDA(ctx, ctx->im_I.ensure(8 * (size_t)np * nfreqs));
        tt.start();
        const uint32_t* len_c = ctx->im_len.as<uint32_t>() + p0;
        unsigned long long* poff = ctx->im_poff.as<unsigned long long>() + p0;
        size_t tmp = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, tmp, len_c, poff, (int64_t)np, ctx->stream);
        MGB_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
        MGB_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->cub_tmp.p, tmp, len_c, poff, (int64_t)np, ctx->stream));
        st.n_launches++;
        ImagePaths ip{poff, ctx->im_nr.as<uint32_t>(), ctx->im_nri.as<uint32_t>(), ctx->im_fac.as<double>(), ctx->im_sh.as<double>(),
                      ctx->im_dz.as<double>(), cap};
        if (!cap) {
            k_image_trace<1><<<(unsigned)(((size_t)np * 32 + 127) / 128), 128, 0, ctx->stream>>>(
                m, g, pix_begin + p0, np, ctx->im_X.as<double>(), ctx->im_Y.as<double>(), ctx->im_len.as<uint32_t>() + p0,
                ctx->im_start.as<uint32_t>() + p0, ctx->im_shift0.as<double>() + p0, ip, nullptr);
            MGB_CUDA(ctx, cudaGetLastError());
            st.n_launches++;
        }
        st.ms_trace += tt.stop();
        tt.start();
        const size_t nitems = (size_t)np * nchunks_f;
        const unsigned blocks = (unsigned)((nitems * 32 + 127) / 128);
        const uint32_t stride = nu_per_point ? nfreqs : 0u;
        if (generic) {
#define MGB_IMAGE_LAUNCH(A, T)                                                                                                          \
    k_image_solve<A, T><<<blocks, 128, 0, ctx->stream>>>(m, np, nfreqs, nchunks_f, ctx->im_nu.as<double>(), stride, len_c,              \
                                                         ctx->im_start.as<uint32_t>() + p0, ctx->im_shift0.as<double>() + p0, ip,         \
                                                         ctx->im_I.as<double>())
            if (approx == APPROX_NONE) { if (optical_depth) MGB_IMAGE_LAUNCH(APPROX_NONE, true); else MGB_IMAGE_LAUNCH(APPROX_NONE, false); }
            else { if (optical_depth) MGB_IMAGE_LAUNCH(APPROX_CLOSELINES, true); else MGB_IMAGE_LAUNCH(APPROX_CLOSELINES, false); }
#undef MGB_IMAGE_LAUNCH
        } else {
            const size_t npos = ne + np; /* every pixel's entries + its start point */
            MGB_CUDA(ctx, grow(ctx->im_hdr, 32 * npos));
            MGB_CUDA(ctx, grow(ctx->im_rec, 32 * npos * std::max(nsel, 1u)));
            ImageRecs ir{ctx->im_hdr.as<double4>(), ctx->im_rec.as<double4>(), slo, nsel};
            const unsigned kblocks = (unsigned)(((size_t)np * 32 + 127) / 128);
            if (approx == APPROX_CLOSELINES)
                k_image_records<true><<<kblocks, 128, 0, ctx->stream>>>(m, np, len_c, ctx->im_start.as<uint32_t>() + p0,
                                                                        ctx->im_shift0.as<double>() + p0, ip, ir);
            else
                k_image_records<false><<<kblocks, 128, 0, ctx->stream>>>(m, np, len_c, ctx->im_start.as<uint32_t>() + p0,
                                                                         ctx->im_shift0.as<double>() + p0, ip, ir);
            MGB_CUDA(ctx, cudaGetLastError());
            st.n_launches++;
            ImageItems it{nfreqs / 32, nfreqs % 32, (nfreqs % 32) ? 32 / (nfreqs % 32) : 0};
            // The chunk is solved in slices of pixels: slice k goes device -> pinned staging on the copy stream while slice k + 1 is
            // solved, and host threads move the staged slices into the caller's buffer (a plain cudaMemcpy of 210 MB into pageable
            // memory took longer than the kernels).
            const size_t chunk_bytes = 8 * (size_t)np * nfreqs;
            const bool staged = chunk_bytes <= kStageMaxBytes;
            const uint32_t nslices = (staged && np >= 8192) ? 8u : 1u;
            const uint32_t align = it.group ? it.group : 1u;
            std::vector<cudaEvent_t> ev(staged ? 2 * nslices : 0, nullptr);
            std::vector<uint32_t> cut(nslices + 1, 0);
            for (uint32_t k = 1; k < nslices; k++) cut[k] = (uint32_t)((size_t)np * k / nslices) / align * align;
            cut[nslices] = np;
            cudaError_t perr = cudaSuccess;
            if (staged) {
                if ((perr = ctx->im_stage.ensure(chunk_bytes)) == cudaSuccess && !ctx->im_copy_stream.s)
                    perr = cudaStreamCreateWithFlags(&ctx->im_copy_stream.s, cudaStreamNonBlocking);
                for (auto& e : ev)
                    if (perr == cudaSuccess) perr = cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
            }
            auto drop_events = [&] { for (auto e : ev) if (e) cudaEventDestroy(e); };
            if (perr != cudaSuccess) { drop_events(); MGB_CUDA(ctx, perr); }
            for (uint32_t k = 0; k < nslices && perr == cudaSuccess; k++) {
                const uint32_t a = cut[k], ns = cut[k + 1] - cut[k];
                if (ns) {
                    const unsigned rblocks = (unsigned)((it.count(ns) * 32 + 127) / 128);
#define MGB_IMAGE_LAUNCH(A, T)                                                                                                          \
    k_image_solve_rec<A, T><<<rblocks, 128, 0, ctx->stream>>>(m, a, ns, nfreqs, it, ctx->im_nu.as<double>(), stride, len_c,             \
                                                              ctx->im_start.as<uint32_t>() + p0, ip, ir.hdr, ir.rec, ir.slo, ir.nsel, ctx->im_I.as<double>())
                    if (approx == APPROX_NONE) { if (optical_depth) MGB_IMAGE_LAUNCH(APPROX_NONE, true); else MGB_IMAGE_LAUNCH(APPROX_NONE, false); }
                    else { if (optical_depth) MGB_IMAGE_LAUNCH(APPROX_CLOSELINES, true); else MGB_IMAGE_LAUNCH(APPROX_CLOSELINES, false); }
#undef MGB_IMAGE_LAUNCH
                    perr = cudaGetLastError();
                    st.n_launches++;
                }
                if (staged && perr == cudaSuccess) {
                    const size_t off = (size_t)a * nfreqs;
                    if ((perr = cudaEventRecord(ev[2 * k], ctx->stream)) == cudaSuccess &&
                        (perr = cudaStreamWaitEvent(ctx->im_copy_stream.s, ev[2 * k], 0)) == cudaSuccess &&
                        (ns == 0 || (perr = cudaMemcpyAsync((double*)ctx->im_stage.p + off, ctx->im_I.as<double>() + off, 8 * (size_t)ns * nfreqs,
                                                            cudaMemcpyDeviceToHost, ctx->im_copy_stream.s)) == cudaSuccess))
                        perr = cudaEventRecord(ev[2 * k + 1], ctx->im_copy_stream.s);
                }
            }
            if (perr != cudaSuccess) { cudaStreamSynchronize(ctx->stream); if (staged) cudaStreamSynchronize(ctx->im_copy_stream.s); drop_events(); MGB_CUDA(ctx, perr); }
            if (staged) {
                const unsigned nthreads = std::max(1u, std::min(8u, std::thread::hardware_concurrency()));
                std::vector<cudaError_t> werr(nthreads, cudaSuccess);
                std::vector<std::thread> workers;
                const int device = ctx->device;
                const double* stage = (const double*)ctx->im_stage.p;
                double* dst = I_out + (size_t)p0 * nfreqs;
                for (unsigned t = 0; t < nthreads; t++)
                    workers.emplace_back([&, t] {
                        cudaSetDevice(device);
                        for (uint32_t k = 0; k < nslices; k++) {
                            const cudaError_t e = cudaEventSynchronize(ev[2 * k + 1]);
                            if (e != cudaSuccess) { werr[t] = e; return; }
                            const size_t n = (size_t)(cut[k + 1] - cut[k]) * nfreqs, b = n * t / nthreads, c = n * (t + 1) / nthreads;
                            memcpy(dst + (size_t)cut[k] * nfreqs + b, stage + (size_t)cut[k] * nfreqs + b, 8 * (c - b));
                        }
                    });
                st.ms_solve += tt.stop(); /* the kernels; the workers drain the slices meanwhile */
                for (auto& w : workers) w.join();
                drop_events();
                for (auto e : werr) MGB_CUDA(ctx, e);
            } else {
                st.ms_solve += tt.stop();
                MGB_CUDA(ctx, cudaMemcpy(I_out + (size_t)p0 * nfreqs, ctx->im_I.p, chunk_bytes, cudaMemcpyDeviceToHost));
            }
        }
        if (generic) {
            MGB_CUDA(ctx, cudaGetLastError());
            st.n_launches++;
            st.ms_solve += tt.stop();
            MGB_CUDA(ctx, cudaMemcpy(I_out + (size_t)p0 * nfreqs, ctx->im_I.p, 8 * (size_t)np * nfreqs, cudaMemcpyDeviceToHost));
        }
        st.n_chunks++;
        p0 = p1;
    }
    st.ms_total = total.stop();
    st.n_pixels = npix; st.n_solves = (uint64_t)npix * nfreqs; st.start_bdy = g.start_bdy;
    if (stats) *stats = st;
    return MGB_OK;
}

extern "C" int mgb_image_feautrier_order_2(mgb_ctx* ctx, int approx, int dimension, const double* ray_dir, uint32_t nx, uint32_t ny,
                                           uint32_t nfreqs, const double* nu, int nu_per_point, uint64_t pix_begin, uint64_t pix_end,
                                           double* ImX_out, double* ImY_out, double* I_out, mgb_image_stats* stats) {
    return image_impl(ctx, false, approx, dimension, ray_dir, nx, ny, nfreqs, nu, nu_per_point, pix_begin, pix_end, ImX_out, ImY_out, I_out, stats);
}

extern "C" int mgb_image_optical_depth(mgb_ctx* ctx, int approx, int dimension, const double* ray_dir, uint32_t nx, uint32_t ny,
                                       uint32_t nfreqs, const double* nu, int nu_per_point, uint64_t pix_begin, uint64_t pix_end,
                                       double* ImX_out, double* ImY_out, double* tau_out, mgb_image_stats* stats) {
    return image_impl(ctx, true, approx, dimension, ray_dir, nx, ny, nfreqs, nu, nu_per_point, pix_begin, pix_end, ImX_out, ImY_out, tau_out, stats);
}
