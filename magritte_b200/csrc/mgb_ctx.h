// mgb_ctx.h -- host-side context of libmagritte_b200.so (internal; the public surface is include/magritte_b200.h).
#pragma once
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h> /* header-only NVTX 3: ranges show up in nsys / ncu timelines, no-ops without a tool attached */

#include "../../include/magritte_b200.h"
#include "mgb_device.cuh"

namespace mgb {

// Owning device allocation (grow-only reuse through ensure()).
struct DevBuf {
    void* p      = nullptr;
    size_t bytes = 0;
    ~DevBuf() { release(); }
    DevBuf() = default;
    DevBuf(const DevBuf&)            = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    void release() {
        if (p) cudaFree(p);
        p = nullptr; bytes = 0;
    }
    cudaError_t ensure(size_t n) {
        if (n <= bytes && p) return cudaSuccess;
        release();
        if (n == 0) n = 8;
        cudaError_t e = cudaMalloc(&p, n);
        if (e == cudaSuccess) bytes = n; else p = nullptr;
        return e;
    }
    template <class T> T* as() const { return static_cast<T*>(p); }
};

struct PathChunk {
    uint32_t o0 = 0, o1 = 0;
    unsigned long long steps = 0;
    bool filled = false;
    DevBuf poff, nr, dZ, shift, shift0;
};

// ---- level populations on the device (mgb_levelpop.cu) ----
struct LpColPar { // CollisionPartner (reference linedata/collisionPartner/collisionPartner.hpp)
    uint32_t num_col_partner = 0, orth_or_para = 0 /* 0 'n', 1 'o', 2 'p' */, ntmp = 0, ncol = 0;
    DevBuf icol, jcol, tmp, Ce, Cd; // Ce, Cd [ntmp][ncol]
};
struct LpSpecies { // LineProducingSpecies + Linedata (lineProducingSpecies.hpp, linedata.hpp)
    uint32_t nlev = 0, nrad = 0, num = 0, line0 = 0;
    bool have_pop = false;
    DevBuf irad, jrad, energy, weight, frequency, A, Ba, Bs;
    std::vector<LpColPar*> colpar;
    DevBuf colpar_views;
    DevBuf pop, prev1, prev2, prev3, trial, trial_prev; // double2 {hi, lo} [npoints][nlev]: the reference's long double populations
    std::vector<DevBuf*> populations, residuals;         // Ng history (lineProducingSpecies.hpp:56-57)
    DevBuf Jlin, Jeff, Jdif;                             // [npoints][nrad]
    DevBuf pop_tot;                                      // [npoints]
    ~LpSpecies() {
        for (auto* c : colpar) delete c;
        for (auto* b : populations) delete b;
        for (auto* b : residuals) delete b;
    }
};

// NVTX range for the lifetime of the object (one per phase of a call: trace, solve, reduce, all-gather, statistical equilibrium ...)
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};

} // namespace mgb

struct mgb_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_a = nullptr, ev_b = nullptr;
    std::string err;
    mgb_params params;
    bool have_geometry = false, have_thermo = false, have_lines = false, have_freqs = false, have_state = false;
    bool points_dirty = true; // sqrtw_max / dshift_max need recomputing
    // host copies of small tables
    std::vector<uint32_t> nrad, nrad_cum, line_spec, line_k;
    std::vector<double> inverse_mass;
    double max_inverse_mass = 0.0;
    uint32_t n_freq_chunks = 0, lines_per_warp = 1; // k_solve work decomposition (build_freq_chunks)
    uint32_t n_freq_chunks2 = 0, lines_per_warp2 = 1; // k_solve2: whole lines per warp, two bins per lane
    int solve_v2 = 4;                                 // solve kernel generation for n_off_diag = 0 (MGB_SOLVE_KERNEL = 1 .. 4; default 4: the pipeline)
    mgb::DevModel m{};
    // device storage
    mgb::DevBuf pos, vel, csr, nbs, dir, weight, antipod, p2b, bcond, btemp;
    mgb::DevBuf temperature, vturb2, sqrtw_max, dshift_max, inv_mass_dev;
    mgb::DevBuf lrow, line, sorted_line, sorted_map, sorted_pos, freq_chunks, freq_chunks2, line_A, line_invmass, line_k_dev, line_spec_dev, line_quadw;
    mgb::DevBuf nu, corr_line;
    mgb::DevBuf stage_a, stage_b, stage_c, stage_d; // H2D staging of the per-iteration line state
    // results
    mgb::DevBuf J, J_lines, J_anis, lam_val, lam_cnt, u_full;
    bool res_anis = false;           // the last run filled J_anis
    mgb::DevBuf lam_list_val, lam_list_nr, scratch_offV, scratch_offN; // banded Lambda operator (n_off_diag > 0)
    uint32_t lam_cap = 1;
    uint32_t u_o0 = 0, u_no = 0;
    uint32_t res_o0 = 0, res_o1 = 0; // origin range of the last run
    uint32_t res_K = 0;              // n_off_diag of the last run (lam_cap is that run's list capacity)
    bool res_valid = false;
    // tracing state
    mgb::DevBuf lens;       // [npoints][nrays]
    uint32_t lens_o0 = 0, lens_o1 = 0;
    std::vector<unsigned long long> origin_steps; // per origin in [lens_o0, lens_o1)
    unsigned long long tot_steps = 0, tot_scanned = 0;
    uint32_t max_len = 0;
    std::vector<mgb::PathChunk*> chunks; // cached path chunks (valid for plan_o0..plan_o1)
    uint32_t plan_o0 = 0, plan_o1 = 0;
    bool plan_cached = false;
    mgb::PathChunk work;    // reusable chunk when paths are not cached
    mgb::DevBuf scratch_u, scratch_L, scratch_fac, counters, origin_tot, cub_tmp;
    // streaming pipeline (mgb_solve4.cuh, MGB_SOLVE_KERNEL=4): list lengths / offsets, ray-entry records, large-shift slots and line integrals
    mgb::DevBuf rec_cnt, rec_off, rec, rec_pts, rec_lcount, rec_slots, rec_trip, side_of, rec_eoff, rec_ni, rec_base;
    bool side_ok = false;                             // every ray is either a `rr` (< hnrays) or the antipode of one, never both
    float ms_v4[4] = {0, 0, 0, 0};                    // last run: records (count + fill), slots + line integrals, sweep, (spare)
    // imager: host copies of what Image::set_coordinates_projection_surface reads (boundary points in boundary order,
    // Points::center, point2boundary[0])
    std::vector<double> h_bpos;      // [nboundary][3]
    std::vector<uint32_t> h_b2p;     // [nboundary]
    double h_center[3] = {0, 0, 0};
    uint32_t h_p2b0 = 0;
    double h_pos_p2b0[3] = {0, 0, 0}; // position[point2boundary[0]] when that is a valid point index
    double h_pos_last[3] = {0, 0, 0}; // position[npoints-1] (start point of a spherically symmetric image)
    mgb::DevBuf im_X, im_Y, im_len, im_start, im_shift0, im_poff, im_nr, im_nri, im_fac, im_sh, im_dz, im_I, im_nu;
    mgb::DevBuf im_hdr, im_rec;      // imager records: per-entry headers and per-(entry, line) interpolated line data (mgb_image.cu)
    std::vector<double> h_sorted_line; // host copy of lines.sorted_line
    uint32_t im_last_max = 0;        // longest pixel walk of the previous image: sizes the single-pass trace's per-pixel capacity
    // imager result path: slices of the image go device -> this pinned staging buffer on a second stream while the next slice is
    // solved; host threads move them on into the caller's (pageable) buffer
    struct PinnedBuf {
        void* p = nullptr; size_t bytes = 0;
        ~PinnedBuf() { if (p) cudaFreeHost(p); }
        cudaError_t ensure(size_t n) {
            if (n <= bytes && p) return cudaSuccess;
            if (p) cudaFreeHost(p);
            p = nullptr; bytes = 0;
            cudaError_t e = cudaHostAlloc(&p, n, cudaHostAllocDefault);
            if (e == cudaSuccess) bytes = n; else p = nullptr;
            return e;
        }
    } im_stage;
    struct OwnedStream {
        cudaStream_t s = nullptr;
        ~OwnedStream() { if (s) cudaStreamDestroy(s); }
    } im_copy_stream;
    // level populations on the device
    std::vector<mgb::LpSpecies*> lp;
    mgb::DevBuf lp_abundance, lp_partials, lp_flags;
    uint32_t lp_nspecs = 0;
    double lp_min_line_opacity = 1.0e-13, lp_min_line_emissivity = 1.0e-50, lp_population_inversion_fraction = -1.0,
           lp_min_rel_pop_for_convergence = 1.0e-10;
    // NCCL
    void* comm = nullptr;
    int nranks = 1, rank = 0;
    std::vector<uint32_t> part_bounds; // [nranks+1] origin range of every rank (set by mgb_partition / mgb_partition_balanced)

    ~mgb_ctx();
    int fail(int code, const char* fmt, ...) {
        char buf[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof buf, fmt, ap);
        va_end(ap);
        err = buf;
        return code;
    }
    void drop_chunks() {
        for (auto* c : chunks) delete c;
        chunks.clear();
        plan_cached = false;
    }
};

int mgb_internal_refresh_points(mgb_ctx* ctx); // mgb_api.cu
int mgb_internal_alloc_results(mgb_ctx* ctx);  // mgb_api.cu

#define MGB_CUDA(ctx, call)                                                                                         \
    do {                                                                                                            \
        cudaError_t e_ = (call);                                                                                    \
        if (e_ != cudaSuccess) return (ctx)->fail(MGB_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                                                  __FILE__, __LINE__);                                              \
    } while (0)
