// mgb_api.cu -- C ABI of libmagritte_b200.so: context, model upload, the Feautrier order-2 run, result download.
// See include/magritte_b200.h for the contract and the reference citations of every entry point.
#include <algorithm>
#include <cstdlib>
#include <numeric>

#include <cub/device/device_scan.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include "mgb_ctx.h"
#ifndef MGB_DEFAULT_SOLVE_KERNEL
#define MGB_DEFAULT_SOLVE_KERNEL 4
#endif
#include "mgb_solve4.cuh"
#include "mgb_shortchar.cuh"

using namespace mgb;

static std::string g_create_error;

mgb_ctx::~mgb_ctx() {
    drop_chunks();
    for (auto* s : lp) delete s;
    if (ev_a) cudaEventDestroy(ev_a);
    if (ev_b) cudaEventDestroy(ev_b);
    if (stream) cudaStreamDestroy(stream);
}

static int upload(mgb_ctx* ctx, DevBuf& b, const void* host, size_t bytes) {
    MGB_CUDA(ctx, b.ensure(bytes));
    if (bytes) MGB_CUDA(ctx, cudaMemcpyAsync(b.p, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return MGB_OK;
}
#define MGB_TRY(x)            \
    do {                      \
        int rc_ = (x);        \
        if (rc_ != MGB_OK) return rc_; \
    } while (0)

extern "C" int mgb_abi_version(void) { return MGB_ABI_VERSION; }

extern "C" int mgb_params_default(mgb_params* p) {
    if (!p) return MGB_EINVAL;
    memset(p, 0, sizeof *p);
    p->min_opacity                       = 1.0e-26;
    p->max_distance_opacity_contribution = 10.0;
    p->max_interpolation_diff            = 1.4;
    p->max_width_fraction                = 0.35;
    p->n_off_diag                        = 0;
    p->store_intensities                 = 0;
    p->path_cache_bytes                  = 48ull << 30;
    p->scratch_bytes                     = 8ull << 30;
    p->cache_paths                       = 1;
    p->anisotropic                       = 0;
    return MGB_OK;
}

extern "C" int mgb_create(int device, mgb_ctx** out) {
    if (!out) return MGB_EINVAL;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) + " (libmagritte_b200 has no CPU fallback)";
        return MGB_ECUDA;
    }
    if (device < 0 || device >= ndev) { g_create_error = "device index out of range"; return MGB_EINVAL; }
    if ((e = cudaSetDevice(device)) != cudaSuccess) { g_create_error = cudaGetErrorString(e); return MGB_ECUDA; }
    mgb_ctx* ctx = new mgb_ctx();
    ctx->device  = device;
    if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) {
        g_create_error = cudaGetErrorString(e);
        delete ctx;
        return MGB_ECUDA;
    }
    mgb_params defaults;
    mgb_params_default(&defaults);
    mgb_set_params(ctx, &defaults);
    *out = ctx;
    return MGB_OK;
}

extern "C" void mgb_destroy(mgb_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    mgb_comm_destroy(ctx);
    delete ctx;
}

extern "C" const char* mgb_last_error(const mgb_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

extern "C" int mgb_set_params(mgb_ctx* ctx, const mgb_params* p) {
    if (!ctx || !p) return MGB_EINVAL;
    if (p->n_off_diag > (uint32_t)kMaxOffDiag) return ctx->fail(MGB_EUNSUPPORTED, "n_off_diag = %u: at most %d off-diagonals are supported", p->n_off_diag, kMaxOffDiag);
    if (p->max_width_fraction != ctx->params.max_width_fraction) ctx->points_dirty = true;
    ctx->params = *p;
    if (ctx->params.path_cache_bytes == 0) ctx->params.path_cache_bytes = 48ull << 30;
    if (ctx->params.scratch_bytes == 0) ctx->params.scratch_bytes = 8ull << 30;
    ctx->m.min_opacity = p->min_opacity;
    ctx->m.max_dist    = p->max_distance_opacity_contribution;
    ctx->m.log_max_interp_diff = (double)logf((float)p->max_interpolation_diff); /* logf(double) as in solver_interp_helper.tpp:68 */
    return MGB_OK;
}

extern "C" int mgb_set_geometry(mgb_ctx* ctx, uint32_t npoints, const double* position, const double* velocity,
                                const uint32_t* n_neighbors, const uint32_t* neighbors, uint64_t n_neighbors_total,
                                uint32_t nrays, const double* direction, const double* weight, const uint32_t* antipod,
                                uint32_t nboundary, const uint32_t* point2boundary, const int32_t* boundary_condition,
                                const double* boundary_temperature, int spherical_symmetry) {
    if (!ctx) return MGB_EINVAL;
    if (!npoints || !position || !velocity || !n_neighbors || !neighbors || !nrays || (nrays & 1) || !direction || !weight ||
        !antipod || !point2boundary || (nboundary && (!boundary_condition || !boundary_temperature)))
        return ctx->fail(MGB_EINVAL, "mgb_set_geometry: null or inconsistent argument");
    cudaSetDevice(ctx->device);
    std::vector<double4> p4(npoints), v4(npoints);
    std::vector<uint2> csr(npoints);
    uint64_t cum = 0;
    for (uint32_t i = 0; i < npoints; i++) {
        p4[i] = make_double4(position[3 * i], position[3 * i + 1], position[3 * i + 2], 0.0);
        v4[i] = make_double4(velocity[3 * i], velocity[3 * i + 1], velocity[3 * i + 2], 0.0);
        csr[i] = make_uint2((uint32_t)cum, n_neighbors[i]);
        cum += n_neighbors[i];
    }
    if (cum != n_neighbors_total) return ctx->fail(MGB_EINVAL, "mgb_set_geometry: sum(n_neighbors) = %llu != n_neighbors_total = %llu",
                                                   (unsigned long long)cum, (unsigned long long)n_neighbors_total);
    if (cum > 0xffffffffull) return ctx->fail(MGB_EINVAL, "mgb_set_geometry: neighbour list exceeds 2^32 entries");
    for (uint64_t i = 0; i < cum; i++)
        if (neighbors[i] >= npoints) return ctx->fail(MGB_EINVAL, "mgb_set_geometry: neighbour index out of range");
    for (uint32_t r = 0; r < nrays; r++)
        if (antipod[r] >= nrays) return ctx->fail(MGB_EINVAL, "mgb_set_geometry: antipod index out of range");
    MGB_TRY(upload(ctx, ctx->pos, p4.data(), sizeof(double4) * npoints));
    MGB_TRY(upload(ctx, ctx->vel, v4.data(), sizeof(double4) * npoints));
    MGB_TRY(upload(ctx, ctx->csr, csr.data(), sizeof(uint2) * npoints));
    MGB_TRY(upload(ctx, ctx->nbs, neighbors, sizeof(uint32_t) * cum));
    MGB_TRY(upload(ctx, ctx->dir, direction, sizeof(double) * 3 * nrays));
    MGB_TRY(upload(ctx, ctx->weight, weight, sizeof(double) * nrays));
    MGB_TRY(upload(ctx, ctx->antipod, antipod, sizeof(uint32_t) * nrays));
    {   /* which side of a ray pair a half ray is (k_large: orientation of the generic line integral): 0 = the pair's ray rr, 1 = its antipode */
        std::vector<uint8_t> side(nrays, 2);
        ctx->side_ok = true;
        for (uint32_t rr = 0; rr < nrays / 2; rr++) {
            if (side[rr] == 1 || antipod[rr] < nrays / 2) ctx->side_ok = false;
            side[rr] = 0; side[antipod[rr]] = 1;
        }
        for (uint32_t r = 0; r < nrays; r++) if (side[r] == 2) { ctx->side_ok = false; side[r] = 1; }
        MGB_TRY(upload(ctx, ctx->side_of, side.data(), nrays));
    }
    MGB_TRY(upload(ctx, ctx->p2b, point2boundary, sizeof(uint32_t) * npoints));
    MGB_TRY(upload(ctx, ctx->bcond, boundary_condition, sizeof(int32_t) * nboundary));
    MGB_TRY(upload(ctx, ctx->btemp, boundary_temperature, sizeof(double) * nboundary));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); /* host vectors go out of scope */
    {   /* what the imager's projection surface needs on the host (image.cpp:221-368, points.cpp:22-42, geometry.tpp:346-364) */
        ctx->h_b2p.assign(nboundary, 0);
        ctx->h_bpos.assign((size_t)nboundary * 3, 0.0);
        double mn[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, mx[3] = {DBL_MIN, DBL_MIN, DBL_MIN}; /* numeric_limits<Real>::min() as in the reference */
        for (uint32_t i = 0; i < npoints; i++) {
            for (int k = 0; k < 3; k++) {
                const double v = position[3 * i + k];
                if (v < mn[k]) mn[k] = v;
                if (v > mx[k]) mx[k] = v;
            }
            const uint32_t b = point2boundary[i];
            if (b < nboundary) {
                ctx->h_b2p[b] = i;
                for (int k = 0; k < 3; k++) ctx->h_bpos[(size_t)b * 3 + k] = position[3 * i + k];
            }
        }
        for (int k = 0; k < 3; k++) ctx->h_center[k] = (mn[k] + mx[k]) / 2;
        ctx->h_p2b0 = point2boundary[0];
        for (int k = 0; k < 3; k++) ctx->h_pos_last[k] = position[3 * (size_t)(npoints - 1) + k];
        if (ctx->h_p2b0 < npoints) for (int k = 0; k < 3; k++) ctx->h_pos_p2b0[k] = position[3 * (size_t)ctx->h_p2b0 + k];
    }
    DevModel& m = ctx->m;
    m.npoints = npoints; m.nrays = nrays; m.hnrays = nrays / 2; m.nboundary = nboundary; m.spherical = spherical_symmetry ? 1 : 0;
    m.adaptive = 0;
    m.pos = ctx->pos.as<double4>(); m.vel = ctx->vel.as<double4>(); m.csr = ctx->csr.as<uint2>(); m.nbs = ctx->nbs.as<uint32_t>();
    m.dir = ctx->dir.as<double>(); m.weight = ctx->weight.as<double>(); m.antipod = ctx->antipod.as<uint32_t>();
    m.p2b = ctx->p2b.as<uint32_t>(); m.bcond = ctx->bcond.as<int32_t>(); m.btemp = ctx->btemp.as<double>();
    ctx->have_geometry = true;
    ctx->have_thermo = ctx->have_freqs = ctx->have_state = false; /* sized by npoints */
    ctx->lens_o0 = ctx->lens_o1 = 0;
    ctx->drop_chunks();
    ctx->points_dirty = true;
    ctx->res_valid = false;
    return MGB_OK;
}

// Rays::use_adaptive_directions (rays.hpp:23-31): a ray set per point. direction [npoints][nrays][3], weight [npoints][nrays]; the antipode
// table stays the one of mgb_set_geometry (the reference derives it from point 0's set and assumes it for every point, rays.cpp:52-63).
extern "C" int mgb_set_adaptive_directions(mgb_ctx* ctx, const double* direction, const double* weight) {
    if (!ctx) return MGB_EINVAL;
    if (!ctx->have_geometry) return ctx->fail(MGB_ESTATE, "mgb_set_adaptive_directions before mgb_set_geometry");
    if (!direction || !weight) return ctx->fail(MGB_EINVAL, "mgb_set_adaptive_directions: null argument");
    cudaSetDevice(ctx->device);
    const size_t n = (size_t)ctx->m.npoints * ctx->m.nrays;
    MGB_TRY(upload(ctx, ctx->dir, direction, sizeof(double) * 3 * n));
    MGB_TRY(upload(ctx, ctx->weight, weight, sizeof(double) * n));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->m.dir = ctx->dir.as<double>(); ctx->m.weight = ctx->weight.as<double>();
    ctx->m.adaptive = 1;
    ctx->lens_o0 = ctx->lens_o1 = 0; /* every path changes */
    ctx->drop_chunks();
    ctx->res_valid = false;
    return MGB_OK;
}

extern "C" int mgb_set_thermodynamics(mgb_ctx* ctx, const double* temperature_gas, const double* vturb2) {
    if (!ctx) return MGB_EINVAL;
    if (!ctx->have_geometry) return ctx->fail(MGB_ESTATE, "mgb_set_thermodynamics before mgb_set_geometry");
    if (!temperature_gas || !vturb2) return ctx->fail(MGB_EINVAL, "mgb_set_thermodynamics: null argument");
    cudaSetDevice(ctx->device);
    const size_t n = ctx->m.npoints;
    MGB_TRY(upload(ctx, ctx->temperature, temperature_gas, 8 * n));
    MGB_TRY(upload(ctx, ctx->vturb2, vturb2, 8 * n));
    MGB_CUDA(ctx, ctx->sqrtw_max.ensure(8 * n));
    MGB_CUDA(ctx, ctx->dshift_max.ensure(8 * n));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->m.temperature = ctx->temperature.as<double>(); ctx->m.vturb2 = ctx->vturb2.as<double>();
    ctx->m.sqrtw_max = ctx->sqrtw_max.as<double>(); ctx->m.dshift_max = ctx->dshift_max.as<double>();
    ctx->have_thermo  = true;
    ctx->points_dirty = true;
    return MGB_OK;
}

extern "C" int mgb_set_lines(mgb_ctx* ctx, uint32_t nlspecs, const uint32_t* nrad, const double* inverse_mass, const double* line,
                             const double* einstein_A, uint32_t nquads, const double* quadrature_weights) {
    if (!ctx) return MGB_EINVAL;
    if (!nlspecs || !nrad || !inverse_mass || !line || !einstein_A || !nquads || !quadrature_weights)
        return ctx->fail(MGB_EINVAL, "mgb_set_lines: null or zero argument");
    cudaSetDevice(ctx->device);
    ctx->nrad.assign(nrad, nrad + nlspecs);
    ctx->inverse_mass.assign(inverse_mass, inverse_mass + nlspecs);
    ctx->nrad_cum.assign(nlspecs, 0);
    uint32_t nlines = 0;
    for (uint32_t l = 0; l < nlspecs; l++) { ctx->nrad_cum[l] = nlines; nlines += nrad[l]; }
    if (!nlines) return ctx->fail(MGB_EINVAL, "mgb_set_lines: no lines");
    ctx->line_spec.resize(nlines); ctx->line_k.resize(nlines);
    std::vector<double> linv(nlines), lquad((size_t)nlines * nquads);
    // max_inverse_mass: Lines::read starts from 0 and keeps the largest species value (lines.cpp:40-46)
    ctx->max_inverse_mass = 0.0;
    for (uint32_t l = 0; l < nlspecs; l++) {
        if (inverse_mass[l] > ctx->max_inverse_mass) ctx->max_inverse_mass = inverse_mass[l];
        for (uint32_t k = 0; k < nrad[l]; k++) {
            const uint32_t lid = ctx->nrad_cum[l] + k;
            ctx->line_spec[lid] = l; ctx->line_k[lid] = k; linv[lid] = inverse_mass[l];
            for (uint32_t z = 0; z < nquads; z++) lquad[(size_t)lid * nquads + z] = quadrature_weights[(size_t)l * nquads + z];
        }
    }
    // sorted line list + index map (lines.cpp:52-64, heapsort with index tracking; stable order for distinct centres)
    std::vector<uint32_t> smap(nlines);
    std::iota(smap.begin(), smap.end(), 0u);
    std::stable_sort(smap.begin(), smap.end(), [&](uint32_t a, uint32_t b) { return line[a] < line[b]; });
    std::vector<double> sline(nlines);
    for (uint32_t i = 0; i < nlines; i++) sline[i] = line[smap[i]];
    MGB_TRY(upload(ctx, ctx->line, line, 8 * (size_t)nlines));
    MGB_TRY(upload(ctx, ctx->line_A, einstein_A, 8 * (size_t)nlines));
    MGB_TRY(upload(ctx, ctx->sorted_line, sline.data(), 8 * (size_t)nlines));
    ctx->h_sorted_line = sline; /* the imager selects the lines an image can reach on the host */
    MGB_TRY(upload(ctx, ctx->sorted_map, smap.data(), 4 * (size_t)nlines));
    std::vector<uint32_t> spos(nlines);
    for (uint32_t i = 0; i < nlines; i++) spos[smap[i]] = i;
    MGB_TRY(upload(ctx, ctx->sorted_pos, spos.data(), 4 * (size_t)nlines));
    MGB_TRY(upload(ctx, ctx->line_invmass, linv.data(), 8 * (size_t)nlines));
    MGB_TRY(upload(ctx, ctx->line_k_dev, ctx->line_k.data(), 4 * (size_t)nlines));
    MGB_TRY(upload(ctx, ctx->line_spec_dev, ctx->line_spec.data(), 4 * (size_t)nlines));
    MGB_TRY(upload(ctx, ctx->line_quadw, lquad.data(), 8 * lquad.size()));
    MGB_TRY(upload(ctx, ctx->inv_mass_dev, inverse_mass, 8 * (size_t)nlspecs));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    DevModel& m = ctx->m;
    m.nlines = nlines; m.nlspecs = nlspecs; m.nquads = nquads;
    m.line = ctx->line.as<double>(); m.line_A = ctx->line_A.as<double>(); m.sorted_line = ctx->sorted_line.as<double>();
    m.sorted_map = ctx->sorted_map.as<uint32_t>(); m.sorted_pos = ctx->sorted_pos.as<uint32_t>(); m.line_invmass = ctx->line_invmass.as<double>();
    m.line_k = ctx->line_k_dev.as<uint32_t>(); m.line_quadw = ctx->line_quadw.as<double>();
    ctx->have_lines = true;
    ctx->have_state = ctx->have_freqs = false;
    ctx->points_dirty = true;
    return MGB_OK;
}

// k_solve's work decomposition: consecutive frequency bins, <= 32 per warp, spanning <= lines_per_warp lines.
static int build_freq_chunks(mgb_ctx* ctx) {
    const uint32_t nq = ctx->m.nquads, nf = ctx->m.nfreqs;
    uint32_t lpw = nq >= 32 ? 2 : (31 + nq - 1) / nq + 1;
    if (lpw > 8) lpw = 8;
    if (lpw > ctx->m.nlines) lpw = ctx->m.nlines;
    std::vector<FreqChunk> ch;
    for (uint32_t f0 = 0; f0 < nf;) {
        const uint32_t l0 = f0 / nq;
        uint32_t f1 = std::min(nf, f0 + 32);
        f1 = std::min(f1, (l0 + lpw) * nq);
        FreqChunk c;
        c.f0 = f0; c.nf = f1 - f0; c.l0 = l0; c.nl = (f1 - 1) / nq - l0 + 1;
        ch.push_back(c);
        f0 = f1;
    }
    MGB_TRY(upload(ctx, ctx->freq_chunks, ch.data(), sizeof(FreqChunk) * ch.size()));
    // k_solve2: whole lines per warp, two bins of one line per lane (lane = k * lpl + j owns bins z0+j and z0+j+lpl of line l0+k)
    std::vector<FreqChunk2> ch2;
    uint32_t lpw2 = 1;
    if (nq <= 64) {
        const uint32_t lpl = (nq + 1) / 2;
        lpw2 = std::max(1u, std::min(std::min(32u / lpl, 8u), ctx->m.nlines));
        for (uint32_t l0 = 0; l0 < ctx->m.nlines; l0 += lpw2) ch2.push_back(FreqChunk2{l0, std::min(lpw2, ctx->m.nlines - l0), 0, nq, lpl});
    } else {
        for (uint32_t l = 0; l < ctx->m.nlines; l++)
            for (uint32_t z0 = 0; z0 < nq; z0 += 64) {
                const uint32_t nz = std::min(64u, nq - z0);
                ch2.push_back(FreqChunk2{l, 1, z0, nz, (nz + 1) / 2});
            }
    }
    MGB_TRY(upload(ctx, ctx->freq_chunks2, ch2.data(), sizeof(FreqChunk2) * ch2.size()));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->n_freq_chunks = (uint32_t)ch.size();
    ctx->lines_per_warp = lpw;
    ctx->n_freq_chunks2 = (uint32_t)ch2.size();
    ctx->lines_per_warp2 = lpw2;
    return MGB_OK;
}

extern "C" int mgb_set_frequencies(mgb_ctx* ctx, uint32_t nfreqs, const double* nu, const uint32_t* corresponding_line) {
    if (!ctx) return MGB_EINVAL;
    if (!ctx->have_geometry || !ctx->have_lines) return ctx->fail(MGB_ESTATE, "mgb_set_frequencies before geometry/lines");
    if (!nu || !corresponding_line) return ctx->fail(MGB_EINVAL, "mgb_set_frequencies: null argument");
    if (nfreqs != ctx->m.nlines * ctx->m.nquads)
        return ctx->fail(MGB_EINVAL, "mgb_set_frequencies: nfreqs = %u but nlines*nquads = %u (the solver needs the line discretisation "
                                     "of Model::compute_spectral_discretisation())", nfreqs, ctx->m.nlines * ctx->m.nquads);
    cudaSetDevice(ctx->device);
    MGB_TRY(upload(ctx, ctx->nu, nu, 8 * (size_t)ctx->m.npoints * nfreqs));
    MGB_TRY(upload(ctx, ctx->corr_line, corresponding_line, 4 * (size_t)nfreqs));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->m.nfreqs = nfreqs; ctx->m.nu = ctx->nu.as<double>(); ctx->m.corr_line = ctx->corr_line.as<uint32_t>();
    MGB_TRY(build_freq_chunks(ctx));
    ctx->have_freqs = true;
    return MGB_OK;
}

extern "C" int mgb_set_line_state(mgb_ctx* ctx, const double* emissivity, const double* opacity, const double* inverse_width,
                                  const double* population_tot) {
    if (!ctx) return MGB_EINVAL;
    if (!ctx->have_geometry || !ctx->have_lines) return ctx->fail(MGB_ESTATE, "mgb_set_line_state before geometry/lines");
    if (!emissivity || !opacity || !inverse_width || !population_tot) return ctx->fail(MGB_EINVAL, "mgb_set_line_state: null argument");
    cudaSetDevice(ctx->device);
    const DevModel& m = ctx->m;
    const size_t n    = (size_t)m.npoints * m.nlines;
    MGB_TRY(upload(ctx, ctx->stage_a, emissivity, 8 * n));
    MGB_TRY(upload(ctx, ctx->stage_b, opacity, 8 * n));
    MGB_TRY(upload(ctx, ctx->stage_c, inverse_width, 8 * n));
    MGB_TRY(upload(ctx, ctx->stage_d, population_tot, 8 * (size_t)m.nlspecs * m.npoints));
    MGB_CUDA(ctx, ctx->lrow.ensure(sizeof(LineRow) * n));
    k_prepare_rows<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(m.npoints, m.nlines, ctx->stage_a.as<double>(),
                                                                         ctx->stage_b.as<double>(), ctx->stage_c.as<double>(),
                                                                         ctx->stage_d.as<double>(), ctx->line_spec_dev.as<uint32_t>(),
                                                                         ctx->lrow.as<LineRow>());
    MGB_CUDA(ctx, cudaGetLastError());
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->m.lrow     = ctx->lrow.as<LineRow>();
    ctx->have_state = true;
    // the sub-step counts depend on the criterion, the geometric paths do not: cached paths stay valid
    return MGB_OK;
}

// ---------------------------------------------------------------------------------------------------
// The run
// ---------------------------------------------------------------------------------------------------
namespace {

struct EventTimer {
    cudaEvent_t a = nullptr, b = nullptr;
    cudaStream_t s;
    explicit EventTimer(cudaStream_t st) : s(st) { cudaEventCreate(&a); cudaEventCreate(&b); }
    ~EventTimer() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
    void start() { cudaEventRecord(a, s); }
    float stop() { /* synchronising: the run is synchronous anyway */
        cudaEventRecord(b, s);
        cudaEventSynchronize(b);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, a, b);
        return ms;
    }
};

int refresh_points(mgb_ctx* ctx) {
    if (!ctx->points_dirty) return MGB_OK;
    const DevModel& m = ctx->m;
    k_prepare_points<<<(m.npoints + 255) / 256, 256, 0, ctx->stream>>>(m.npoints, m.nlspecs, m.temperature, m.vturb2,
                                                                      ctx->inv_mass_dev.as<double>(), ctx->max_inverse_mass,
                                                                      ctx->params.max_width_fraction, ctx->sqrtw_max.as<double>(),
                                                                      ctx->dshift_max.as<double>());
    MGB_CUDA(ctx, cudaGetLastError());
    ctx->points_dirty = false;
    return MGB_OK;
}

// Count pass: geometric length of every half-ray of origins [o0,o1) (once per geometry and range).
int ensure_lens(mgb_ctx* ctx, uint32_t o0, uint32_t o1, mgb_stats* st, uint32_t& launches) {
    const DevModel& m = ctx->m;
    if (ctx->lens_o1 > ctx->lens_o0 && o0 >= ctx->lens_o0 && o1 <= ctx->lens_o1) return MGB_OK;
    MGB_CUDA(ctx, ctx->lens.ensure(4 * (size_t)m.npoints * m.nrays));
    MGB_CUDA(ctx, ctx->counters.ensure(8 * 8));
    MGB_CUDA(ctx, cudaMemsetAsync(ctx->counters.p, 0, 64, ctx->stream));
    const uint32_t no = o1 - o0;
    const size_t nw   = (size_t)no * m.nrays;
    NvtxRange nvtx("mgb: ray lengths (k_trace count pass)");
    EventTimer t(ctx->stream);
    t.start();
    PathView pv{};
    k_trace<false><<<(unsigned)((nw * 32 + 127) / 128), 128, 0, ctx->stream>>>(m, o0, no, ctx->lens.as<uint32_t>(), pv,
                                                                               ctx->counters.as<unsigned long long>());
    MGB_CUDA(ctx, cudaGetLastError());
    MGB_CUDA(ctx, ctx->origin_tot.ensure(8 * (size_t)no));
    k_origin_totals<<<(no + 255) / 256, 256, 0, ctx->stream>>>(o0, no, m.nrays, ctx->lens.as<uint32_t>(),
                                                               ctx->origin_tot.as<unsigned long long>());
    MGB_CUDA(ctx, cudaGetLastError());
    launches += 2;
    const float ms = t.stop();
    if (st) st->ms_trace_count += ms;
    ctx->origin_steps.resize(no);
    unsigned long long c[3];
    MGB_CUDA(ctx, cudaMemcpy(ctx->origin_steps.data(), ctx->origin_tot.p, 8 * (size_t)no, cudaMemcpyDeviceToHost));
    MGB_CUDA(ctx, cudaMemcpy(c, ctx->counters.p, 24, cudaMemcpyDeviceToHost));
    ctx->tot_steps = c[0]; ctx->tot_scanned = c[1]; ctx->max_len = (uint32_t)c[2];
    ctx->lens_o0 = o0; ctx->lens_o1 = o1;
    for (auto* c : ctx->chunks) c->filled = false; /* buffers stay: cudaFree/cudaMalloc of GB-sized buffers costs ~0.3 s once NCCL has
                                                      peer mappings live (profiles/comm_diag_r2.log) */
    ctx->plan_cached = false;
    return MGB_OK;
}

int fill_chunk(mgb_ctx* ctx, PathChunk& ch, mgb_stats* st, uint32_t& launches) {
    const DevModel& m = ctx->m;
    const uint32_t no = ch.o1 - ch.o0;
    const size_t nhr  = (size_t)no * m.nrays;
    MGB_CUDA(ctx, ch.poff.ensure(8 * nhr));
    MGB_CUDA(ctx, ch.shift0.ensure(8 * nhr));
    MGB_CUDA(ctx, ch.nr.ensure(4 * (size_t)ch.steps));
    MGB_CUDA(ctx, ch.dZ.ensure(8 * (size_t)ch.steps));
    MGB_CUDA(ctx, ch.shift.ensure(8 * (size_t)ch.steps));
    NvtxRange nvtx("mgb: ray paths (k_trace fill pass)");
    EventTimer t(ctx->stream);
    t.start();
    const uint32_t* lens_chunk = ctx->lens.as<uint32_t>() + (size_t)ch.o0 * m.nrays;
    size_t tmp = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp, lens_chunk, ch.poff.as<unsigned long long>(), (int64_t)nhr, ctx->stream);
    MGB_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
    MGB_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->cub_tmp.p, tmp, lens_chunk, ch.poff.as<unsigned long long>(), (int64_t)nhr,
                                                ctx->stream));
    PathView pv{};
    pv.o0 = ch.o0; pv.no = no; pv.len = ctx->lens.as<uint32_t>(); pv.poff = ch.poff.as<unsigned long long>();
    pv.nr = ch.nr.as<uint32_t>(); pv.dZ = ch.dZ.as<double>(); pv.shift = ch.shift.as<double>(); pv.shift0 = ch.shift0.as<double>();
    k_trace<true><<<(unsigned)((nhr * 32 + 127) / 128), 128, 0, ctx->stream>>>(m, ch.o0, no, nullptr, pv, nullptr);
    MGB_CUDA(ctx, cudaGetLastError());
    launches += 2;
    const float ms = t.stop();
    if (st) st->ms_trace_fill += ms;
    ch.filled = true;
    return MGB_OK;
}

template <int APPROX, bool OFFD>
cudaError_t launch_solve2(mgb_ctx* ctx, const PathView& pv, const ScratchView& sc, uint32_t nchunks_f) {
    const size_t nitems = (size_t)pv.no * ctx->m.hnrays * nchunks_f;
    const size_t blocks = (nitems + kSolveWarps - 1) / kSolveWarps;
    const size_t smem   = kSolveWarps * solve_smem_per_warp(ctx->lines_per_warp);
    cudaError_t e = cudaFuncSetAttribute(k_solve<APPROX, OFFD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_solve<APPROX, OFFD><<<(unsigned)blocks, 32 * kSolveWarps, smem, ctx->stream>>>(ctx->m, pv, sc, ctx->freq_chunks.as<FreqChunk>(), nchunks_f,
                                                                        ctx->lines_per_warp);
    return cudaGetLastError();
}
// n_off_diag = 0: the solve kernel leaves L_diag in sc.L and k_reduce_lambda2 applies the origin's factors. MGB_SOLVE_V2=1 in the environment
// selects k_solve2 (mgb_solve2.cuh: two bins per lane, straight-line common step; bit-identical, measured SLOWER on B200 -- see DESIGN.md 6).
template <int APPROX>
cudaError_t launch_solve_v2(mgb_ctx* ctx, const PathView& pv, const ScratchView& sc) {
    const size_t nitems = (size_t)pv.no * ctx->m.hnrays * ctx->n_freq_chunks2;
    const size_t blocks = (nitems + kSolve2Warps - 1) / kSolve2Warps;
    const size_t smem   = kSolve2Warps * solve2_smem_per_warp(ctx->lines_per_warp2);
    cudaError_t e = cudaFuncSetAttribute(k_solve2<APPROX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_solve2<APPROX><<<(unsigned)blocks, 32 * kSolve2Warps, smem, ctx->stream>>>(ctx->m, pv, sc, ctx->freq_chunks2.as<FreqChunk2>(), ctx->n_freq_chunks2,
                                                                               ctx->lines_per_warp2);
    return cudaGetLastError();
}
// MGB_SOLVE_KERNEL=3: k_solve3 (mgb_solve3.cuh: k_solve's work items, straight-line common step, range-free erff)
template <int APPROX>
cudaError_t launch_solve_v3(mgb_ctx* ctx, const PathView& pv, const ScratchView& sc, uint32_t nchunks_f) {
    const size_t nitems = (size_t)pv.no * ctx->m.hnrays * nchunks_f;
    const size_t blocks = (nitems + kSolve3Warps - 1) / kSolve3Warps;
    const size_t smem   = kSolve3Warps * solve_smem_per_warp(ctx->lines_per_warp);
    cudaError_t e = cudaFuncSetAttribute(k_solve3<APPROX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_solve3<APPROX><<<(unsigned)blocks, 32 * kSolve3Warps, smem, ctx->stream>>>(ctx->m, pv, sc, ctx->freq_chunks.as<FreqChunk>(), nchunks_f,
                                                                               ctx->lines_per_warp);
    return cudaGetLastError();
}
// MGB_SOLVE_KERNEL=4: the streaming pipeline of mgb_solve4.cuh on one path chunk. The chunk's origins are taken in sub-ranges whose
// records fit the record budget (MGB_REC_BYTES, default 40 GB for the 80-byte records; the line integrals of the large-shift
// segments come on top, 24 * nquads bytes each).
struct CastU64 {
    __host__ __device__ unsigned long long operator()(uint32_t v) const { return v; }
};
template <int APPROX>
int run_pipeline_v4(mgb_ctx* ctx, const PathView& pv, const ScratchView& sc, const PathChunk* ch, uint32_t nchunks_f, uint32_t& launches) {
    const DevModel& m = ctx->m;
    if (!ctx->side_ok) return ctx->fail(MGB_EUNSUPPORTED, "MGB_SOLVE_KERNEL=4 needs antipod[rr] >= nrays/2 for every rr < nrays/2");
    const unsigned long long budget = [] { /* read per run: tests shrink it to force many record ranges */
        const char* e = getenv("MGB_REC_BYTES");
        const unsigned long long v = e ? strtoull(e, nullptr, 10) : 0ull;
        return v ? v : (40ull << 30);
    }();
    const uint32_t no = pv.no;
    const size_t lists_per_o = (size_t)m.nrays * m.nlines;
    auto est_bytes = [&](uint32_t ol) { /* records of one origin: steps x lines (x 1.3 for sub-steps) + one origin record per list */
        const unsigned long long so = ctx->origin_steps[ch->o0 + ol - ctx->lens_o0];
        return (unsigned long long)(96.0 * 1.3 * (double)so * m.nlines) + 96ull * lists_per_o;
    };
    MGB_CUDA(ctx, ctx->rec_lcount.ensure(8));
    uint32_t a = 0;
    while (a < no) {
        uint32_t b = a;
        unsigned long long est = 0;
        while (b < no && (b == a || est + est_bytes(b) <= budget)) { est += est_bytes(b); b++; }
        unsigned long long nrec = 0, nelem = 0, nlarge = 0;
        RecView rv{};
        DenseView dv{};
        for (;;) { /* elements and records of the range; halve the range when the records would not fit */
            const uint32_t nhr = (b - a) * m.nrays;
            MGB_CUDA(ctx, ctx->rec_cnt.ensure(4 * ((size_t)nhr + 1)));
            MGB_CUDA(ctx, ctx->rec_eoff.ensure(8 * ((size_t)nhr + 1)));
            EventTimer t0(ctx->stream);
            t0.start();
            k_ecount<<<(nhr + 1 + 255) / 256, 256, 0, ctx->stream>>>(nhr, m.nrays, m.nlines, pv.o0 + a, pv.len, ctx->rec_cnt.as<uint32_t>());
            MGB_CUDA(ctx, cudaGetLastError());
            size_t tmp = 0;
            {
                cub::TransformInputIterator<unsigned long long, CastU64, const uint32_t*> in(ctx->rec_cnt.as<uint32_t>(), CastU64());
                cub::DeviceScan::ExclusiveSum(nullptr, tmp, in, ctx->rec_eoff.as<unsigned long long>(), (int64_t)nhr + 1, ctx->stream);
                MGB_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
                MGB_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->cub_tmp.p, tmp, in, ctx->rec_eoff.as<unsigned long long>(), (int64_t)nhr + 1, ctx->stream));
            }
            MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); /* the context's stream does not synchronise with the default stream's copy */
            MGB_CUDA(ctx, cudaMemcpy(&nelem, ctx->rec_eoff.as<unsigned long long>() + nhr, 8, cudaMemcpyDeviceToHost));
            MGB_CUDA(ctx, ctx->rec_ni.ensure(4 * ((size_t)nelem + 1)));
            MGB_CUDA(ctx, ctx->rec_base.ensure(8 * ((size_t)nelem + 1)));
            MGB_CUDA(ctx, cudaMemsetAsync(ctx->rec_ni.as<uint32_t>() + nelem, 0, 4, ctx->stream));
            dv.eoff = ctx->rec_eoff.as<unsigned long long>(); dv.ni = ctx->rec_ni.as<uint32_t>(); dv.base = ctx->rec_base.as<unsigned long long>();
            k_ni<<<(unsigned)(((size_t)nhr * 32 + 127) / 128), 128, 0, ctx->stream>>>(m, pv, a, b - a, dv);
            MGB_CUDA(ctx, cudaGetLastError());
            {
                cub::TransformInputIterator<unsigned long long, CastU64, const uint32_t*> in(ctx->rec_ni.as<uint32_t>(), CastU64());
                cub::DeviceScan::ExclusiveSum(nullptr, tmp, in, ctx->rec_base.as<unsigned long long>(), (int64_t)nelem + 1, ctx->stream);
                MGB_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
                MGB_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->cub_tmp.p, tmp, in, ctx->rec_base.as<unsigned long long>(), (int64_t)nelem + 1, ctx->stream));
            }
            ctx->ms_v4[0] += t0.stop();
            launches += 4;
            MGB_CUDA(ctx, cudaMemcpy(&nrec, ctx->rec_base.as<unsigned long long>() + nelem, 8, cudaMemcpyDeviceToHost));
            if (nrec * 96ull > budget + budget / 4 && b - a > 1) { b = a + (b - a) / 2; continue; }
            const size_t nlists = (size_t)(b - a) * lists_per_o;
            MGB_CUDA(ctx, ctx->rec_off.ensure(8 * (nlists + 1)));
            MGB_CUDA(ctx, ctx->rec.ensure(64 * (size_t)nrec));
            MGB_CUDA(ctx, ctx->rec_pts.ensure(16 * (size_t)nrec));
            rv.ola = a; rv.nob = b - a; rv.off = ctx->rec_off.as<unsigned long long>(); rv.lcount = ctx->rec_lcount.as<unsigned long long>();
            rv.rec = ctx->rec.as<double>(); rv.pts = ctx->rec_pts.as<double2>();
            MGB_CUDA(ctx, ctx->rec_slots.ensure(sizeof(SlotInfo) * (size_t)nrec)); /* capacity: every record could end a large-shift segment */
            rv.slots = ctx->rec_slots.as<SlotInfo>(); rv.side_of = ctx->side_of.as<uint8_t>();
            {
                EventTimer t1(ctx->stream);
                t1.start();
                MGB_CUDA(ctx, cudaMemsetAsync(ctx->rec_lcount.p, 0, 8, ctx->stream));
                k_off<<<(unsigned)((nlists + 1 + 255) / 256), 256, 0, ctx->stream>>>(m, pv, a, b - a, dv, nelem, ctx->rec_off.as<unsigned long long>());
                MGB_CUDA(ctx, cudaGetLastError());
                k_fill<<<(unsigned)(((size_t)(b - a) * m.nrays * 32 + 127) / 128), 128, 0, ctx->stream>>>(m, pv, rv, dv);
                MGB_CUDA(ctx, cudaGetLastError());
                ctx->ms_v4[0] += t1.stop();
                launches += 2;
                MGB_CUDA(ctx, cudaMemcpy(&nlarge, ctx->rec_lcount.p, 8, cudaMemcpyDeviceToHost));
            }
            /* the line integrals of the flagged segments (24 * nquads bytes each) have the same budget: a range in which most segments
               take the large-shift branch is halved and redone */
            if (nlarge * 24ull * m.nquads > budget + budget / 4 && b - a > 1) { b = a + (b - a) / 2; continue; }
            break;
        }
        const size_t nlists = (size_t)(b - a) * lists_per_o;
        {
            EventTimer t2(ctx->stream);
            t2.start();
            if (nlarge) {
                MGB_CUDA(ctx, ctx->rec_trip.ensure(24 * (size_t)m.nquads * (size_t)nlarge));
                const unsigned spb = 128 / m.nquads; /* slots per block */
                if (!spb) return ctx->fail(MGB_EUNSUPPORTED, "MGB_SOLVE_KERNEL=4: nquads > 128");
                k_large<APPROX><<<(unsigned)((nlarge + spb - 1) / spb), 128, 0, ctx->stream>>>(m, rv.rec, rv.pts, ctx->rec_slots.as<SlotInfo>(), nlarge, ctx->rec_trip.as<double>());
                MGB_CUDA(ctx, cudaGetLastError());
                launches += 1;
            }
            ctx->ms_v4[1] += t2.stop();
        }
        {
            EventTimer t3(ctx->stream);
            t3.start();
            const size_t nitems = (size_t)(b - a) * m.hnrays * nchunks_f;
            k_sweep<APPROX><<<(unsigned)((nitems + kSweepWarps - 1) / kSweepWarps), 32 * kSweepWarps, 0, ctx->stream>>>(
                m, pv, sc, ctx->freq_chunks.as<FreqChunk>(), nchunks_f, rv, ctx->rec_trip.as<double>());
            MGB_CUDA(ctx, cudaGetLastError());
            ctx->ms_v4[2] += t3.stop();
            ctx->ms_v4[3] += 1.f; /* record ranges (= k_sweep launches) of the run */
            launches++;
        }
        a = b;
    }
    return MGB_OK;
}

template <int APPROX>
cudaError_t launch_solve(mgb_ctx* ctx, const PathView& pv, const ScratchView& sc, uint32_t nchunks_f) {
    if (sc.K) return launch_solve2<APPROX, true>(ctx, pv, sc, nchunks_f);
    if (ctx->solve_v2 == 3) return launch_solve_v3<APPROX>(ctx, pv, sc, nchunks_f);
    return ctx->solve_v2 == 2 ? launch_solve_v2<APPROX>(ctx, pv, sc) : launch_solve2<APPROX, false>(ctx, pv, sc, nchunks_f);
}
template <int APPROX>
cudaError_t launch_shortchar(mgb_ctx* ctx, const PathView& pv, const ScratchView& sc, uint32_t nchunks_f, double min_dtau) {
    const size_t nitems = (size_t)pv.no * ctx->m.hnrays * nchunks_f;
    const size_t blocks = (nitems + kSolveWarps - 1) / kSolveWarps;
    const size_t smem   = kSolveWarps * solve_smem_per_warp(ctx->lines_per_warp);
    cudaError_t e = cudaFuncSetAttribute(k_shortchar<APPROX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_shortchar<APPROX><<<(unsigned)blocks, 32 * kSolveWarps, smem, ctx->stream>>>(ctx->m, pv, sc, ctx->freq_chunks.as<FreqChunk>(), nchunks_f,
                                                                                ctx->lines_per_warp, min_dtau);
    return cudaGetLastError();
}
template <int APPROX>
cudaError_t launch_reduce_lambda2(mgb_ctx* ctx, uint32_t o0, uint32_t no, const ScratchView& sc) {
    const size_t nl = (size_t)no * ctx->m.nlines;
    k_reduce_lambda2<APPROX><<<(unsigned)((nl + 127) / 128), 128, 0, ctx->stream>>>(ctx->m, o0, no, ctx->lens.as<uint32_t>(), sc.L, ctx->lam_val.as<double>(),
                                                                                   ctx->lam_cnt.as<uint32_t>(), ctx->scratch_fac.as<double>(),
                                                                                   ctx->scratch_fac.as<double>() + (size_t)no * ctx->m.nfreqs);
    return cudaGetLastError();
}

} // namespace

// shared with mgb_image.cu: per-point derived tables (sqrtw_max, Model::set_dshift_max) up to date
int mgb_internal_refresh_points(mgb_ctx* ctx) { return refresh_points(ctx); }

// Result arrays [npoints][...] of the Feautrier run (J, J per line, diagonal Lambda element and list length).
int mgb_internal_alloc_results(mgb_ctx* ctx) {
    const DevModel& m = ctx->m;
    const size_t prow = m.npoints;
    MGB_CUDA(ctx, ctx->J.ensure(8 * prow * m.nfreqs));
    MGB_CUDA(ctx, ctx->lam_val.ensure(8 * prow * m.nlines));
    MGB_CUDA(ctx, ctx->J_lines.ensure(8 * prow * m.nlines));
    MGB_CUDA(ctx, ctx->lam_cnt.ensure(4 * prow * m.nlines));
    return MGB_OK;
}

// Contiguous origin ranges of equal predicted cost: cost(o) = sum over rays of the geometric ray length (the count pass), plus one
// entry per ray for the origin itself. Every rank computes the same bounds from the same replicated geometry.
extern "C" int mgb_partition_balanced(mgb_ctx* ctx, int nranks, int rank, uint32_t* o_begin, uint32_t* o_end) {
    if (!ctx || !o_begin || !o_end || nranks < 1 || rank < 0 || rank >= nranks) return MGB_EINVAL;
    if (!(ctx->have_geometry && ctx->have_thermo && ctx->have_lines)) return ctx->fail(MGB_ESTATE, "mgb_partition_balanced: model incomplete");
    cudaSetDevice(ctx->device);
    const uint32_t N = ctx->m.npoints;
    uint32_t launches = 0;
    MGB_TRY(refresh_points(ctx));
    MGB_TRY(ensure_lens(ctx, 0, N, nullptr, launches));
    std::vector<unsigned long long> pre(N + 1, 0);
    for (uint32_t o = 0; o < N; o++) pre[o + 1] = pre[o] + ctx->origin_steps[o] + ctx->m.nrays;
    ctx->part_bounds.assign(nranks + 1, N);
    ctx->part_bounds[0] = 0;
    for (int r = 1; r < nranks; r++) {
        const unsigned long long target = pre[N] / nranks * r;
        ctx->part_bounds[r] = (uint32_t)(std::lower_bound(pre.begin(), pre.end(), target) - pre.begin());
        if (ctx->part_bounds[r] < ctx->part_bounds[r - 1]) ctx->part_bounds[r] = ctx->part_bounds[r - 1];
        if (ctx->part_bounds[r] > N) ctx->part_bounds[r] = N;
    }
    *o_begin = ctx->part_bounds[rank]; *o_end = ctx->part_bounds[rank + 1];
    return MGB_OK;
}

extern "C" int mgb_invalidate_paths(mgb_ctx* ctx) {
    if (!ctx) return MGB_EINVAL;
    cudaSetDevice(ctx->device);
    for (auto* c : ctx->chunks) c->filled = false;
    ctx->work.filled = false;
    ctx->lens_o0 = ctx->lens_o1 = 0;
    ctx->plan_cached = false; /* the chunk objects keep their device buffers for the next plan */
    return MGB_OK;
}

extern "C" int mgb_timer_start(mgb_ctx* ctx) {
    if (!ctx) return MGB_EINVAL;
    cudaSetDevice(ctx->device);
    if (!ctx->ev_a) { MGB_CUDA(ctx, cudaEventCreate(&ctx->ev_a)); MGB_CUDA(ctx, cudaEventCreate(&ctx->ev_b)); }
    MGB_CUDA(ctx, cudaEventRecord(ctx->ev_a, ctx->stream));
    return MGB_OK;
}

extern "C" int mgb_timer_stop(mgb_ctx* ctx, float* ms) {
    if (!ctx || !ms) return MGB_EINVAL;
    if (!ctx->ev_a) return ctx->fail(MGB_ESTATE, "mgb_timer_stop without mgb_timer_start");
    cudaSetDevice(ctx->device);
    MGB_CUDA(ctx, cudaEventRecord(ctx->ev_b, ctx->stream));
    MGB_CUDA(ctx, cudaEventSynchronize(ctx->ev_b));
    MGB_CUDA(ctx, cudaEventElapsedTime(ms, ctx->ev_a, ctx->ev_b));
    return MGB_OK;
}

// The radiation-field run of both solvers: `shortchar` = false is Solver::solve_feautrier_order_2 (solver.tpp:458-540), true is
// Solver::solve_shortchar_order_0 (solver.tpp:150-235) on the same traced paths, origin chunks and J / Lambda result arrays.
static int run_radiation_field(mgb_ctx* ctx, int approx, uint32_t o_begin, uint32_t o_end, mgb_stats* stats, bool shortchar, double min_dtau) {
    const char* who = shortchar ? "mgb_run_shortchar_order_0" : "mgb_run_feautrier_order_2";
    if (!ctx) return MGB_EINVAL;
    if (!(ctx->have_geometry && ctx->have_thermo && ctx->have_lines && ctx->have_freqs && ctx->have_state))
        return ctx->fail(MGB_ESTATE, "%s: model incomplete (geometry %d thermo %d lines %d freqs %d state %d)", who,
                         ctx->have_geometry, ctx->have_thermo, ctx->have_lines, ctx->have_freqs, ctx->have_state);
    if (shortchar && !(min_dtau > 0.0)) return ctx->fail(MGB_EINVAL, "%s: min_dtau must be positive (Parameters::min_dtau = 1e-15)", who);
    if (approx < 0 || approx > 2) return ctx->fail(MGB_EINVAL, "approx must be 0 (None), 1 (OneLine) or 2 (CloseLines)");
    DevModel& m = ctx->m;
    if (o_begin > o_end || o_end > m.npoints) return ctx->fail(MGB_EINVAL, "origin range [%u,%u) invalid for npoints = %u", o_begin, o_end, m.npoints);
    cudaSetDevice(ctx->device);
    mgb_stats st;
    memset(&st, 0, sizeof st);
    if (o_begin == o_end) { /* a rank that owns no origin (nranks > npoints, or a very skewed cost partition): a successful no-op that still
                               owns result buffers, so that the peers' collective in mgb_allgather_results does not hang on it */
        MGB_TRY(mgb_internal_alloc_results(ctx));
        ctx->res_o0 = ctx->res_o1 = o_begin; ctx->res_K = shortchar ? 0 : ctx->params.n_off_diag; ctx->res_valid = true;
        if (stats) *stats = st;
        return MGB_OK;
    }
    uint32_t launches = 0;
    NvtxRange nvtx_run(who);
    EventTimer total(ctx->stream);
    total.start();
    MGB_TRY(refresh_points(ctx));
    MGB_TRY(ensure_lens(ctx, o_begin, o_end, &st, launches));

    const uint32_t nchunks_f = ctx->n_freq_chunks;
    // ---- plan origin chunks: bounded scratch (16 B per solve) and bounded path bytes (20 B per step) ----
    const uint32_t K = shortchar ? 0 : ctx->params.n_off_diag; /* the short-characteristics solver only has the local Lambda element */
    const size_t per_origin_scratch = (size_t)m.hnrays * m.nfreqs * (shortchar ? 24 : 16 + 24 * (size_t)K);
    const uint32_t max_o_scratch    = (uint32_t)std::max<size_t>(1, ctx->params.scratch_bytes / per_origin_scratch);
    unsigned long long range_steps  = 0;
    for (uint32_t o = o_begin; o < o_end; o++) range_steps += ctx->origin_steps[o - ctx->lens_o0];
    const bool cacheable = ctx->params.cache_paths && range_steps * 20ull + (size_t)(o_end - o_begin) * m.nrays * 16ull <= ctx->params.path_cache_bytes;
    const unsigned long long chunk_step_budget = cacheable ? ~0ull : std::max<unsigned long long>(1ull << 20, ctx->params.path_cache_bytes / 20ull);
    if (!(ctx->plan_cached && ctx->plan_o0 == o_begin && ctx->plan_o1 == o_end && cacheable)) {
        struct Plan { uint32_t o0, o1; unsigned long long steps; };
        std::vector<Plan> plan;
        uint32_t o = o_begin;
        while (o < o_end) {
            Plan pl{o, o, 0};
            while (o < o_end && (o - pl.o0) < max_o_scratch) {
                const unsigned long long so = ctx->origin_steps[o - ctx->lens_o0];
                if (o > pl.o0 && pl.steps + so > chunk_step_budget) break;
                pl.steps += so; o++;
            }
            pl.o1 = o;
            plan.push_back(pl);
        }
        bool same = plan.size() == ctx->chunks.size();
        for (size_t i = 0; same && i < plan.size(); i++)
            same = ctx->chunks[i]->o0 == plan[i].o0 && ctx->chunks[i]->o1 == plan[i].o1 && ctx->chunks[i]->steps == plan[i].steps;
        if (!same) { /* a different decomposition: new chunk objects (an unchanged one keeps its device buffers and its paths) */
            ctx->drop_chunks();
            for (const Plan& pl : plan) {
                PathChunk* ch = new PathChunk();
                ch->o0 = pl.o0; ch->o1 = pl.o1; ch->steps = pl.steps;
                ctx->chunks.push_back(ch);
            }
        }
        ctx->plan_o0 = o_begin; ctx->plan_o1 = o_end; ctx->plan_cached = cacheable;
    }
    // ---- outputs: full-size arrays on every rank, each rank fills the rows of its origins (mgb_allgather_results completes them) ----
    const size_t prow = m.npoints;
    MGB_TRY(mgb_internal_alloc_results(ctx));
    double* u_full = nullptr;
    if (ctx->params.store_intensities) {
        MGB_CUDA(ctx, ctx->u_full.ensure(8 * (size_t)m.hnrays * (o_end - o_begin) * m.nfreqs));
        u_full = ctx->u_full.as<double>(); ctx->u_o0 = o_begin; ctx->u_no = o_end - o_begin;
    }
    const bool anis = ctx->params.anisotropic && !shortchar;
    if (anis) MGB_CUDA(ctx, ctx->J_anis.ensure(8 * 6 * prow * m.nlines));
    uint32_t max_no = 0;
    for (auto* c : ctx->chunks) max_no = std::max(max_no, c->o1 - c->o0);
    MGB_CUDA(ctx, ctx->scratch_u.ensure(8 * (size_t)m.hnrays * max_no * m.nfreqs));
    MGB_CUDA(ctx, ctx->scratch_L.ensure(8 * (size_t)m.hnrays * max_no * m.nfreqs));
    MGB_CUDA(ctx, ctx->counters.ensure(64));
    MGB_CUDA(ctx, cudaMemsetAsync(ctx->counters.p, 0, 64, ctx->stream));
    {
        const char* v2 = getenv("MGB_SOLVE_V2");
        const char* vk = getenv("MGB_SOLVE_KERNEL"); /* 1: k_solve, 2: k_solve2, 3: k_solve3, 4: the streaming pipeline (mgb_solve4.cuh) */
        ctx->ms_v4[0] = ctx->ms_v4[1] = ctx->ms_v4[2] = ctx->ms_v4[3] = 0.f;
        ctx->solve_v2 = (vk && vk[0] >= '1' && vk[0] <= '4') ? vk[0] - '0' : ((v2 && v2[0] == '1') ? 2 : MGB_DEFAULT_SOLVE_KERNEL);
    }
    const bool lambda2 = !K && !shortchar; /* the diagonal operator: L_diag from the solve kernel, factors applied by k_reduce_lambda2 */
    if (shortchar) MGB_CUDA(ctx, ctx->scratch_offV.ensure(16 * (size_t)m.hnrays * max_no * m.nfreqs)); /* the two half-ray factors of each (rr, o, f) */
    if (lambda2) MGB_CUDA(ctx, ctx->scratch_fac.ensure(16 * (size_t)max_no * m.nfreqs));
    if (K) { /* banded operator: at most 1 + 2K distinct points per ray pair (lambda.tpp:79-89 lists) */
        ctx->lam_cap = 1 + 2 * K * m.hnrays;
        MGB_CUDA(ctx, ctx->lam_list_val.ensure(8 * prow * m.nlines * ctx->lam_cap));
        MGB_CUDA(ctx, ctx->lam_list_nr.ensure(4 * prow * m.nlines * ctx->lam_cap));
        MGB_CUDA(ctx, ctx->scratch_offV.ensure(8 * (size_t)m.hnrays * max_no * m.nfreqs * 2 * K));
        MGB_CUDA(ctx, ctx->scratch_offN.ensure(4 * (size_t)m.hnrays * max_no * m.nfreqs * 2 * K));
    } else ctx->lam_cap = 1;

    bool all_cached = true;
    for (auto* cp : ctx->chunks) {
        PathChunk* ch = cp;
        if (!cacheable) { /* reuse one workspace for every chunk */
            ctx->work.o0 = cp->o0; ctx->work.o1 = cp->o1; ctx->work.steps = cp->steps; ctx->work.filled = false;
            ch = &ctx->work;
        }
        if (!ch->filled) { all_cached = false; MGB_TRY(fill_chunk(ctx, *ch, &st, launches)); }
        const uint32_t no = ch->o1 - ch->o0;
        PathView pv{};
        pv.o0 = ch->o0; pv.no = no; pv.len = ctx->lens.as<uint32_t>(); pv.poff = ch->poff.as<unsigned long long>();
        pv.nr = ch->nr.as<uint32_t>(); pv.dZ = ch->dZ.as<double>(); pv.shift = ch->shift.as<double>(); pv.shift0 = ch->shift0.as<double>();
        ScratchView sc{ctx->scratch_u.as<double>(), ctx->scratch_L.as<double>(), ctx->counters.as<unsigned long long>(),
                       ctx->scratch_offV.as<double>(), ctx->scratch_offN.as<uint32_t>(), K};
        {
        NvtxRange nvtx_solve("mgb: k_solve (eta/chi + Feautrier sweep)");
        EventTimer ts(ctx->stream);
        ts.start();
        if (!shortchar && !K && ctx->solve_v2 == 4) {
            if (approx == APPROX_NONE) MGB_TRY(run_pipeline_v4<APPROX_NONE>(ctx, pv, sc, ch, nchunks_f, launches));
            else if (approx == APPROX_ONELINE) MGB_TRY(run_pipeline_v4<APPROX_ONELINE>(ctx, pv, sc, ch, nchunks_f, launches));
            else MGB_TRY(run_pipeline_v4<APPROX_CLOSELINES>(ctx, pv, sc, ch, nchunks_f, launches));
        } else if (shortchar) {
            if (approx == APPROX_NONE) MGB_CUDA(ctx, launch_shortchar<APPROX_NONE>(ctx, pv, sc, nchunks_f, min_dtau));
            else if (approx == APPROX_ONELINE) MGB_CUDA(ctx, launch_shortchar<APPROX_ONELINE>(ctx, pv, sc, nchunks_f, min_dtau));
            else MGB_CUDA(ctx, launch_shortchar<APPROX_CLOSELINES>(ctx, pv, sc, nchunks_f, min_dtau));
        } else if (approx == APPROX_NONE) MGB_CUDA(ctx, launch_solve<APPROX_NONE>(ctx, pv, sc, nchunks_f));
        else if (approx == APPROX_ONELINE) MGB_CUDA(ctx, launch_solve<APPROX_ONELINE>(ctx, pv, sc, nchunks_f));
        else MGB_CUDA(ctx, launch_solve<APPROX_CLOSELINES>(ctx, pv, sc, nchunks_f));
        st.ms_solve += ts.stop();
        }
        NvtxRange nvtx_reduce("mgb: J / Lambda reduction");
        st.n_solve_launches++;
        launches++;
        EventTimer tr(ctx->stream);
        tr.start();
        const size_t nj = (size_t)no * m.nfreqs, nl = (size_t)no * m.nlines;
        k_reduce_J<<<(unsigned)((nj + 255) / 256), 256, 0, ctx->stream>>>(m, ch->o0, no, sc.u, ctx->J.as<double>(), u_full, ctx->u_o0, ctx->u_no);
        k_reduce_J_lines<<<(unsigned)((nl + 255) / 256), 256, 0, ctx->stream>>>(m, ch->o0, no, sc.u, ctx->J_lines.as<double>());
        if (anis) { k_reduce_anis<<<(unsigned)((nl + 255) / 256), 256, 0, ctx->stream>>>(m, ch->o0, no, sc.u, ctx->J_anis.as<double>()); launches++; }
        if (shortchar) {
            const unsigned nb = (unsigned)((nl + 127) / 128);
            if (approx == APPROX_NONE) k_reduce_lambda_sc<APPROX_NONE><<<nb, 128, 0, ctx->stream>>>(m, ch->o0, no, ctx->lens.as<uint32_t>(), sc.offV, ctx->lam_val.as<double>(), ctx->lam_cnt.as<uint32_t>());
            else if (approx == APPROX_ONELINE) k_reduce_lambda_sc<APPROX_ONELINE><<<nb, 128, 0, ctx->stream>>>(m, ch->o0, no, ctx->lens.as<uint32_t>(), sc.offV, ctx->lam_val.as<double>(), ctx->lam_cnt.as<uint32_t>());
            else k_reduce_lambda_sc<APPROX_CLOSELINES><<<nb, 128, 0, ctx->stream>>>(m, ch->o0, no, ctx->lens.as<uint32_t>(), sc.offV, ctx->lam_val.as<double>(), ctx->lam_cnt.as<uint32_t>());
        } else if (K)
            k_reduce_lambda_banded<<<(unsigned)((nl + 127) / 128), 128, 0, ctx->stream>>>(
                m, ch->o0, no, K, ctx->lam_cap, ctx->lens.as<uint32_t>(), sc.L, sc.offV, sc.offN, ctx->lam_list_val.as<double>(),
                ctx->lam_list_nr.as<uint32_t>(), ctx->lam_val.as<double>(), ctx->lam_cnt.as<uint32_t>(), ctx->counters.as<unsigned long long>() + 4);
        else if (lambda2) {
            if (approx == APPROX_NONE) MGB_CUDA(ctx, launch_reduce_lambda2<APPROX_NONE>(ctx, ch->o0, no, sc));
            else if (approx == APPROX_ONELINE) MGB_CUDA(ctx, launch_reduce_lambda2<APPROX_ONELINE>(ctx, ch->o0, no, sc));
            else MGB_CUDA(ctx, launch_reduce_lambda2<APPROX_CLOSELINES>(ctx, ch->o0, no, sc));
        } else
            k_reduce_lambda<<<(unsigned)((nl + 255) / 256), 256, 0, ctx->stream>>>(m, ch->o0, no, ctx->lens.as<uint32_t>(), sc.L,
                                                                                   ctx->lam_val.as<double>(), ctx->lam_cnt.as<uint32_t>());
        MGB_CUDA(ctx, cudaGetLastError());
        st.ms_reduce += tr.stop();
        launches += 3;
    }
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    st.ms_total = total.stop();
    unsigned long long c[5];
    MGB_CUDA(ctx, cudaMemcpy(c, ctx->counters.p, 40, cudaMemcpyDeviceToHost));
    st.n_entries = c[0]; st.n_large_shift = c[1];
    if (K && c[4]) return ctx->fail(MGB_ENOMEM, "banded Lambda operator: %llu list insertions exceeded the capacity %u", c[4], ctx->lam_cap);
    if (ctx->solve_v2 == 4 && getenv("MGB_V4_DEBUG"))
        fprintf(stderr, "[mgb v4] records %.2f ms, slots + line integrals %.2f ms, sweep %.2f ms\n", ctx->ms_v4[0], ctx->ms_v4[1], ctx->ms_v4[2]);
    st.n_steps = range_steps;
    st.n_nbs_scanned = ctx->tot_scanned; /* of the counted range */
    st.n_solves = (uint64_t)(o_end - o_begin) * m.hnrays * m.nfreqs;
    st.max_ray_length = ctx->max_len;
    st.n_chunks = (uint32_t)ctx->chunks.size();
    st.n_launches = launches;
    st.paths_cached = all_cached ? 1 : 0;
    ctx->res_o0 = o_begin; ctx->res_o1 = o_end; ctx->res_K = K; ctx->res_valid = true; ctx->res_anis = anis;
    if (stats) *stats = st;
    return MGB_OK;
}

extern "C" int mgb_solve_pipeline_info(mgb_ctx* ctx, int* kernel, float* ms_records, float* ms_line_integrals, float* ms_sweep, uint32_t* n_ranges) {
    if (!ctx) return MGB_EINVAL;
    if (kernel) *kernel = ctx->solve_v2;
    if (ms_records) *ms_records = ctx->ms_v4[0];
    if (ms_line_integrals) *ms_line_integrals = ctx->ms_v4[1];
    if (ms_sweep) *ms_sweep = ctx->ms_v4[2];
    if (n_ranges) *n_ranges = (uint32_t)ctx->ms_v4[3];
    return MGB_OK;
}

extern "C" int mgb_run_feautrier_order_2(mgb_ctx* ctx, int approx, uint32_t o_begin, uint32_t o_end, mgb_stats* stats) {
    return run_radiation_field(ctx, approx, o_begin, o_end, stats, false, 0.0);
}

extern "C" int mgb_run_shortchar_order_0(mgb_ctx* ctx, int approx, double min_dtau, uint32_t o_begin, uint32_t o_end, mgb_stats* stats) {
    return run_radiation_field(ctx, approx, o_begin, o_end, stats, true, min_dtau);
}

static int check_range(mgb_ctx* ctx, uint32_t o0, uint32_t o1, const char* who) {
    if (!ctx->res_valid) return ctx->fail(MGB_ESTATE, "%s: no results (run first)", who);
    if (o0 >= o1 || o0 < ctx->res_o0 || o1 > ctx->res_o1) /* mgb_allgather_results widens the valid range to [0, npoints) */
        return ctx->fail(MGB_EINVAL, "%s: range [%u,%u) outside the computed range [%u,%u)", who, o0, o1, ctx->res_o0, ctx->res_o1);
    return MGB_OK;
}

extern "C" int mgb_get_J(mgb_ctx* ctx, uint32_t o_begin, uint32_t o_end, double* J) {
    if (!ctx || !J) return MGB_EINVAL;
    MGB_TRY(check_range(ctx, o_begin, o_end, "mgb_get_J"));
    cudaSetDevice(ctx->device);
    const size_t F = ctx->m.nfreqs;
    MGB_CUDA(ctx, cudaMemcpy(J, ctx->J.as<double>() + (size_t)o_begin * F, 8 * (size_t)(o_end - o_begin) * F, cudaMemcpyDeviceToHost));
    return MGB_OK;
}

extern "C" int mgb_get_J_lines(mgb_ctx* ctx, uint32_t o_begin, uint32_t o_end, double* J_lines) {
    if (!ctx || !J_lines) return MGB_EINVAL;
    MGB_TRY(check_range(ctx, o_begin, o_end, "mgb_get_J_lines"));
    cudaSetDevice(ctx->device);
    const size_t L = ctx->m.nlines;
    MGB_CUDA(ctx, cudaMemcpy(J_lines, ctx->J_lines.as<double>() + (size_t)o_begin * L, 8 * (size_t)(o_end - o_begin) * L, cudaMemcpyDeviceToHost));
    return MGB_OK;
}

extern "C" int mgb_get_J_anis(mgb_ctx* ctx, uint32_t o_begin, uint32_t o_end, double* out) {
    if (!ctx || !out) return MGB_EINVAL;
    MGB_TRY(check_range(ctx, o_begin, o_end, "mgb_get_J_anis"));
    if (!ctx->res_anis || !ctx->J_anis.p) return ctx->fail(MGB_ESTATE, "mgb_get_J_anis: the last run did not have params.anisotropic set");
    if (ctx->nranks > 1 && (int)ctx->part_bounds.size() == ctx->nranks + 1 && (o_begin < ctx->part_bounds[ctx->rank] || o_end > ctx->part_bounds[ctx->rank + 1]))
        return ctx->fail(MGB_EUNSUPPORTED, "mgb_get_J_anis: the anisotropic tensor is not part of the all-gather; ask for this rank's own rows");
    cudaSetDevice(ctx->device);
    const size_t L = ctx->m.nlines, plane = (size_t)ctx->m.npoints * L, rows = (size_t)(o_end - o_begin) * L;
    for (int q = 0; q < 6; q++)
        MGB_CUDA(ctx, cudaMemcpy(out + q * rows, ctx->J_anis.as<double>() + q * plane + (size_t)o_begin * L, 8 * rows, cudaMemcpyDeviceToHost));
    return MGB_OK;
}

extern "C" int mgb_get_lambda(mgb_ctx* ctx, uint32_t o_begin, uint32_t o_end, double* lambda_val, uint32_t* lambda_cnt) {
    if (!ctx || !lambda_val || !lambda_cnt) return MGB_EINVAL;
    MGB_TRY(check_range(ctx, o_begin, o_end, "mgb_get_lambda"));
    cudaSetDevice(ctx->device);
    const size_t L = ctx->m.nlines, n = (size_t)(o_end - o_begin) * L;
    MGB_CUDA(ctx, cudaMemcpy(lambda_val, ctx->lam_val.as<double>() + (size_t)o_begin * L, 8 * n, cudaMemcpyDeviceToHost));
    MGB_CUDA(ctx, cudaMemcpy(lambda_cnt, ctx->lam_cnt.as<uint32_t>() + (size_t)o_begin * L, 4 * n, cudaMemcpyDeviceToHost));
    return MGB_OK;
}

// A radiation field computed elsewhere (another run, a checkpoint, the reference) as the result of the "last run": what
// mgb_lp_compute_Jeff and the getters then see. Any of J / J_lines may be NULL (left as it is).
extern "C" int mgb_set_radiation_field(mgb_ctx* ctx, const double* J, const double* J_lines, const double* lambda_val, const uint32_t* lambda_cnt) {
    if (!ctx || !lambda_val || !lambda_cnt) return MGB_EINVAL;
    if (!(ctx->have_geometry && ctx->have_lines && ctx->have_freqs)) return ctx->fail(MGB_ESTATE, "mgb_set_radiation_field: geometry / lines / frequencies not set");
    cudaSetDevice(ctx->device);
    MGB_TRY(mgb_internal_alloc_results(ctx));
    const DevModel& m = ctx->m;
    const size_t N = m.npoints;
    if (J) MGB_CUDA(ctx, cudaMemcpyAsync(ctx->J.p, J, 8 * N * m.nfreqs, cudaMemcpyHostToDevice, ctx->stream));
    if (J_lines) MGB_CUDA(ctx, cudaMemcpyAsync(ctx->J_lines.p, J_lines, 8 * N * m.nlines, cudaMemcpyHostToDevice, ctx->stream));
    MGB_CUDA(ctx, cudaMemcpyAsync(ctx->lam_val.p, lambda_val, 8 * N * m.nlines, cudaMemcpyHostToDevice, ctx->stream));
    MGB_CUDA(ctx, cudaMemcpyAsync(ctx->lam_cnt.p, lambda_cnt, 4 * N * m.nlines, cudaMemcpyHostToDevice, ctx->stream));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->res_o0 = 0; ctx->res_o1 = m.npoints; ctx->res_K = 0; ctx->lam_cap = 1; ctx->res_valid = true;
    return MGB_OK;
}

extern "C" int mgb_lambda_capacity(mgb_ctx* ctx, uint32_t* cap) {
    if (!ctx || !cap) return MGB_EINVAL;
    /* the capacity of the lists the LAST RUN produced (what mgb_get_lambda_lists copies), not of the current parameters */
    *cap = ctx->res_valid ? ctx->lam_cap : (ctx->params.n_off_diag ? 1 + 2 * ctx->params.n_off_diag * ctx->m.hnrays : 1);
    return MGB_OK;
}

extern "C" int mgb_get_lambda_lists(mgb_ctx* ctx, uint32_t o_begin, uint32_t o_end, double* list_val, uint32_t* list_nr, uint32_t* list_cnt) {
    if (!ctx || !list_val || !list_nr || !list_cnt) return MGB_EINVAL;
    MGB_TRY(check_range(ctx, o_begin, o_end, "mgb_get_lambda_lists"));
    if (!ctx->res_K || !ctx->lam_list_val.p) return ctx->fail(MGB_ESTATE, "mgb_get_lambda_lists: the last run had n_off_diag = 0 (use mgb_get_lambda)");
    cudaSetDevice(ctx->device);
    const size_t L = ctx->m.nlines, n = (size_t)(o_end - o_begin) * L, cap = ctx->lam_cap;
    MGB_CUDA(ctx, cudaMemcpy(list_val, ctx->lam_list_val.as<double>() + (size_t)o_begin * L * cap, 8 * n * cap, cudaMemcpyDeviceToHost));
    MGB_CUDA(ctx, cudaMemcpy(list_nr, ctx->lam_list_nr.as<uint32_t>() + (size_t)o_begin * L * cap, 4 * n * cap, cudaMemcpyDeviceToHost));
    MGB_CUDA(ctx, cudaMemcpy(list_cnt, ctx->lam_cnt.as<uint32_t>() + (size_t)o_begin * L, 4 * n, cudaMemcpyDeviceToHost));
    return MGB_OK;
}

extern "C" int mgb_get_u(mgb_ctx* ctx, uint32_t o_begin, uint32_t o_end, double* u) {
    if (!ctx || !u) return MGB_EINVAL;
    if (!ctx->params.store_intensities || !ctx->u_full.p) return ctx->fail(MGB_ESTATE, "mgb_get_u: store_intensities was not set for the last run");
    if (o_begin != ctx->u_o0 || o_end != ctx->u_o0 + ctx->u_no) return ctx->fail(MGB_EINVAL, "mgb_get_u: range must equal the run's range");
    cudaSetDevice(ctx->device);
    MGB_CUDA(ctx, cudaMemcpy(u, ctx->u_full.p, 8 * (size_t)ctx->m.hnrays * ctx->u_no * ctx->m.nfreqs, cudaMemcpyDeviceToHost));
    return MGB_OK;
}

extern "C" int mgb_solve_feautrier_order_2(mgb_ctx* ctx, int approx, const double* emissivity, const double* opacity,
                                           const double* inverse_width, const double* population_tot, uint32_t o_begin,
                                           uint32_t o_end, double* J, double* lambda_val, uint32_t* lambda_cnt, double* u,
                                           mgb_stats* stats) {
    if (!ctx) return MGB_EINVAL;
    MGB_TRY(mgb_set_line_state(ctx, emissivity, opacity, inverse_width, population_tot));
    MGB_TRY(mgb_run_feautrier_order_2(ctx, approx, o_begin, o_end, stats));
    if (J) MGB_TRY(mgb_get_J(ctx, o_begin, o_end, J));
    if (lambda_val && lambda_cnt) MGB_TRY(mgb_get_lambda(ctx, o_begin, o_end, lambda_val, lambda_cnt));
    if (u) MGB_TRY(mgb_get_u(ctx, o_begin, o_end, u));
    return MGB_OK;
}

// Model::set_column (model.cpp:1143-1154) -> Solver::set_column / get_column (solver.tpp:2953-2991).
extern "C" int mgb_set_column(mgb_ctx* ctx, const double* density, uint32_t o_begin, uint32_t o_end, double* column, float* ms_kernel) {
    if (!ctx || !density || !column) return MGB_EINVAL;
    if (!ctx->have_geometry) return ctx->fail(MGB_ESTATE, "mgb_set_column before mgb_set_geometry");
    const DevModel& m = ctx->m;
    if (o_begin >= o_end || o_end > m.npoints) return ctx->fail(MGB_EINVAL, "mgb_set_column: bad origin range [%u,%u)", o_begin, o_end);
    cudaSetDevice(ctx->device);
    const uint32_t no = o_end - o_begin;
    const size_t nw   = (size_t)no * m.nrays;
    DevBuf rho, col;
    MGB_CUDA(ctx, rho.ensure(8 * (size_t)m.npoints));
    MGB_CUDA(ctx, col.ensure(8 * nw));
    MGB_CUDA(ctx, cudaMemcpyAsync(rho.p, density, 8 * (size_t)m.npoints, cudaMemcpyHostToDevice, ctx->stream));
    EventTimer t(ctx->stream);
    t.start();
    k_column<<<(unsigned)((nw * 32 + 127) / 128), 128, 0, ctx->stream>>>(m, o_begin, no, rho.as<double>(), col.as<double>());
    MGB_CUDA(ctx, cudaGetLastError());
    const float ms = t.stop();
    if (ms_kernel) *ms_kernel = ms;
    MGB_CUDA(ctx, cudaMemcpyAsync(column, col.p, 8 * nw, cudaMemcpyDeviceToHost, ctx->stream));
    MGB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return MGB_OK;
}

extern "C" int mgb_result_ptrs_dev(mgb_ctx* ctx, void** J_dev, void** lambda_val_dev, void** lambda_cnt_dev) {
    if (!ctx) return MGB_EINVAL;
    if (!ctx->J.p) return ctx->fail(MGB_ESTATE, "mgb_result_ptrs_dev: no results yet");
    if (J_dev) *J_dev = ctx->J.p;
    if (lambda_val_dev) *lambda_val_dev = ctx->lam_val.p;
    if (lambda_cnt_dev) *lambda_cnt_dev = ctx->lam_cnt.p;
    return MGB_OK;
}
