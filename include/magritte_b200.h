/*
 * magritte_b200.h -- C ABI of libmagritte_b200.so: the B200 implementation of Magritte's
 * Feautrier order-2 formal-solver hot path.
 *
 * This is the drop-in boundary. The reference has no FFI layer of its own
 * for this path: `Solver` is a header-only template struct instantiated inside
 * src/model/model.cpp:353-390 and reached from Python through
 * src/bindings/pybindings.cpp:185 (`compute_radiation_field_feautrier_order_2`). The entry points
 * below are what a `Model` method would bind instead of constructing that `Solver`; every one
 * cites the reference code it replaces. Plain pointers and sizes only; all pointers are HOST
 * pointers unless a name ends in `_dev`. Every function returns 0 on success and a negative
 * MGB_E* code on failure; mgb_last_error() then describes the failure. There is no CPU fallback:
 * without a CUDA device mgb_create() fails.
 *
 * Array layouts are the reference's (row-major Paracabs Matrix/Tensor,
 * dependencies/Paracabs/Paracabs/datatypes/matrix.hpp:74, tensor.hpp:77; Vector3D = 3 doubles).
 */
#ifndef MAGRITTE_B200_H
#define MAGRITTE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MGB_ABI_VERSION 2

enum { MGB_OK = 0, MGB_EINVAL = -1, MGB_ECUDA = -2, MGB_ESTATE = -3, MGB_ENOMEM = -4, MGB_EUNSUPPORTED = -5, MGB_ENCCL = -6 };

/* ApproximationType of the reference (src/solver/solver.hpp: enum ApproximationType {None, OneLine, CloseLines});
 * chosen by Model::compute_radiation_field_feautrier_order_2 from the parameters (model.cpp:366-389). */
enum { MGB_APPROX_NONE = 0, MGB_APPROX_ONELINE = 1, MGB_APPROX_CLOSELINES = 2 };

/* BoundaryCondition (src/model/geometry/boundary/boundary.hpp: enum BoundaryCondition {Zero, Thermal, CMB}). */
enum { MGB_BC_ZERO = 0, MGB_BC_THERMAL = 1, MGB_BC_CMB = 2 };

typedef struct mgb_ctx mgb_ctx;

/* The tunables of Parameters that the path reads (src/model/parameters/parameters.hpp:66-124). */
typedef struct mgb_params {
    double min_opacity;                       /* parameters.hpp: min_opacity = 1e-26 */
    double max_distance_opacity_contribution; /* = 10 (CloseLines window, solver.tpp:1219) */
    double max_interpolation_diff;            /* = 1.4 (sub-step criterion, solver_interp_helper.tpp:68) */
    double max_width_fraction;                /* = 0.35 (Model::set_dshift_max, model.cpp:1158-1178) */
    uint32_t n_off_diag;                      /* = 0; banded Lambda operator up to 4 off-diagonals (results: mgb_get_lambda_lists) */
    uint32_t store_intensities;               /* keep u(rr,o,f) (Radiation::u, solver.tpp:461-465,503) */
    /* implementation knobs (no reference counterpart) */
    uint64_t path_cache_bytes;                /* HBM budget for cached ray paths; 0 = library default */
    uint64_t scratch_bytes;                   /* HBM budget for per-chunk u/Lambda scratch; 0 = default */
    uint32_t cache_paths;                     /* 1: keep traced paths across calls when they fit the budget */
    uint32_t anisotropic;                     /* 1: also accumulate the anisotropic radiation-field tensor (mgb_get_J_anis) */
} mgb_params;

/* Work and timing counters of the last mgb_run_feautrier_order_2 (Sigma S, Sigma D, Sigma N). */
typedef struct mgb_stats {
    uint64_t n_steps;        /* Sigma S: geometric ray steps (valid get_next results), both directions */
    uint64_t n_nbs_scanned;  /* Sigma D: neighbours inspected by get_next */
    uint64_t n_entries;      /* Sigma N: ray entries summed over (o, rr, line) (sub-steps included) */
    uint64_t n_solves;       /* (o, rr, f) tridiagonal systems solved */
    uint64_t n_large_shift;  /* segments x frequencies that took the line-integrated (erff) branch */
    uint32_t max_ray_length; /* longest geometric half-ray pair (steps) */
    uint32_t n_chunks;       /* origin chunks the call was split into */
    uint32_t n_launches;     /* kernels launched by the call */
    uint32_t paths_cached;   /* 1 if the trace was served from the path cache */
    float ms_trace_count;    /* device time, CUDA events on the launching stream */
    float ms_trace_fill;
    float ms_solve;          /* the dominant kernel: eta/chi + two-sided Thomas sweep */
    float ms_reduce;         /* J and Lambda accumulation in reference order */
    float ms_total;          /* first launch to last completion */
    uint32_t n_solve_launches;
    uint32_t reserved;
} mgb_stats;

/* ---- life cycle ----------------------------------------------------------------------------- */
int mgb_abi_version(void);
/* Create a context on CUDA device `device` (cudaSetDevice). Replaces the per-call `Solver solver;`
 * of model.cpp:356 and Paracabs' accelerator state. Fails with MGB_ECUDA when there is no device. */
int mgb_create(int device, mgb_ctx** out);
void mgb_destroy(mgb_ctx* ctx);
const char* mgb_last_error(const mgb_ctx* ctx); /* ctx may be NULL: error of the last failed mgb_create */

/* ---- model upload (the device mirrors of src/model/{geometry,thermodynamics,lines,radiation}) -- */
/* Geometry: points.position/velocity (points.hpp), CSR neighbour lists n_neighbors/neighbors
 * (points.hpp:15-17; cum_n_neighbors is rebuilt here), rays.direction/weight/antipod (rays.hpp),
 * boundary.point2boundary / boundary_condition / boundary_temperature (boundary.hpp).
 * velocity is v/c as in the reference. point2boundary[p] == npoints marks an interior point. */
int mgb_set_geometry(mgb_ctx* ctx, uint32_t npoints, const double* position, const double* velocity,
                     const uint32_t* n_neighbors, const uint32_t* neighbors, uint64_t n_neighbors_total,
                     uint32_t nrays, const double* direction, const double* weight, const uint32_t* antipod,
                     uint32_t nboundary, const uint32_t* point2boundary, const int32_t* boundary_condition,
                     const double* boundary_temperature, int spherical_symmetry);
/* Rays::use_adaptive_directions (rays.hpp:23-31, rays.cpp:37-100): one ray set PER POINT, direction [npoints][nrays][3] and weight
 * [npoints][nrays] (what Rays::read finds when the direction array has nrays*npoints rows). Call after mgb_set_geometry, whose
 * direction / weight arguments then hold point 0's set (the antipode table is derived from it for all points, like the reference).
 * Selects the <..., true> instantiations of Geometry::get_next / get_shift and Rays::get_weight (model.cpp:357-389). */
int mgb_set_adaptive_directions(mgb_ctx* ctx, const double* direction, const double* weight);
/* thermodynamics.temperature.gas, thermodynamics.turbulence.vturb2 (both [npoints]). */
int mgb_set_thermodynamics(mgb_ctx* ctx, const double* temperature_gas, const double* vturb2);
/* Static line tables (Lines::read, src/model/lines/lines.cpp:11-69): per species nrad and
 * linedata.inverse_mass; per line (species-major order, lid = nrad_cum[lspec] + k) the centre
 * frequency lines.line and Einstein A; quadrature.weights [nlspecs][nquads]. The sorted line list
 * of lines.cpp:52-64 is rebuilt inside. */
int mgb_set_lines(mgb_ctx* ctx, uint32_t nlspecs, const uint32_t* nrad, const double* inverse_mass,
                  const double* line, const double* einstein_A, uint32_t nquads, const double* quadrature_weights);
/* radiation.frequencies.nu [npoints][nfreqs] with f = lid*nquads + z (Model::compute_spectral_discretisation,
 * model.cpp:70-161) and corresponding_line [nfreqs] (frequencies.hpp:13-22; only read by OneLine). */
int mgb_set_frequencies(mgb_ctx* ctx, uint32_t nfreqs, const double* nu, const uint32_t* corresponding_line);
/* Per-iteration line state: lines.emissivity / opacity / inverse_width [npoints][nlines]
 * (Lines::set_emissivity_and_opacity, lines.tpp:29-41; set_inverse_width, :45-59) and
 * lineProducingSpecies[l].population_tot [nlspecs][npoints]. Also derives the InterpHelper criterion
 * (solver_interp_helper.hpp:15-30) and Model::dshift_max (model.cpp:1158-1178) on the device. */
int mgb_set_line_state(mgb_ctx* ctx, const double* emissivity, const double* opacity, const double* inverse_width,
                       const double* population_tot);
int mgb_params_default(mgb_params* p);
int mgb_set_params(mgb_ctx* ctx, const mgb_params* p);

/* ---- the hot path --------------------------------------------------------------------------- */
/* Solver::setup + Solver::solve_feautrier_order_2<approx>(Model&) (solver.tpp:1-13, 458-540) for origins
 * o in [o_begin, o_end) and all hnrays ray pairs: trace, eta/chi, Feautrier order-2 solve, J and Lambda
 * accumulation, results left in device memory. Synchronous. `stats` may be NULL. */
int mgb_run_feautrier_order_2(mgb_ctx* ctx, int approx, uint32_t o_begin, uint32_t o_end, mgb_stats* stats);
/* Solver::setup + Solver::solve_shortchar_order_0<approx>(Model&) (solver.tpp:150-235; the half-ray integrations
 * solve_shortchar_order_0_ray_forward / _backward, :1564-1654 / :1664-1757; Model::compute_radiation_field_shortchar_order_0,
 * model.cpp:313-349): same traced paths, chunks and result arrays as the Feautrier run -- J, u = (I_f + I_b) / 2 (with
 * store_intensities), and the LOCAL Lambda element only (n_off_diag is ignored, like the reference does). min_dtau is
 * Parameters::min_dtau (1e-15). A half ray without a segment contributes the factor 0 to its Lambda element (the
 * reference reads an uninitialised dtau there). */
int mgb_run_shortchar_order_0(mgb_ctx* ctx, int approx, double min_dtau, uint32_t o_begin, uint32_t o_end, mgb_stats* stats);
/* Which kernel generation served the last run for n_off_diag = 0 (environment MGB_SOLVE_KERNEL: 1 k_solve, 2 k_solve2, 3 k_solve3,
 * 4 = default: the streaming pipeline k_ni / k_fill -> k_large -> k_sweep of mgb_solve4.cuh; all bit-identical) and, for the
 * pipeline, the device time of its three stages and the number of record ranges (k_sweep launches). No reference counterpart:
 * measurement aid for bench.py. Any pointer may be NULL. */
int mgb_solve_pipeline_info(mgb_ctx* ctx, int* kernel, float* ms_records, float* ms_line_integrals, float* ms_sweep, uint32_t* n_ranges);
/* Copy results of the last run to the host. J: rows [o_begin,o_end) of radiation.J, (o_end-o_begin)*nfreqs.
 * Lambda (n_off_diag = 0: the diagonal operator): lambda_val [(o_end-o_begin)][nlines] holds Ls for the
 * element (p=o, k, nr=o) that Lambda::add_element accumulated (lambda.tpp:79-89), lambda_cnt the list
 * length (0 or 1). u: radiation.u rows, [hnrays][(o_end-o_begin)][nfreqs], only if store_intensities. */
int mgb_get_J(mgb_ctx* ctx, uint32_t o_begin, uint32_t o_end, double* J);
int mgb_get_lambda(mgb_ctx* ctx, uint32_t o_begin, uint32_t o_end, double* lambda_val, uint32_t* lambda_cnt);
/* The memory-sparse solver's output (Solver::solve_feautrier_order_2_sparse, solver.tpp:297-363; what
 * Model::compute_radiation_field_feautrier_order_2_sparse leaves in lspec.J): J integrated over each line's quadrature,
 * J_lines [(o_end-o_begin)][nlines] (line = global line index; species s owns columns nrad_cum[s] .. +nrad[s]). Every run
 * accumulates it next to J from the same solves, in the reference's rr > line > z order. */
int mgb_get_J_lines(mgb_ctx* ctx, uint32_t o_begin, uint32_t o_end, double* J_lines);
int mgb_get_u(mgb_ctx* ctx, uint32_t o_begin, uint32_t o_end, double* u);
/* Solver::solve_feautrier_order_2_anis (solver.tpp:365-456; Model::compute_radiation_field_feautrier_order_2_anis, model.cpp:435-470): the
 * same solves accumulated per line into lspec.J and the five anisotropic tensor elements J2_0, J2_1_Re, J2_1_Im, J2_2_Re, J2_2_Im
 * (lineProducingSpecies.hpp:29-35) with du = w_z * weight(o,rr) * u, J += 2 du, J2_x += wt_x(n) * du in the reference's rr > line > z
 * order. Needs params.anisotropic = 1 for the run. out [6][(o_end-o_begin)][nlines] in that order of quantities. */
int mgb_get_J_anis(mgb_ctx* ctx, uint32_t o_begin, uint32_t o_end, double* out);
/* n_off_diag = K > 0 (banded approximate Lambda operator, Solver::update_Lambda solver.tpp:1806-1850 with L_lower/L_upper of
 * :2320-2355): the (p, line) lists of Lambda::add_element (lambda.tpp:79-89) in the reference's insertion order.
 * capacity = 1 + 2*K*hnrays entries per list; list_val / list_nr [(o_end-o_begin)][nlines][capacity] (unused tail: 0 / 0xffffffff),
 * list_cnt [(o_end-o_begin)][nlines]. With K > 0 mgb_get_lambda returns the diagonal element and the list length. */
int mgb_lambda_capacity(mgb_ctx* ctx, uint32_t* capacity);
int mgb_get_lambda_lists(mgb_ctx* ctx, uint32_t o_begin, uint32_t o_end, double* list_val, uint32_t* list_nr, uint32_t* list_cnt);
/* One call, host buffers in and out (what Model::compute_radiation_field_feautrier_order_2 binds):
 * mgb_set_line_state + mgb_run + mgb_get_J + mgb_get_lambda (+ mgb_get_u if u != NULL). */
int mgb_solve_feautrier_order_2(mgb_ctx* ctx, int approx, const double* emissivity, const double* opacity,
                                const double* inverse_width, const double* population_tot, uint32_t o_begin,
                                uint32_t o_end, double* J, double* lambda_val, uint32_t* lambda_cnt, double* u,
                                mgb_stats* stats);
/* Install a radiation field computed elsewhere (a checkpoint, another run) as the result of the last run: radiation.J [npoints][nfreqs],
 * J per line [npoints][nlines] (either may be NULL: left as it is), the diagonal Lambda element and list length [npoints][nlines]. */
int mgb_set_radiation_field(mgb_ctx* ctx, const double* J, const double* J_lines, const double* lambda_val, const uint32_t* lambda_cnt);
/* Device addresses of the result buffers (full-size [npoints][nfreqs] / [npoints][nlines]) for in-place
 * collectives: replaces Radiation::MPI_reduce_J / Lambda::MPI_gather (radiation.cpp:76-91, lambda.tpp:121-206). */
int mgb_result_ptrs_dev(mgb_ctx* ctx, void** J_dev, void** lambda_val_dev, void** lambda_cnt_dev);
/* Drop the cached ray lengths and paths (geometry unchanged): the next run repeats the count pass and the trace. */
int mgb_invalidate_paths(mgb_ctx* ctx);
/* Device-side stopwatch on the context's stream (CUDA events): start, then stop returns elapsed milliseconds.
 * For benchmarking a sequence of calls without host clocks. */
int mgb_timer_start(mgb_ctx* ctx);
int mgb_timer_stop(mgb_ctx* ctx, float* ms);

/* ---- the new imager ------------------------------------------------------ */
/* Work and timing counters of the last mgb_image_feautrier_order_2. */
typedef struct mgb_image_stats {
    uint64_t n_pixels;       /* pixels of this call's range */
    uint64_t n_entries;      /* ray entries (sub-steps included) summed over the pixels, start points excluded */
    uint64_t n_solves;       /* pixels x frequencies */
    uint32_t max_ray_length; /* longest pixel ray (entries): Solver::get_ray_lengths_max_new_imager */
    uint32_t start_bdy;      /* Geometry::get_closest_bdy_point_in_custom_raydir */
    uint32_t n_chunks, n_launches;
    float ms_trace, ms_solve, ms_total;
    uint32_t reserved;
} mgb_image_stats;
/* Model::compute_image_new(rx, ry, rz, Nxpix, Nypix) (model.cpp:999-1035) -> Solver::image_feautrier_order_2_new_imager<approx>
 * (solver.tpp:576-627): Image::set_coordinates_projection_surface (image.cpp:221-368), per pixel the start search
 * (trace_ray_imaging_get_start, solver.tpp:808-825), the Rest-frame imaging trace with all-lines sub-stepping
 * (trace_ray_imaging :840-897, set_data :1003-1043) and image_feautrier_order_2 (:2462-2574) for every frequency.
 * ray_dir need not be normed (it is normed like model.cpp:1014). dimension = Parameters::dimension() (1 or 3).
 * Frequencies: nu [nfreqs] when nu_per_point == 0 (Model::compute_spectral_discretisation(nu_min, nu_max, n) gives every
 * point the same bins, model.cpp:257-300), else nu [npoints][nfreqs] (radiation.frequencies.nu; the pixel uses the row of
 * its start point, solver.tpp:611-614). Pixels [pix_begin, pix_end) of the nx*ny image (index y + ny*x) are computed.
 * Outputs (host): ImX, ImY [nx*ny] (may be NULL), I [(pix_end-pix_begin)][nfreqs]. Needs geometry, thermodynamics, lines and
 * line state; errors mirror the reference's exceptions: zero direction, OneLine (not supported for imaging), and a 1D
 * model imaged along anything but (0,1,0) (image.cpp:80-86) return MGB_EINVAL / MGB_EUNSUPPORTED with the same text.
 * Host side of a call: I may be ordinary pageable memory -- the image leaves the device in slices through a pinned staging buffer
 * of the context (at most 1 GiB, allocated on first use, freed by mgb_destroy) and up to 8 short-lived host threads copy the
 * slices on while the next slice is solved. The call is synchronous: I is complete when it returns. */
int mgb_image_feautrier_order_2(mgb_ctx* ctx, int approx, int dimension, const double* ray_dir, uint32_t nx, uint32_t ny,
                                uint32_t nfreqs, const double* nu, int nu_per_point, uint64_t pix_begin, uint64_t pix_end,
                                double* ImX, double* ImY, double* I, mgb_image_stats* stats);
/* Model::compute_image_optical_depth_new(raydir, Nxpix, Nypix) (model.cpp:1102-1127) -> Solver::image_optical_depth_new_imager<approx>
 * (solver.tpp:688-733) with Solver::image_optical_depth (:2718-2760): same surface, start search and trace as above; the image holds
 * the optical depth along each pixel's ray per frequency bin (0 for a ray without interior points). Same arguments and errors
 * (the OneLine text is the reference's "One line approximation is currently not supported for imaging."). */
int mgb_image_optical_depth(mgb_ctx* ctx, int approx, int dimension, const double* ray_dir, uint32_t nx, uint32_t ny,
                                uint32_t nfreqs, const double* nu, int nu_per_point, uint64_t pix_begin, uint64_t pix_end,
                                double* ImX, double* ImY, double* tau, mgb_image_stats* stats);

/* ---- column densities ---------------------------------------------------- */
/* Model::set_column (model.cpp:1143-1154) -> Solver::set_column / get_column (solver.tpp:2953-2991): for origins o in
 * [o_begin, o_end) and every ray r, the trapezoid sum of 0.5 * (density[crt] + density[nxt]) * dZ over the geometric steps of
 * the half-ray, outward from the origin (0 for a ray that leaves the model at once). density [npoints] (Model::density, host);
 * column [nrays][o_end-o_begin] (rows of Model::column, host). Needs the geometry only. ms_kernel (may be NULL): device time
 * of the kernel. A ray that dead-ends inside the model stops there (the reference would index density[npoints]). */
int mgb_set_column(mgb_ctx* ctx, const double* density, uint32_t o_begin, uint32_t o_end, double* column, float* ms_kernel);
/* mgb_set_column without any device allocation. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). Same inputs, outputs
 * and results as mgb_set_column (host density [npoints] in, host column [nrays][o_end-o_begin] out), but the density and the column
 * block are staged through a file-scope static device array of 4 Mi doubles (32 MiB, held by every GPU that loads the library until
 * the process exits) instead of per-call buffers: the call makes no cudaMalloc / cudaFree and uses no allocating library routine.
 * LIMIT: npoints (rounded up to a multiple of 32) + nrays * (o_end - o_begin) <= 4 194 304, else MGB_EINVAL with the sizes in
 * mgb_last_error -- split the origin range or use mgb_set_column. All calls in the process share the array per device: calls that
 * overlap on one GPU race and must be serialised by the caller. Synchronous. */
int mgb_set_column_noalloc(mgb_ctx* ctx, const double* density, uint32_t o_begin, uint32_t o_end, double* column);

/* ---- multi-GPU: origin partition + NCCL all-gather ---------------------------- */
/* 128-byte ncclUniqueId made on rank 0 and distributed by the host application. */
int mgb_nccl_unique_id(void* id128);
int mgb_comm_init(mgb_ctx* ctx, int nranks, int rank, const void* id128);
/* Destroy the communicator (ncclCommDestroy) and return the context to single-GPU operation; mgb_destroy does it too. */
int mgb_comm_destroy(mgb_ctx* ctx);
/* Even split of [0,npoints) into nranks contiguous ranges of equal element count (the last padded). */
int mgb_partition(mgb_ctx* ctx, int nranks, int rank, uint32_t* o_begin, uint32_t* o_end);
/* Contiguous ranges of equal predicted cost (sum of geometric ray lengths per origin, from the cached count pass). */
int mgb_partition_balanced(mgb_ctx* ctx, int nranks, int rank, uint32_t* o_begin, uint32_t* o_end);
/* The same split without a context (pure host arithmetic): rank r owns rows [r*per, min((r+1)*per, npoints)), per = ceil(npoints/nranks). */
int mgb_partition_rows(uint32_t npoints, int nranks, int rank, uint32_t* o_begin, uint32_t* o_end, uint32_t* rows_per_rank);
/* In-place all-gather (one grouped ncclBroadcast per owner, so ranges may differ in size and may be empty) of the J, J-per-line and
 * Lambda slices every rank computed for its mgb_partition / mgb_partition_balanced range; afterwards rows [0, npoints) are valid on every
 * rank. Collective: EVERY rank of the communicator must call it, also one whose range was empty; a rank whose run failed must not enter
 * alone -- the callers agree on the error first (the host Model all-reduces the status). Measured 0.5 ms for 485 MB on 2 B200s
 * (970 GB/s, P2P over NVLink; profiles/comm_diag_r2.log). */
int mgb_allgather_results(mgb_ctx* ctx);

/* Minimum of `status` over the ranks of the communicator (ncclAllReduce, min): how the ranks agree that every run succeeded
 * before they enter mgb_allgather_results together. nranks == 1: *min_status = status. */
int mgb_comm_agree(mgb_ctx* ctx, int status, int* min_status);

/* ---- level populations on the device ----------------------------------
 * The level-population half of an ALI iteration (Model::compute_level_populations, model.cpp:654-747) next to the radiation
 * field, so that an iteration moves nothing across PCIe: J / Lambda are read where the run left them, the per-(point, line)
 * emissivity / opacity / criterion table the solver reads is rewritten in place. Populations are the reference's long double
 * values (lineProducingSpecies.hpp:46) carried as double-double pairs: x = hi + lo (lo may be NULL / is 0 for plain doubles).
 * Block-diagonal statistical equilibrium only (n_off_diag = 0), as Model::compute_level_populations_sparse requires too. */
typedef struct mgb_lp_convergence { /* LineProducingSpecies::fraction_not_converged, relative_change_mean, relative_change_max */
    double fraction_not_converged, relative_change_mean, relative_change_max;
} mgb_lp_convergence;
/* Parameters::min_line_opacity, min_line_emissivity, population_inversion_fraction, min_rel_pop_for_convergence (parameters.hpp:66-124). */
int mgb_lp_set_parameters(mgb_ctx* ctx, double min_line_opacity, double min_line_emissivity, double population_inversion_fraction,
                          double min_rel_pop_for_convergence);
/* Linedata of line-producing species l (linedata.hpp; Linedata::read, linedata.cpp:12-71): species_index = linedata.num (column of
 * the abundance table), levels energy / weight [nlev], transitions irad / jrad / frequency / A / Ba / Bs [nrad]. After mgb_set_lines. */
int mgb_lp_set_species(mgb_ctx* ctx, uint32_t l, uint32_t species_index, uint32_t nlev, uint32_t nrad, const uint32_t* irad,
                       const uint32_t* jrad, const double* energy, const double* weight, const double* frequency, const double* A,
                       const double* Ba, const double* Bs);
/* CollisionPartner (collisionPartner.hpp; read at collisionPartner.cpp:10-45): orth_or_para 0 'n', 1 'o', 2 'p'; Ce, Cd [ntmp][ncol]. */
int mgb_lp_add_collision_partner(mgb_ctx* ctx, uint32_t l, uint32_t num_col_partner, int orth_or_para, uint32_t ntmp, uint32_t ncol,
                                 const uint32_t* icol, const uint32_t* jcol, const double* tmp, const double* Ce, const double* Cd);
/* chemistry.species.abundance [npoints][nspecs]. */
int mgb_lp_set_abundance(mgb_ctx* ctx, uint32_t nspecs, const double* abundance);
/* lineProducingSpecies[l].population [npoints][nlev] (+ population_tot [npoints]); seeds the Ng history with it, as
 * LineProducingSpecies::update_using_LTE does (lineProducingSpecies.tpp:49-69). */
int mgb_lp_set_populations(mgb_ctx* ctx, uint32_t l, const double* population_hi, const double* population_lo, const double* population_tot);
int mgb_lp_get_populations(mgb_ctx* ctx, uint32_t l, double* population_hi, double* population_lo);
/* Lines::set_emissivity_and_opacity + set_inverse_width (lines.tpp:29-59) + InterpHelper criterion (solver_interp_helper.hpp:15-30)
 * from the device populations into the solver's table; replaces the per-iteration mgb_set_line_state upload. */
int mgb_lp_refresh_line_state(mgb_ctx* ctx);
/* The table back on the host: lines.emissivity / opacity / inverse_width [npoints][nlines] (any may be NULL). */
int mgb_lp_get_line_state(mgb_ctx* ctx, double* emissivity, double* opacity, double* inverse_width);
/* Model::compute_Jeff (model.cpp:514-545; sparse != 0: compute_Jeff_sparse :549-570, from J per line) for every species, reading
 * J / Lambda of the last run in place. Needs results for all points (single GPU, or after mgb_allgather_results). */
int mgb_lp_compute_Jeff(mgb_ctx* ctx, int sparse);
int mgb_lp_get_Jeff(mgb_ctx* ctx, uint32_t l, double* Jlin, double* Jeff); /* lspec.Jlin / Jeff [npoints][nrad] */
/* Lines::iteration_using_statistical_equilibrium (lines.cpp:126-137) for every species: population_prev shift and (keep_history != 0)
 * the Ng history push, LineProducingSpecies::update_using_statistical_equilibrium (lineProducingSpecies.tpp:377-581, one nlev x nlev
 * block per point), correct_negative_populations (:711-802), check_for_convergence(pop_prec) (:128-165) -> conv [nlspecs],
 * set_emissivity_and_opacity. ms (may be NULL): device time. A singular block fails with the reference's "Eigen solver ERROR.". */
int mgb_lp_update_statistical_equilibrium(mgb_ctx* ctx, int keep_history, double pop_prec, mgb_lp_convergence* conv, float* ms);
/* Ng acceleration, LineProducingSpecies::update_using_acceleration (lineProducingSpecies.tpp:266-314) and _trial (:316-371), split
 * at the order x order solve, which stays with the host's long double QR: (1) mgb_lp_ng_gram pushes the history entry and returns the
 * Gram matrix of the last `order` residuals; (2) the host solves RTR c = 1, normalises c; (3) mgb_lp_ng_combine forms the new
 * population (trial: trial_population) and trims the history like the reference; (4) mgb_lp_finish_update (real step:
 * correct_negative_populations + check_for_convergence + set_emissivity_and_opacity, lines.cpp:106-117) or
 * mgb_lp_trial_convergence (check_for_convergence_trial, :171-201). order <= 16. */
int mgb_lp_ng_gram(mgb_ctx* ctx, uint32_t l, uint32_t order, int trial, double* gram_hi, double* gram_lo);
int mgb_lp_ng_combine(mgb_ctx* ctx, uint32_t l, uint32_t order, int trial, const double* coef_hi, const double* coef_lo);
int mgb_lp_finish_update(mgb_ctx* ctx, double pop_prec, mgb_lp_convergence* conv);
int mgb_lp_trial_convergence(mgb_ctx* ctx, double pop_prec, mgb_lp_convergence* conv);

/* ---- test hooks ------------------------------------------------------------------------------ */
/* The ray of (origin o, ray pair rr, line l) exactly as Solver::trace_ray_for_line fills nr_/nr_interp_/
 * interp_factor_/shift_/dZ_ between first_ and last_ (solver.tpp:484-496, 773-800, 1057-1097).
 * Arrays have capacity `cap`; returns n_tot in *n_out and the centre's offset from first_ in *centre_out. */
int mgb_trace_dump(mgb_ctx* ctx, uint32_t o, uint32_t rr, uint32_t l, uint32_t cap, uint32_t* nr, uint32_t* nr_interp,
                   double* interp_factor, double* shift, double* dZ, uint32_t* n_out, uint32_t* centre_out);
/* Device libm self-test: evaluates fn (0 expf, 1 logf, 2 erff, 3 expm1f, 4 powf) on n host-supplied
 * inputs (y only for powf) with the device implementation; the caller compares with its own libm. */
int mgb_selftest_libm(mgb_ctx* ctx, int fn, uint64_t n, const float* x, const float* y, float* out);
/* Sustained DFMA rate of the device in TFLOP/s (2 flops per DFMA): the FP64 roofline denominator. */
int mgb_measure_fp64_peak(mgb_ctx* ctx, double* tflops);
/* Device division self-test: out[i] = the kernels' scaled IEEE division of a[i] / b[i] (must equal a[i] / b[i] bit for bit). */
int mgb_selftest_ddiv(mgb_ctx* ctx, uint64_t n, const double* a, const double* b, double* out);

#ifdef __cplusplus
}
#endif
#endif /* MAGRITTE_B200_H */
