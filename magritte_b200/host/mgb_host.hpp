// mgb_host.hpp -- host side of the drop-in: a C++ `Model` with the method names of magritte.core's Model for the Feautrier
// order-2 path (reference src/model/model.hpp:19-122, bound in src/bindings/pybindings.cpp:123-271). It owns the host arrays
// in the reference's layouts and drives the GPU exclusively through the C ABI of include/magritte_b200.h; there is no CPU
// implementation of the radiation field here (without a CUDA device compute_radiation_field_feautrier_order_2 throws).
//
// What is implemented on the host, in the reference's operation order and number types (Real = double, populations in
// long double like Eigen's VectorXld, lineProducingSpecies.hpp:46):
//   compute_inverse_line_widths     Lines::set_inverse_width                 lines.tpp:45-59
//   compute_spectral_discretisation Model::compute_spectral_discretisation   model.cpp:70-161 (lines) and :257-300 (image bins)
//   compute_LTE_level_populations   Lines::iteration_using_LTE               lines.cpp:82-91, lineProducingSpecies.tpp:49-69
//   compute_Jeff                    Model::compute_Jeff                      model.cpp:514-545
//   compute_level_populations_from_stateq / compute_level_populations       model.cpp:655-748, lines.cpp:93-137,
//                                   lineProducingSpecies.tpp:128-165 (convergence), 266-314 / 316-364 (Ng acceleration),
//                                   377-581 (statistical equilibrium), 711-802 (negative populations / inversions)
// The statistical-equilibrium system of the reference is one global sparse LU (Eigen::SparseLU, long double). With the
// diagonal Lambda operator of this path (n_off_diag = 0) it is block diagonal -- one nlev x nlev block per point -- and is
// solved here block by block with a dense long-double LU with partial pivoting: same equations, different pivot order.
// With a banded Lambda operator (n_off_diag > 0) the blocks couple; the same global system is then solved in long double by
// iterative refinement on the original equations with block-Jacobi preconditioned GMRES corrections, down to a component-wise
// backward error of a few long-double epsilons (LineProducingSpecies::solve_statistical_equilibrium_banded).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <ctime>
#include <limits>
#include <memory>
#include <sys/mman.h>
#include <cstdlib>
#include <numeric>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/magritte_b200.h"

namespace mgbh {

typedef double Real;   // reference tools/types.hpp:25
typedef uint32_t Size; // reference tools/types.hpp:26
typedef long double LD;

// reference src/tools/constants.hpp
constexpr double CC = 2.99792458E+8, HH = 6.62607004E-34, KB = 1.38064852E-23, AMU = 1.66053904E-27, T_CMB = 2.7254800;
constexpr double CC_SQUARED = 8.98755179E+16;
constexpr double TWO_KB_OVER_AMU_CC_SQUARED = 2.0 * KB / (AMU * CC_SQUARED);
const double FOUR_PI = 4.0 * M_PI;
const double HH_OVER_FOUR_PI = HH / FOUR_PI;

struct Parameters { // reference src/model/parameters/parameters.hpp:66-124 (the tunables this path reads)
    Size n_off_diag = 0;
    Real max_width_fraction = 0.35, convergence_fraction = 0.995, min_rel_pop_for_convergence = 1.0e-10, pop_prec = 1.0e-6;
    Real min_opacity = 1.0e-26;
    Real min_dtau = 1.0e-15; // lower bound of a segment's optical depth in the short-characteristics solver (solver.tpp:1608)
    LD min_line_opacity = 1.0e-13, min_line_emissivity = 1.0e-50;
    Real max_interpolation_diff = 1.4, max_distance_opacity_contribution = 10.0;
    Real population_inversion_fraction = -1.0;
    bool store_intensities = false, use_adaptive_Ng_acceleration = true, adaptive_Ng_acceleration_use_max_criterion = true;
    Size Ng_acceleration_mem_limit = 9, adaptive_Ng_acceleration_min_order = 2, Ng_acceleration_remove_N_its = 0;
    bool one_line_approximation = false, sum_opacity_emissivity_over_all_lines = false;
    bool spherical_symmetry = false;
    Size dimension = 3;
};

struct CollisionPartner { // reference linedata/collisionPartner/collisionPartner.hpp
    Size num_col_partner = 0;
    std::string orth_or_para_H2 = "n";
    std::vector<Size> icol, jcol;
    std::vector<Real> tmp;
    std::vector<std::vector<Real>> Ce, Cd; // [ntmp][ncol]
    std::vector<Real> Ce_intpld, Cd_intpld;

    // reference src/tools/interpolation.hpp: search()
    static Size search(const std::vector<Real>& x, const Real value) {
        Size start = 0, stop = (Size)x.size() - 1;
        if (value >= x[stop]) return stop;
        else if (value <= x[start]) return start;
        while (stop > start + 1) {
            const Size middle = (stop + start) / 2;
            if (value > x[middle]) start = middle;
            else if (value < x[middle]) stop = middle;
            else return middle;
        }
        return stop;
    }
    // collisionPartner.tpp: interpolate_collision_coefficients
    void interpolate_collision_coefficients(const Real temperature_gas) {
        const Size ntmp = (Size)tmp.size(), ncol = (Size)icol.size();
        const Size t = search(tmp, temperature_gas);
        if (t == 0) { Ce_intpld = Ce[0]; Cd_intpld = Cd[0]; }
        else if (t == ntmp - 1) { Ce_intpld = Ce[ntmp - 1]; Cd_intpld = Cd[ntmp - 1]; }
        else {
            const Real step = (temperature_gas - tmp[t - 1]) / (tmp[t] - tmp[t - 1]);
            Ce_intpld.resize(ncol); Cd_intpld.resize(ncol);
            for (Size k = 0; k < ncol; k++) {
                Ce_intpld[k] = Ce[t - 1][k] + (Ce[t][k] - Ce[t - 1][k]) * step;
                Cd_intpld[k] = Cd[t - 1][k] + (Cd[t][k] - Cd[t - 1][k]) * step;
            }
        }
    }
    // collisionPartner.tpp: adjust_abundance_for_ortho_or_para
    void adjust_abundance_for_ortho_or_para(const Real temperature_gas, Real& abundance) const {
        if (orth_or_para_H2 != "n") {
            const Real frac_H2_para = 1.0 / (1.0 + 9.0 * expf(-170.5 / temperature_gas));
            if (orth_or_para_H2 == "o") abundance *= (1.0 - frac_H2_para);
            if (orth_or_para_H2 == "p") abundance *= frac_H2_para;
        }
    }
};

struct Linedata { // reference linedata/linedata.hpp
    Size num = 0, nlev = 0, nrad = 0;
    Real inverse_mass = 0.0;
    std::vector<Size> irad, jrad;
    std::vector<Real> energy, weight, frequency, A, Ba, Bs;
    std::vector<CollisionPartner> colpar;
};

typedef std::vector<LD> VecLD;
inline LD dot(const VecLD& a, const VecLD& b) { LD s = 0; for (size_t i = 0; i < a.size(); i++) s += a[i] * b[i]; return s; }

// Dense long-double LU with partial pivoting, in place: A [n][n] row-major, b [n] -> solution. false on a zero pivot.
inline bool lu_solve(std::vector<LD>& A, std::vector<LD>& b, const Size n) {
    for (Size c = 0; c < n; c++) {
        Size piv = c;
        LD big = fabsl(A[(size_t)c * n + c]);
        for (Size r = c + 1; r < n; r++) if (fabsl(A[(size_t)r * n + c]) > big) { big = fabsl(A[(size_t)r * n + c]); piv = r; }
        if (big == 0.0L) return false;
        if (piv != c) { for (Size k = 0; k < n; k++) std::swap(A[(size_t)c * n + k], A[(size_t)piv * n + k]); std::swap(b[c], b[piv]); }
        const LD inv = 1.0L / A[(size_t)c * n + c];
        for (Size r = c + 1; r < n; r++) {
            const LD f = A[(size_t)r * n + c] * inv;
            if (f == 0.0L) continue;
            for (Size k = c + 1; k < n; k++) A[(size_t)r * n + k] -= f * A[(size_t)c * n + k];
            b[r] -= f * b[c];
        }
    }
    for (Size r = n; r-- > 0;) {
        LD s = b[r];
        for (Size k = r + 1; k < n; k++) s -= A[(size_t)r * n + k] * b[k];
        b[r] = s / A[(size_t)r * n + r];
    }
    return true;
}

// The same factorisation kept for repeated solves: A [n][n] row-major -> L\U in place, piv [n]. false on a zero pivot.
inline bool lu_factor(LD* A, Size* piv, const Size n) {
    for (Size c = 0; c < n; c++) {
        Size pr = c;
        LD big = fabsl(A[(size_t)c * n + c]);
        for (Size r = c + 1; r < n; r++) if (fabsl(A[(size_t)r * n + c]) > big) { big = fabsl(A[(size_t)r * n + c]); pr = r; }
        if (big == 0.0L) return false;
        piv[c] = pr;
        if (pr != c) for (Size k = 0; k < n; k++) std::swap(A[(size_t)c * n + k], A[(size_t)pr * n + k]);
        const LD inv = 1.0L / A[(size_t)c * n + c];
        for (Size r = c + 1; r < n; r++) {
            const LD f = A[(size_t)r * n + c] * inv;
            A[(size_t)r * n + c] = f;
            if (f == 0.0L) continue;
            for (Size k = c + 1; k < n; k++) A[(size_t)r * n + k] -= f * A[(size_t)c * n + k];
        }
    }
    return true;
}
inline void lu_apply(const LD* A, const Size* piv, LD* b, const Size n) {
    for (Size c = 0; c < n; c++) {
        if (piv[c] != c) std::swap(b[c], b[piv[c]]);
        const LD bc = b[c];
        if (bc != 0.0L) for (Size r = c + 1; r < n; r++) b[r] -= A[(size_t)r * n + c] * bc;
    }
    for (Size r = n; r-- > 0;) {
        LD t = b[r];
        for (Size k = r + 1; k < n; k++) t -= A[(size_t)r * n + k] * b[k];
        b[r] = t / A[(size_t)r * n + r];
    }
}

// Band matrix (kl sub-, ku super-diagonals) in LAPACK band storage with room for the fill of row pivoting: element (i, j) at
// AB[(kl + ku + i - j) + ldab * j], ldab = 2 kl + ku + 1. LU with partial pivoting (the unblocked dgbtf2 scheme) and the solve of
// one right-hand side, in place. false on a zero pivot.
inline bool band_lu_solve(std::vector<LD>& AB, const size_t n, const size_t kl, const size_t ku, std::vector<LD>& b) {
    const size_t kv = kl + ku, ldab = 2 * kl + ku + 1;
    std::vector<size_t> ipiv(n);
    auto at = [&](size_t i, size_t j) -> LD& { return AB[(kv + i - j) + ldab * j]; };
    size_t ju = 0;
    for (size_t j = 0; j < n; j++) {
        const size_t km = std::min(kl, n - 1 - j);
        size_t jp = 0;
        LD big = fabsl(at(j, j));
        for (size_t i = 1; i <= km; i++) if (fabsl(at(j + i, j)) > big) { big = fabsl(at(j + i, j)); jp = i; }
        ipiv[j] = j + jp;
        if (big == 0.0L) return false;
        ju = std::max(ju, std::min(j + ku + jp, n - 1));
        if (jp) for (size_t c = j; c <= ju; c++) std::swap(at(j, c), at(j + jp, c));
        const LD inv = 1.0L / at(j, j);
        for (size_t i = 1; i <= km; i++) at(j + i, j) *= inv;
        for (size_t c = j + 1; c <= ju; c++) {
            const LD u = at(j, c);
            if (u == 0.0L) continue;
            for (size_t i = 1; i <= km; i++) at(j + i, c) -= at(j + i, j) * u;
        }
    }
    for (size_t j = 0; j < n; j++) { /* L y = P b */
        if (ipiv[j] != j) std::swap(b[j], b[ipiv[j]]);
        const size_t km = std::min(kl, n - 1 - j);
        const LD bj = b[j];
        if (bj != 0.0L) for (size_t i = 1; i <= km; i++) b[j + i] -= at(j + i, j) * bj;
    }
    for (size_t j = n; j-- > 0;) { /* U x = y */
        b[j] /= at(j, j);
        const LD bj = b[j];
        const size_t lo = j > kv ? j - kv : 0;
        if (bj != 0.0L) for (size_t i = lo; i < j; i++) b[i] -= at(i, j) * bj;
    }
    return true;
}

// Least-squares / square solve by Householder QR with column pivoting (what RTR.colPivHouseholderQr().solve(ones) does in the
// reference's Ng acceleration, lineProducingSpecies.tpp:300-302); rank decided with Eigen's default threshold n * epsilon.
inline VecLD colpiv_qr_solve(std::vector<LD> A, VecLD b, const Size n) {
    std::vector<Size> perm(n);
    std::iota(perm.begin(), perm.end(), 0u);
    std::vector<LD> rdiag(n, 0.0L);
    Size rank = n;
    LD maxpivot = 0.0L;
    for (Size k = 0; k < n; k++) {
        Size best = k; LD bn = -1.0L;
        for (Size c = k; c < n; c++) { LD s = 0; for (Size r = k; r < n; r++) s += A[(size_t)r * n + c] * A[(size_t)r * n + c]; if (s > bn) { bn = s; best = c; } }
        if (best != k) { for (Size r = 0; r < n; r++) std::swap(A[(size_t)r * n + k], A[(size_t)r * n + best]); std::swap(perm[k], perm[best]); }
        LD norm = sqrtl(bn);
        if (k == 0) maxpivot = norm;
        if (norm <= maxpivot * (LD)n * std::numeric_limits<LD>::epsilon()) { rank = k; break; }
        const LD alpha = (A[(size_t)k * n + k] > 0) ? -norm : norm;
        std::vector<LD> v(n - k);
        for (Size r = k; r < n; r++) v[r - k] = A[(size_t)r * n + k];
        v[0] -= alpha;
        LD vn = 0; for (LD x : v) vn += x * x;
        if (vn > 0) {
            for (Size c = k; c < n; c++) {
                LD s = 0; for (Size r = k; r < n; r++) s += v[r - k] * A[(size_t)r * n + c];
                s = 2.0L * s / vn;
                for (Size r = k; r < n; r++) A[(size_t)r * n + c] -= s * v[r - k];
            }
            LD s = 0; for (Size r = k; r < n; r++) s += v[r - k] * b[r];
            s = 2.0L * s / vn;
            for (Size r = k; r < n; r++) b[r] -= s * v[r - k];
        }
        rdiag[k] = A[(size_t)k * n + k];
    }
    VecLD y(n, 0.0L), x(n, 0.0L);
    for (Size r = rank; r-- > 0;) {
        LD s = b[r];
        for (Size c = r + 1; c < rank; c++) s -= A[(size_t)r * n + c] * y[c];
        y[r] = s / A[(size_t)r * n + r];
    }
    for (Size k = 0; k < n; k++) x[perm[k]] = y[k];
    return x;
}

struct LineProducingSpecies { // reference lineProducingSpecies.hpp / .tpp
    const Parameters* parameters = nullptr;
    Size npoints = 0;
    Linedata linedata;
    std::vector<Real> quadrature_roots, quadrature_weights;
    VecLD population, population_prev1, population_prev2, population_prev3, trial_population, trial_population_prev;
    std::vector<Real> population_tot;
    std::vector<VecLD> populations, residuals;
    std::vector<Real> Jlin, Jeff, Jdif;      // [npoints][nrad]
    std::vector<Real> J;                     // [npoints][nrad] line-integrated mean intensity of the memory-sparse solver (lspec.J)
    std::vector<Real> J2_0, J2_1_Re, J2_1_Im, J2_2_Re, J2_2_Im; // [npoints][nrad] anisotropic radiation-field tensor (lineProducingSpecies.hpp:29-35)
    std::vector<Real> lambda_Ls;             // [npoints][nrad] the diagonal ALO element (nr == p), lambda.tpp:79-89
    std::vector<Size> lambda_cnt;            // [npoints][nrad] list length (0 or 1)
    // n_off_diag > 0: the whole (p, k) lists of Lambda::add_element (lambda.tpp:79-89) in insertion order, CSR over (p, k)
    bool banded = false;
    std::vector<size_t> ll_off;              // [npoints * nrad + 1]
    std::vector<Size> ll_nr;                 // Lambda.nr[p][k][m]
    std::vector<Real> ll_Ls;                 // Lambda.Ls[p][k][m]
    int stateq_iterations = 0;               // Krylov iterations of the last banded statistical-equilibrium solve
    LD stateq_residual = 0;                  // its final preconditioned residual (max norm, in units of the point's total population)
    Real fraction_not_converged = 0, relative_change_mean = 0, relative_change_max = 0;

    Size index(const Size p, const Size i) const { return i + p * linedata.nlev; }

    void resize(const Size np) {
        npoints = np;
        const size_t n = (size_t)np * linedata.nlev;
        population.assign(n, 0); population_prev1.assign(n, 0); population_prev2.assign(n, 0); population_prev3.assign(n, 0);
        trial_population.assign(n, 0); trial_population_prev.assign(n, 0);
        population_tot.assign(np, 0);
        Jlin.assign((size_t)np * linedata.nrad, 0); Jeff = Jlin; Jdif = Jlin; lambda_Ls = Jlin; J = Jlin;
        lambda_cnt.assign((size_t)np * linedata.nrad, 0);
    }
    // lineProducingSpecies.tpp:23-44
    Real get_emissivity(const Size p, const Size k) const {
        const Size i = index(p, linedata.irad[k]);
        const Real freq = linedata.frequency[k];
        return std::max(freq * HH_OVER_FOUR_PI * linedata.A[k] * population[i], parameters->min_line_emissivity);
    }
    Real get_opacity(const Size p, const Size k) const {
        const Size i = index(p, linedata.irad[k]), j = index(p, linedata.jrad[k]);
        const Real freq = linedata.frequency[k];
        return std::max(freq * HH_OVER_FOUR_PI * (population[j] * linedata.Ba[k] - population[i] * linedata.Bs[k]), parameters->min_line_opacity);
    }
    // lineProducingSpecies.tpp:49-69
    void update_using_LTE(const std::vector<Real>& abundance, const Size nspecs, const std::vector<Real>& temperature) {
        for (Size p = 0; p < npoints; p++) {
            population_tot[p] = abundance[(size_t)p * nspecs + linedata.num];
            Real partition_function = 0.0;
            for (Size i = 0; i < linedata.nlev; i++) {
                const Size ind = index(p, i);
                population[ind] = linedata.weight[i] * exp(-linedata.energy[i] / (KB * temperature[p]));
                partition_function += population[ind];
            }
            for (Size i = 0; i < linedata.nlev; i++) population[index(p, i)] *= population_tot[p] / partition_function;
        }
        populations.push_back(population);
    }
    // lineProducingSpecies.tpp:75-92
    void set_LTE(const std::vector<Real>& temperature, const Size p, const Size k) {
        const Size i = index(p, linedata.irad[k]), j = index(p, linedata.jrad[k]);
        Real population_tot_line = population[i] + population[j];
        Real partition_function  = 0.0;
        population[i] = linedata.weight[linedata.irad[k]] * exp(-linedata.energy[linedata.irad[k]] / (KB * temperature[p]));
        population[j] = linedata.weight[linedata.jrad[k]] * exp(-linedata.energy[linedata.jrad[k]] / (KB * temperature[p]));
        partition_function = population[i] + population[j];
        population[i] *= population_tot_line / partition_function;
        population[j] *= population_tot_line / partition_function;
    }
    // lineProducingSpecies.tpp:98-121
    void set_LTE_specific_levels(const std::vector<Real>& temperature, const Size p, const std::vector<char>& level_mask) {
        Real population_tot_levels = 0.0, partition_function = 0.0;
        for (Size i = 0; i < linedata.nlev; i++)
            if (level_mask[i]) {
                const Size ind = index(p, i);
                population_tot_levels += population[ind];
                population[ind] = linedata.weight[i] * exp(-linedata.energy[i] / (KB * temperature[p]));
                partition_function += population[ind];
            }
        for (Size i = 0; i < linedata.nlev; i++)
            if (level_mask[i]) population[index(p, i)] *= population_tot_levels / partition_function;
    }
    // lineProducingSpecies.tpp:128-165 (the reference reduces with OpenMP; the order of that sum is not fixed there either)
    void check_for_convergence(const Real pop_prec) {
        const Real weight = 1.0 / (npoints * linedata.nlev);
        Real fnc = 0.0, rcm = 0.0, rcmax = 0.0;
        for (Size p = 0; p < npoints; p++)
            for (Size i = 0; i < linedata.nlev; i++) {
                const Size ind = index(p, i);
                if (population[ind] > parameters->min_rel_pop_for_convergence * population_tot[p]) {
                    const Real relative_change = 2.0 * std::abs((population[ind] - population_prev1[ind]) / (population[ind] + population_prev1[ind]));
                    if (relative_change > pop_prec) fnc += weight;
                    rcm += (weight * relative_change);
                    rcmax = std::max(relative_change, rcmax);
                }
            }
        fraction_not_converged = fnc; relative_change_mean = rcm; relative_change_max = rcmax;
    }
    // lineProducingSpecies.tpp:171-201
    void check_for_convergence_trial(const Real pop_prec) {
        const Real weight = 1.0 / (npoints * linedata.nlev);
        Real fnc = 0.0, rcm = 0.0, rcmax = 0.0;
        for (Size p = 0; p < npoints; p++)
            for (Size i = 0; i < linedata.nlev; i++) {
                const Size ind = index(p, i);
                const Real relative_change = 2.0 * std::abs((trial_population[ind] - trial_population_prev[ind]) / (trial_population[ind] + trial_population_prev[ind]));
                if (relative_change > pop_prec) fnc += weight;
                rcm += (weight * relative_change);
                rcmax = std::max(relative_change, rcmax);
            }
        fraction_not_converged = fnc; relative_change_mean = rcm; relative_change_max = rcmax;
    }
    VecLD ng_combination(const Size order) const { // the shared core of update_using_acceleration[_trial], :287-309 / :340-361
        const Size popsize = (Size)populations.size(), ressize = (Size)residuals.size();
        std::vector<LD> RTR((size_t)order * order);
        for (Size i = 0; i < order; i++)
            for (Size j = 0; j < order; j++) RTR[(size_t)i * order + j] = dot(residuals[ressize - order + i], residuals[ressize - order + j]);
        VecLD coef = colpiv_qr_solve(RTR, VecLD(order, 1.0L), order);
        LD sum = 0; for (LD c : coef) sum += c;
        for (LD& c : coef) c /= sum;
        VecLD out(population.size(), 0.0L);
        for (Size i = 0; i < order; i++) {
            const VecLD& P = populations[popsize - 1 - i];
            const LD c = coef[order - 1 - i];
            for (size_t n = 0; n < out.size(); n++) out[n] += P[n] * c;
        }
        return out;
    }
    static VecLD diff(const VecLD& a, const VecLD& b) { VecLD r(a.size()); for (size_t i = 0; i < a.size(); i++) r[i] = a[i] - b[i]; return r; }
    // lineProducingSpecies.tpp:266-314
    void update_using_acceleration(const Size order) {
        population_prev3 = population_prev2; population_prev2 = population_prev1; population_prev1 = population;
        residuals.push_back(diff(population, populations.back()));
        populations.push_back(population);
        if (populations.size() < (residuals.size() + 1)) throw std::runtime_error("Error during ng acceleration; inconstent sizes.");
        if (residuals.size() < order) throw std::runtime_error("Error during ng acceleration; order of acceleration greather than data stored.");
        population = ng_combination(order);
        residuals.clear();
        populations.erase(populations.begin(), populations.end() - 1);
    }
    // lineProducingSpecies.tpp:316-371
    void update_using_acceleration_trial(const Size order) {
        residuals.push_back(diff(population, populations.back()));
        populations.push_back(population);
        trial_population_prev = trial_population;
        if (populations.size() < (residuals.size() + 1)) throw std::runtime_error("Error during ng acceleration; inconstent sizes.");
        if (populations.size() < order) throw std::runtime_error("Error during ng acceleration; order of acceleration greather than data stored.");
        trial_population = ng_combination(order);
        residuals.pop_back();
        populations.pop_back();
    }
    // lineProducingSpecies.tpp:377-581, block by block (see the header comment). Needs Jeff and the diagonal Lambda element.
    // `sparse`: update_using_statistical_equilibrium_sparse (lineProducingSpecies.tpp:589-705), the same per-point matrix solved by a
    // column-pivoted Householder QR as the reference does there.
    void update_using_statistical_equilibrium(const std::vector<Real>& abundance, const Size nspecs, const std::vector<Real>& temperature,
                                              const bool sparse = false) {
        if (sparse && parameters->n_off_diag != 0) throw std::runtime_error("parameters->n_off_diag != 0 so cannot use sparse update.");
        const Size nlev = linedata.nlev;
        population_prev3 = population_prev2; population_prev2 = population_prev1; population_prev1 = population;
        residuals.push_back(diff(population, populations.back()));
        populations.push_back(population);
        if (banded && !sparse) { solve_statistical_equilibrium_banded(abundance, nspecs, temperature); return; }
        bool failed = false;
#pragma omp parallel
        {
            std::vector<LD> M((size_t)nlev * nlev), y(nlev);
            std::vector<CollisionPartner> cps = linedata.colpar; /* interpolation buffers are per thread */
#pragma omp for schedule(static)
            for (Size p = 0; p < npoints; p++) {
                std::fill(M.begin(), M.end(), 0.0L);
                std::fill(y.begin(), y.end(), 0.0L);
                auto add = [&](Size r, Size c, LD v) { M[(size_t)r * nlev + c] += v; };
                for (Size k = 0; k < linedata.nrad; k++) {
                    const LD v_IJ = linedata.A[k] + linedata.Bs[k] * Jeff[(size_t)p * linedata.nrad + k];
                    const LD v_JI = linedata.Ba[k] * Jeff[(size_t)p * linedata.nrad + k];
                    const Size I = linedata.irad[k], J = linedata.jrad[k];
                    if (J != nlev - 1) { add(J, I, +v_IJ); add(J, J, -v_JI); }
                    if (I != nlev - 1) { add(I, J, +v_JI); add(I, I, -v_IJ); }
                }
                for (Size k = 0; k < linedata.nrad; k++)
                    for (Size m = 0; m < lambda_cnt[(size_t)p * linedata.nrad + k]; m++) { /* nr == p: the diagonal element */
                        const Real freq = linedata.frequency[k];
                        const LD v_IJ   = -lambda_Ls[(size_t)p * linedata.nrad + k] * get_opacity(p, k) / freq;
                        const Size I = linedata.irad[k], J = linedata.jrad[k];
                        if (J != nlev - 1) add(J, I, +v_IJ);
                        if (I != nlev - 1) add(I, I, -v_IJ);
                    }
                for (CollisionPartner& colpar : cps) {
                    Real abn = abundance[(size_t)p * nspecs + colpar.num_col_partner];
                    Real tmp = temperature[p];
                    colpar.adjust_abundance_for_ortho_or_para(tmp, abn);
                    colpar.interpolate_collision_coefficients(tmp);
                    for (Size k = 0; k < colpar.icol.size(); k++) {
                        const LD v_IJ = colpar.Cd_intpld[k] * abn;
                        const LD v_JI = colpar.Ce_intpld[k] * abn;
                        const Size I = colpar.icol[k], J = colpar.jcol[k];
                        if (J != nlev - 1) { add(J, I, +v_IJ); add(J, J, -v_JI); }
                        if (I != nlev - 1) { add(I, J, +v_JI); add(I, I, -v_IJ); }
                    }
                }
                for (Size i = 0; i < nlev; i++) add(nlev - 1, i, 1.0L);
                y[nlev - 1] = population_tot[p];
                if (sparse) {
                    const VecLD x = colpiv_qr_solve(M, y, nlev);
                    for (Size i = 0; i < nlev; i++) population[index(p, i)] = x[i];
                    continue;
                }
                if (!lu_solve(M, y, nlev)) { failed = true; continue; }
                for (Size i = 0; i < nlev; i++) population[index(p, i)] = y[i];
            }
        }
        if (failed) throw std::runtime_error("Eigen solver ERROR."); /* same text as lineProducingSpecies.tpp:551 */
    }
    // lineProducingSpecies.tpp:377-581 with a banded Lambda operator: the off-diagonal elements couple level irad of point nr to the
    // rate equation of level jrad at point p, so the system is one global sparse matrix A (npoints * nlev unknowns), which the
    // reference hands to Eigen::SparseLU in long double. Here: the same equations, assembled as the per-point nlev x nlev blocks D_p
    // plus a list of off-block entries O; unknowns scaled to fractions of the point's total population; solved by restarted GMRES
    // in long double on the block-Jacobi preconditioned system D^-1 (D + O) f = D^-1 y, started from the block-diagonal solution and
    // run until the preconditioned residual (i.e. the error estimate in fraction units) reaches long-double rounding level.
    void solve_statistical_equilibrium_banded(const std::vector<Real>& abundance, const Size nspecs, const std::vector<Real>& temperature) {
        const Size nlev = linedata.nlev, nrad = linedata.nrad;
        const size_t B = (size_t)nlev * nlev, n = (size_t)npoints * nlev;
        std::vector<LD> D(B * npoints, 0.0L);
        std::vector<Size> piv(n);
        VecLD y(n, 0.0L);
        std::vector<LD> scale(npoints);
        for (Size p = 0; p < npoints; p++) scale[p] = population_tot[p] > 0.0 ? (LD)population_tot[p] : 1.0L;
        // per-unknown column scale: the level's current population (the solution is expected within a few orders of it), floored at
        // 1e-60 of the point's total -- the Lambda elements of lines between nearly empty levels are huge numbers multiplying tiny ones
        VecLD cscale(n);
        for (Size p = 0; p < npoints; p++)
            for (Size i = 0; i < nlev; i++)
                cscale[(size_t)p * nlev + i] = std::max((LD)fabsl(population[index(p, i)]), 1.0e-60L * scale[p]);
        // ---- local parts of the blocks (radiative, collisional, conservation row), as in the block-diagonal path ----
#pragma omp parallel
        {
            std::vector<CollisionPartner> cps = linedata.colpar;
#pragma omp for schedule(static)
            for (Size p = 0; p < npoints; p++) {
                LD* M = &D[B * p];
                auto add = [&](Size r, Size c, LD v) { M[(size_t)r * nlev + c] += v; };
                for (Size k = 0; k < nrad; k++) {
                    const LD v_IJ = linedata.A[k] + linedata.Bs[k] * Jeff[(size_t)p * nrad + k];
                    const LD v_JI = linedata.Ba[k] * Jeff[(size_t)p * nrad + k];
                    const Size I = linedata.irad[k], J = linedata.jrad[k];
                    if (J != nlev - 1) { add(J, I, +v_IJ); add(J, J, -v_JI); }
                    if (I != nlev - 1) { add(I, J, +v_JI); add(I, I, -v_IJ); }
                }
                for (CollisionPartner& colpar : cps) {
                    Real abn = abundance[(size_t)p * nspecs + colpar.num_col_partner];
                    Real tmp = temperature[p];
                    colpar.adjust_abundance_for_ortho_or_para(tmp, abn);
                    colpar.interpolate_collision_coefficients(tmp);
                    for (Size k = 0; k < colpar.icol.size(); k++) {
                        const LD v_IJ = colpar.Cd_intpld[k] * abn;
                        const LD v_JI = colpar.Ce_intpld[k] * abn;
                        const Size I = colpar.icol[k], J = colpar.jcol[k];
                        if (J != nlev - 1) { add(J, I, +v_IJ); add(J, J, -v_JI); }
                        if (I != nlev - 1) { add(I, J, +v_JI); add(I, I, -v_IJ); }
                    }
                }
                for (Size i = 0; i < nlev; i++) add(nlev - 1, i, 1.0L);
                y[index(p, nlev - 1)] = population_tot[p];
            }
        }
        // ---- the approximated Lambda operator (:439-460): (J = (p, jrad), I = (nr, irad)) += v, (I, I) -= v with I at point nr ----
        struct Off { Size row, nr, col; LD v; }; /* row level at point p, column level at point nr */
        std::vector<size_t> o_off(npoints + 1, 0);
        std::vector<Off> offs;
        for (Size p = 0; p < npoints; p++) {
            for (Size k = 0; k < nrad; k++) {
                const Real freq = linedata.frequency[k];
                const Real opac = get_opacity(p, k);
                for (size_t m = ll_off[(size_t)p * nrad + k]; m < ll_off[(size_t)p * nrad + k + 1]; m++) {
                    const Size nr   = ll_nr[m];
                    const LD v_IJ   = -ll_Ls[m] * opac / freq;
                    const Size I = linedata.irad[k], J = linedata.jrad[k];
                    if (J != nlev - 1) {
                        if (nr == p) D[B * p + (size_t)J * nlev + I] += v_IJ;
                        else offs.push_back(Off{J, nr, I, v_IJ});
                    }
                    if (I != nlev - 1) D[B * nr + (size_t)I * nlev + I] -= v_IJ; /* a diagonal element of point nr's block */
                }
            }
            o_off[p + 1] = offs.size();
        }
        // ---- unknowns as fractions of the point's total: column scaling of every block and off-block entry; right-hand side -> 1 ----
#pragma omp parallel for schedule(static)
        for (Size p = 0; p < npoints; p++) {
            for (Size r = 0; r < nlev; r++)
                for (Size c = 0; c < nlev; c++) D[B * p + (size_t)r * nlev + c] *= cscale[(size_t)p * nlev + c];
            for (size_t e = o_off[p]; e < o_off[p + 1]; e++) offs[e].v *= cscale[(size_t)offs[e].nr * nlev + offs[e].col];
        }
        std::vector<LD> Dm = D; /* the blocks themselves (matrix-vector products); D becomes their LU factors */
        bool failed = false;
#pragma omp parallel for schedule(static)
        for (Size p = 0; p < npoints; p++)
            if (!lu_factor(&D[B * p], &piv[(size_t)p * nlev], nlev)) failed = true;
        if (failed) throw std::runtime_error("Eigen solver ERROR."); /* same text as lineProducingSpecies.tpp:551 */
        auto precond = [&](VecLD& v) { /* v <- D^-1 v */
#pragma omp parallel for schedule(static)
            for (Size p = 0; p < npoints; p++) lu_apply(&D[B * p], &piv[(size_t)p * nlev], &v[(size_t)p * nlev], nlev);
        };
        auto applyA = [&](const VecLD& f, VecLD& out, VecLD* mag) { /* out <- (D + O) f; mag <- (|D| + |O|) |f| when asked for */
#pragma omp parallel for schedule(static)
            for (Size p = 0; p < npoints; p++) {
                const LD* M = &Dm[B * p];
                for (Size r = 0; r < nlev; r++) {
                    LD t = 0.0L, a = 0.0L;
                    for (Size c = 0; c < nlev; c++) { const LD x = M[(size_t)r * nlev + c] * f[(size_t)p * nlev + c]; t += x; a += fabsl(x); }
                    out[(size_t)p * nlev + r] = t;
                    if (mag) (*mag)[(size_t)p * nlev + r] = a;
                }
                for (size_t e = o_off[p]; e < o_off[p + 1]; e++) {
                    const LD x = offs[e].v * f[(size_t)offs[e].nr * nlev + offs[e].col];
                    out[(size_t)p * nlev + offs[e].row] += x;
                    if (mag) (*mag)[(size_t)p * nlev + offs[e].row] += fabsl(x);
                }
            }
        };
        auto dotp = [&](const VecLD& a, const VecLD& b) {
            LD t = 0.0L;
#pragma omp parallel for schedule(static) reduction(+ : t)
            for (size_t i = 0; i < n; i++) t += a[i] * b[i];
            return t;
        };
        // ---- iterative refinement on the ORIGINAL equations: r = y - A f in long double, component-wise backward error
        //      omega = max_i |r_i| / ((|A| |f|)_i + |y_i|) (Oettli-Prager: f solves a system whose every entry is off by at most omega
        //      relative -- what a backward-stable direct solver delivers at omega ~ epsilon); the correction A d = r is solved by one
        //      GMRES(restart) cycle on the block-Jacobi preconditioned system. The blocks can be badly conditioned (a point collects the
        //      Lambda elements of every list that names it), which is why convergence is NOT judged on the preconditioned residual. ----
        // ---- start of the refinement. When the lists only name points within a few indices of their own (1D models: the points are
        //      ordered along the radius and a ray couples a shell to its n_off_diag neighbours on either side) the system is a narrow
        //      band matrix and a direct band LU with partial pivoting gives the solution outright; otherwise (3D point clouds: any
        //      ordering leaves a wide band) the block-diagonal solution starts the Krylov refinement. ----
        size_t pband = 0;
        for (Size p = 0; p < npoints; p++)
            for (size_t e = o_off[p]; e < o_off[p + 1]; e++) pband = std::max<size_t>(pband, offs[e].nr > p ? offs[e].nr - p : p - offs[e].nr);
        const size_t kl = (pband + 1) * nlev - 1;
        const char* want = getenv("MGBH_STATEQ_SOLVER"); /* "direct" | "krylov": tests */
        const bool direct = want ? std::string(want) == "direct" : ((pband + 1) * nlev <= 512 && (3 * kl + 1) * n * sizeof(LD) <= (4ull << 30));
        VecLD f = y;
        if (direct) {
            const size_t ldab = 3 * kl + 1, kv = 2 * kl;
            std::vector<LD> AB(ldab * n, 0.0L);
            for (Size p = 0; p < npoints; p++) {
                for (Size r = 0; r < nlev; r++)
                    for (Size c = 0; c < nlev; c++) {
                        const size_t i = (size_t)p * nlev + r, j = (size_t)p * nlev + c;
                        AB[(kv + i - j) + ldab * j] = Dm[B * p + (size_t)r * nlev + c];
                    }
                for (size_t e = o_off[p]; e < o_off[p + 1]; e++) {
                    const size_t i = (size_t)p * nlev + offs[e].row, j = (size_t)offs[e].nr * nlev + offs[e].col;
                    AB[(kv + i - j) + ldab * j] += offs[e].v;
                }
            }
            if (!band_lu_solve(AB, n, kl, kl, f)) throw std::runtime_error("Eigen solver ERROR.");
        } else {
            precond(f); /* the block-diagonal solution */
        }
        VecLD r(n), w(n), mag(n);
        if (getenv("MGBH_STATEQ_DEBUG")) {
            LD fmax = 0.0L; size_t bad = 0, where = 0;
            for (size_t i = 0; i < n; i++) { if (!std::isfinite((double)f[i])) bad++; else if (fabsl(f[i]) > fmax) { fmax = fabsl(f[i]); where = i; } }
            LD omax = 0.0L; for (const auto& o : offs) omax = std::max(omax, fabsl(o.v));
            fprintf(stderr, "[stateq] %s start (point band %zu): max |f| %.3Le at point %zu level %zu, %zu non-finite; %zu off-block entries, largest %.3Le\n", direct ? "direct band LU" : "block-diagonal", pband, fmax, where / nlev, where % nlev, bad, offs.size(), omax);
        }
        const int restart = 40, max_outer = 80;
        const LD eps = std::numeric_limits<LD>::epsilon(), tol = 8.0L * eps;
        std::vector<VecLD> V;
        std::vector<std::vector<LD>> H(restart + 1, std::vector<LD>(restart, 0.0L));
        std::vector<LD> cs(restart), sn(restart), g(restart + 1);
        stateq_iterations = 0;
        LD omega = 0.0L, best = std::numeric_limits<LD>::infinity();
        VecLD fbest;
        int since_best = 0;
        const bool debug = getenv("MGBH_STATEQ_DEBUG") != nullptr;
        for (int outer = 0; outer <= max_outer; outer++) {
            applyA(f, w, &mag);
            omega = 0.0L;
            size_t worst_row = 0;
            for (size_t i = 0; i < n; i++) {
                r[i] = y[i] - w[i];
                const LD den = mag[i] + fabsl(y[i]);
                if (r[i] != 0.0L) { const LD o_i = den > 0.0L ? fabsl(r[i]) / den : std::numeric_limits<LD>::infinity(); if (o_i > omega) { omega = o_i; worst_row = i; } }
            }
            if (debug) fprintf(stderr, "[stateq]   worst row: point %zu level %zu, y %.3Le, (|A||f|) %.3Le, r %.3Le\n", worst_row / nlev, worst_row % nlev, y[worst_row], mag[worst_row], r[worst_row]);
            if (debug) fprintf(stderr, "[stateq] refinement step %d, Krylov iterations %d, backward error %.3Le\n", outer, stateq_iterations, omega);
            if (omega < best) { best = omega; fbest = f; since_best = 0; } else since_best++;
            if (best <= tol || outer == max_outer || since_best >= 3) break;
            precond(r); /* right-hand side of the preconditioned correction equation */
            const LD beta = sqrtl(dotp(r, r));
            if (!(beta > 0.0L)) break;
            V.assign(1, r);
            for (LD& x : V[0]) x /= beta;
            std::fill(g.begin(), g.end(), 0.0L);
            g[0] = beta;
            int j = 0;
            for (; j < restart; j++) {
                applyA(V[j], w, nullptr);
                precond(w);
                stateq_iterations++;
                for (int pass = 0; pass < 2; pass++) /* modified Gram-Schmidt, twice */
                    for (int i = 0; i <= j; i++) {
                        const LD h = dotp(w, V[i]);
                        if (pass == 0) H[i][j] = h; else H[i][j] += h;
                        const LD* vi = V[i].data();
#pragma omp parallel for schedule(static)
                        for (size_t q = 0; q < n; q++) w[q] -= h * vi[q];
                    }
                H[j + 1][j] = sqrtl(dotp(w, w));
                for (int i = 0; i < j; i++) { /* the previous Givens rotations on the new column */
                    const LD t = cs[i] * H[i][j] + sn[i] * H[i + 1][j];
                    H[i + 1][j] = -sn[i] * H[i][j] + cs[i] * H[i + 1][j];
                    H[i][j] = t;
                }
                const LD den = sqrtl(H[j][j] * H[j][j] + H[j + 1][j] * H[j + 1][j]);
                if (den == 0.0L) { j++; break; }
                cs[j] = H[j][j] / den; sn[j] = H[j + 1][j] / den;
                H[j][j] = den;
                g[j + 1] = -sn[j] * g[j];
                g[j] = cs[j] * g[j];
                const bool breakdown = H[j + 1][j] == 0.0L;
                if (!breakdown) { V.push_back(w); for (LD& x : V.back()) x /= H[j + 1][j]; }
                H[j + 1][j] = 0.0L;
                if (breakdown || fabsl(g[j + 1]) <= 1.0e-10L * beta) { j++; break; }
            }
            if (debug) fprintf(stderr, "[stateq]   GMRES cycle: %d iterations, preconditioned residual reduced to %.3Le of its start\n", j, fabsl(g[j]) / beta);
            std::vector<LD> yk(j, 0.0L);
            for (int i = j; i-- > 0;) {
                LD t = g[i];
                for (int k = i + 1; k < j; k++) t -= H[i][k] * yk[k];
                yk[i] = t / H[i][i];
            }
            for (int i = 0; i < j; i++) {
                const LD c = yk[i];
                const LD* vi = V[i].data();
#pragma omp parallel for schedule(static)
                for (size_t q = 0; q < n; q++) f[q] += c * vi[q];
            }
        }
        if (!fbest.empty()) f.swap(fbest);
        stateq_residual = best;
        if (!(best <= 1.0e-15L)) {
            char msg[400];
            snprintf(msg, sizeof msg, "Eigen solver ERROR. (banded Lambda operator: the statistical-equilibrium system of %zu unknowns did not converge, "
                     "backward error %.2Le after %d Krylov iterations; the block-Jacobi preconditioned refinement needs blocks that are regular on "
                     "their own -- n_off_diag = 0 or a 1D model, which is solved directly, always works)", n, best, stateq_iterations);
            throw std::runtime_error(msg); /* the text starts like lineProducingSpecies.tpp:551 */
        }
        for (Size p = 0; p < npoints; p++)
            for (Size i = 0; i < nlev; i++) population[index(p, i)] = f[(size_t)p * nlev + i] * cscale[(size_t)p * nlev + i];
    }
    // lineProducingSpecies.tpp:711-802
    void correct_negative_populations(const std::vector<Real>& temperature) {
        const Size nlev = linedata.nlev;
        for (Size p = 0; p < npoints; p++)
            if (population_tot[p] > 0.0) {
                Real total_positive_population = 0.0;
                for (Size i = 0; i < nlev; i++) {
                    if (population[p * nlev + i] < 0.0) population[p * nlev + i] = 0.0;
                    else total_positive_population += population[p * nlev + i];
                }
                for (Size i = 0; i < nlev; i++) population[p * nlev + i] = population[p * nlev + i] * population_tot[p] / total_positive_population;
            }
        for (Size p = 0; p < npoints; p++) {
            std::vector<char> modified_level(nlev, false);
            Size next_n_modified_levels = 0, curr_n_modified_levels = 0;
            auto inverted = [&](Size k) {
                const Size i = index(p, linedata.irad[k]), j = index(p, linedata.jrad[k]);
                return population[j] < linedata.Bs[k] / linedata.Ba[k] * population[i] * parameters->population_inversion_fraction;
            };
            do {
                curr_n_modified_levels = next_n_modified_levels;
                bool population_inversion_at_point = false;
                for (Size k = linedata.nrad - 1; k != static_cast<Size>(-1); k--)
                    if (inverted(k)) { set_LTE(temperature, p, k); modified_level[linedata.irad[k]] = true; modified_level[linedata.jrad[k]] = true; population_inversion_at_point = true; }
                for (Size k = 0; k < linedata.nrad; k++)
                    if (inverted(k)) { set_LTE(temperature, p, k); modified_level[linedata.irad[k]] = true; modified_level[linedata.jrad[k]] = true; population_inversion_at_point = true; }
                if (population_inversion_at_point) set_LTE_specific_levels(temperature, p, modified_level);
                next_n_modified_levels = std::accumulate(modified_level.begin(), modified_level.end(), static_cast<Size>(0));
            } while (next_n_modified_levels != curr_n_modified_levels);
        }
    }
};

struct Image { // reference src/model/image/image.hpp (the fields the new imager fills)
    Size nx = 0, ny = 0, nfreqs = 0;
    bool optical_depth = false; // ImageType::OpticalDepth (image.hpp) instead of Intensity
    size_t pix_begin = 0, pix_end = 0; // the pixels this rank computed (all of them on one GPU; the other rows of I stay 0)
    double ray_direction[3] = {0, 0, 0};
    std::vector<double> ImX, ImY, freqs;
    // Image.I [nx*ny][nfreqs]: 210 MB for 512^2 x 100 -- allocated WITHOUT the zero fill of std::vector (the device writes every row of this
    // rank's pixel range; only the rows outside it are cleared)
    struct Pixels {
        struct Free { void operator()(double* q) const { free(q); } };
        std::unique_ptr<double[], Free> p;
        size_t n = 0;
        // 2 MB-aligned, transparent huge pages requested, pages touched by all host threads: first-touching 210 MB from one thread
        // (inside the D2H copy) cost 85 ms, more than the image's kernels
        void allocate(size_t count) {
            const size_t huge = (size_t)2 << 20, bytes = ((count * sizeof(double) + huge - 1) / huge) * huge;
            void* q = nullptr;
            if (posix_memalign(&q, huge, bytes ? bytes : huge) != 0) throw std::bad_alloc();
            madvise(q, bytes, MADV_HUGEPAGE);
            p.reset(static_cast<double*>(q)); n = count;
            char* c = static_cast<char*>(q);
#pragma omp parallel for schedule(static)
            for (long long off = 0; off < (long long)bytes; off += 4096) c[off] = 0;
        }
        double* data() { return p.get(); }
        const double* data() const { return p.get(); }
        size_t size() const { return n; }
    } I;
};

enum SpectralDiscretisation { SD_None, SD_Lines, SD_Image }; // model.hpp

class Model {
  public:
    Parameters parameters;
    // geometry (reference src/model/geometry): points, rays, boundary
    Size npoints = 0, nrays = 0, nboundary = 0, nspecs = 0;
    std::vector<double> position, velocity;      // [npoints][3]
    std::vector<Size> n_neighbors, neighbors;
    std::vector<double> direction, weight;       // [nrays][3], [nrays]
    std::vector<Size> antipod, boundary2point, point2boundary;
    std::vector<int32_t> boundary_condition;
    std::vector<double> boundary_temperature;
    // thermodynamics, chemistry
    std::vector<Real> temperature_gas, vturb2, abundance; // abundance [npoints][nspecs]
    // lines
    std::vector<LineProducingSpecies> lineProducingSpecies;
    std::vector<Size> nrad_cum;
    std::vector<Real> line;                                  // [nlines] (Lines::line, lines.cpp:28-38)
    std::vector<Real> emissivity, opacity, inverse_width;    // [npoints][nlines]
    // radiation
    Size nfreqs = 0;
    std::vector<Real> nu, J;                                 // [npoints][nfreqs]
    std::vector<Size> corresponding_line;
    SpectralDiscretisation spectralDiscretisation = SD_None;
    std::vector<Image> images;
    std::vector<double> error_mean, error_max;
    mgb_stats last_stats{};
    mgb_image_stats last_image_stats{};
    double seconds_radiation_field = 0.0, seconds_statistical_equilibrium = 0.0; // Timer totals of model.cpp:706-724

    Model() = default;
    Model(const Model&) = delete;
    Model& operator=(const Model&) = delete;
    ~Model() { if (ctx) mgb_destroy(ctx); }

    // ---- B200 knobs (no reference counterpart) ----
    // true (default): statistical equilibrium, Ng acceleration, compute_Jeff and the emissivity/opacity refresh of
    // compute_level_populations[_sparse] run on the GPU (mgb_lp_*), an iteration moves nothing across PCIe; false: the host twin
    // below (long double, the reference's code path), with the line state uploaded and J / Lambda downloaded every iteration.
    bool device_level_populations = true;
    // One process per GPU: origins are partitioned over the ranks (cost-balanced), J / Lambda slices all-gathered over NCCL before
    // the level-population update (replaces Radiation::MPI_reduce_J / Lambda::MPI_gather, solver.tpp:530-539). id128 from
    // nccl_unique_id() of rank 0, distributed by the application.
    static std::string nccl_unique_id() {
        std::string id(128, '\0');
        if (mgb_nccl_unique_id(&id[0]) != MGB_OK) throw std::runtime_error("mgb_nccl_unique_id failed (libnccl.so.2 not loadable?)");
        return id;
    }
    void set_communicator(const int nranks_, const int rank_, const std::string& id128) {
        if (nranks_ < 1 || rank_ < 0 || rank_ >= nranks_ || id128.size() != 128) throw std::runtime_error("set_communicator: bad arguments");
        nranks = nranks_; rank = rank_; comm_id = id128; comm_dirty = true;
    }
    int nranks = 1, rank = 0;
    Size o_begin = 0, o_end = 0; // this rank's origins in the last radiation-field call

    Size nlines() const { Size n = 0; for (auto& l : lineProducingSpecies) n += l.linedata.nrad; return n; }
    Size nlspecs() const { return (Size)lineProducingSpecies.size(); }
    Size nquads() const { return lineProducingSpecies.empty() ? 0 : (Size)lineProducingSpecies[0].quadrature_roots.size(); }
    int device = 0;

    // ---- set-up (what Model::read / the pybind setters establish in the reference) ----
    void set_geometry(Size np, const double* pos, const double* vel, const Size* n_nbs, const Size* nbs, size_t n_nbs_tot, Size nr,
                      const double* dir, const double* wgt, const Size* anti, Size nb, const Size* b2p, const int32_t* bcond,
                      const double* btemp) {
        npoints = np; nrays = nr; nboundary = nb; use_adaptive_directions = false;
        position.assign(pos, pos + 3 * (size_t)np); velocity.assign(vel, vel + 3 * (size_t)np);
        n_neighbors.assign(n_nbs, n_nbs + np); neighbors.assign(nbs, nbs + n_nbs_tot);
        direction.assign(dir, dir + 3 * (size_t)nr); weight.assign(wgt, wgt + nr); antipod.assign(anti, anti + nr);
        boundary2point.assign(b2p, b2p + nb);
        point2boundary.assign(np, np); /* boundary.cpp:28-37: npoints marks an interior point */
        for (Size b = 0; b < nb; b++) point2boundary[b2p[b]] = b;
        boundary_condition.assign(bcond, bcond + nb); boundary_temperature.assign(btemp, btemp + nb);
        dirty_geometry = true;
    }
    // Rays::use_adaptive_directions (rays.hpp:23-31): dir [npoints][nrays][3], wgt [npoints][nrays]; after set_geometry (point 0's set)
    std::vector<double> direction_adaptive, weight_adaptive;
    bool use_adaptive_directions = false;
    void set_adaptive_directions(const double* dir, const double* wgt) {
        direction_adaptive.assign(dir, dir + 3 * (size_t)npoints * nrays); weight_adaptive.assign(wgt, wgt + (size_t)npoints * nrays);
        use_adaptive_directions = true; dirty_geometry = true;
    }
    void upload_geometry() {
        check(mgb_set_geometry(ctx, npoints, position.data(), velocity.data(), n_neighbors.data(), neighbors.data(), neighbors.size(), nrays,
                               direction.data(), weight.data(), antipod.data(), nboundary, point2boundary.data(), boundary_condition.data(),
                               boundary_temperature.data(), parameters.spherical_symmetry ? 1 : 0), "mgb_set_geometry");
        if (use_adaptive_directions) check(mgb_set_adaptive_directions(ctx, direction_adaptive.data(), weight_adaptive.data()), "mgb_set_adaptive_directions");
    }
    void set_thermodynamics(const double* T, const double* vt2) {
        temperature_gas.assign(T, T + npoints); vturb2.assign(vt2, vt2 + npoints);
        dirty_thermo = true;
    }
    void set_abundance(const double* ab, Size nsp) { nspecs = nsp; abundance.assign(ab, ab + (size_t)npoints * nsp); }
    // Lines::read (lines.cpp:11-69): line list in species-major order
    void add_species(LineProducingSpecies&& ls) {
        lineProducingSpecies.push_back(std::move(ls));
        finalize_lines();
    }
    void finalize_lines() {
        nrad_cum.clear(); line.clear();
        Size c = 0;
        for (auto& l : lineProducingSpecies) {
            l.parameters = &parameters; /* the vector may have moved */
            nrad_cum.push_back(c); c += l.linedata.nrad;
            for (Size k = 0; k < l.linedata.nrad; k++) line.push_back(l.linedata.frequency[k]);
            if (l.npoints != npoints) l.resize(npoints);
        }
        emissivity.assign((size_t)npoints * c, 0); opacity = emissivity; inverse_width = emissivity;
        dirty_lines = true;
    }

    Real profile_width(const Real inverse_mass, const Size p) const { // thermodynamics.tpp:32-35
        return sqrt(TWO_KB_OVER_AMU_CC_SQUARED * inverse_mass * temperature_gas[p] + vturb2[p]);
    }
    // ---- Lines::set_inverse_width (lines.tpp:45-59) ----
    int compute_inverse_line_widths() {
        const Size L = nlines();
        state_serial++;
        for (Size p = 0; p < npoints; p++)
            for (Size l = 0; l < nlspecs(); l++)
                for (Size k = 0; k < lineProducingSpecies[l].linedata.nrad; k++) {
                    const Real invr_mass = lineProducingSpecies[l].linedata.inverse_mass;
                    const Real frequency = lineProducingSpecies[l].linedata.frequency[k];
                    inverse_width[(size_t)p * L + nrad_cum[l] + k] = (Real)1.0 / (frequency * profile_width(invr_mass, p));
                }
        return 0;
    }
    // ---- Lines::set_emissivity_and_opacity (lines.tpp:29-41) ----
    void set_emissivity_and_opacity() {
        const Size L = nlines();
        state_serial++;
#pragma omp parallel for schedule(static)
        for (Size p = 0; p < npoints; p++)
            for (Size l = 0; l < nlspecs(); l++)
                for (Size k = 0; k < lineProducingSpecies[l].linedata.nrad; k++) {
                    emissivity[(size_t)p * L + nrad_cum[l] + k] = lineProducingSpecies[l].get_emissivity(p, k);
                    opacity[(size_t)p * L + nrad_cum[l] + k]    = lineProducingSpecies[l].get_opacity(p, k);
                }
    }
    // ---- Model::compute_spectral_discretisation() (model.cpp:70-161): f = lid*nquads + z ----
    int compute_spectral_discretisation() {
        const Size nq = nquads(), L = nlines();
        nfreqs = L * nq;
        nu.assign((size_t)npoints * nfreqs, 0); corresponding_line.assign(nfreqs, 0);
        for (Size p = 0; p < npoints; p++) {
            Size index0 = 0, index1 = 0;
            for (Size l = 0; l < nlspecs(); l++) {
                const Real inverse_mass = lineProducingSpecies[l].linedata.inverse_mass;
                for (Size k = 0; k < lineProducingSpecies[l].linedata.nrad; k++) {
                    const Real freqs_line = line[index0];
                    const Real width      = freqs_line * profile_width(inverse_mass, p);
                    for (Size z = 0; z < nq; z++) {
                        const Real root = lineProducingSpecies[l].quadrature_roots[z];
                        nu[(size_t)p * nfreqs + index1] = freqs_line + width * root;
                        index1++;
                    }
                    index0++;
                }
            }
        }
        for (Size f = 0; f < nfreqs; f++) corresponding_line[f] = f / nq;
        spectralDiscretisation = SD_Lines;
        dirty_freqs = true;
        return 0;
    }
    // ---- Model::compute_spectral_discretisation(nu_min, nu_max, n) (model.cpp:257-300): the same bins at every point ----
    int compute_spectral_discretisation(const Real nu_min, const Real nu_max, const Size n_image_freqs) {
        if (n_image_freqs < 1) throw std::runtime_error("At least a single frequency is needed to compute a spectral discretization.");
        nfreqs = n_image_freqs;
        image_nu.assign(nfreqs, 0);
        if (n_image_freqs == 1) image_nu[0] = (Real)(nu_min + nu_max) / 2.0;
        else {
            const long double dnu = (nu_max - nu_min) / (n_image_freqs - 1);
            for (Size f = 0; f < n_image_freqs; f++) image_nu[f] = (Real)(nu_min + f * dnu);
        }
        spectralDiscretisation = SD_Image;
        return 0;
    }
    int compute_spectral_discretisation(const Real nu_min, const Real nu_max) { return compute_spectral_discretisation(nu_min, nu_max, nlines() * nquads()); }
    std::vector<Real> image_nu;

    // ---- Lines::iteration_using_LTE (lines.cpp:82-91) ----
    int compute_LTE_level_populations() {
        for (auto& lspec : lineProducingSpecies) lspec.update_using_LTE(abundance, nspecs, temperature_gas);
        set_emissivity_and_opacity();
        return 0;
    }

    int approx() const { // model.cpp:366-389
        if (parameters.one_line_approximation) return MGB_APPROX_ONELINE;
        if (parameters.sum_opacity_emissivity_over_all_lines) return MGB_APPROX_NONE;
        return MGB_APPROX_CLOSELINES;
    }
    // ---- Model::compute_radiation_field_feautrier_order_2 (model.cpp:353-390) -> the C ABI ----
    int compute_radiation_field_feautrier_order_2() {
        run_radiation_field(true);
        if (parameters.n_off_diag) fetch_J_and_lambda_lists(); /* banded Lambda: the lists couple neighbouring points (host statistical equilibrium) */
        else fetch_J_and_lambda();
        return 0;
    }
    // The banded lists [npoints][nlines][cap] (unused tail ignored) -> every species' CSR lists; lambda_Ls / lambda_cnt keep the
    // diagonal element and the list length, as mgb_get_lambda reports them for n_off_diag > 0.
    void set_lambda_lists(const double* val, const uint32_t* nr, const uint32_t* cnt, const size_t cap) {
        const Size L = nlines();
        for (Size l = 0; l < nlspecs(); l++) {
            auto& ls = lineProducingSpecies[l];
            const Size nrad = ls.linedata.nrad;
            ls.banded = true;
            ls.ll_off.assign((size_t)npoints * nrad + 1, 0);
            ls.ll_nr.clear(); ls.ll_Ls.clear();
            for (Size p = 0; p < npoints; p++)
                for (Size k = 0; k < nrad; k++) {
                    const size_t row = (size_t)p * L + nrad_cum[l] + k;
                    const Size c = cnt[row];
                    if (c > cap) throw std::runtime_error("Lambda list longer than its capacity.");
                    ls.lambda_cnt[(size_t)p * nrad + k] = c;
                    ls.lambda_Ls[(size_t)p * nrad + k]  = 0.0;
                    for (Size m = 0; m < c; m++) {
                        ls.ll_nr.push_back(nr[row * cap + m]);
                        ls.ll_Ls.push_back(val[row * cap + m]);
                        if (nr[row * cap + m] >= npoints) throw std::runtime_error("Lambda list names a point outside the model.");
                        if (nr[row * cap + m] == p) ls.lambda_Ls[(size_t)p * nrad + k] = val[row * cap + m];
                    }
                    ls.ll_off[(size_t)p * nrad + k + 1] = ls.ll_nr.size();
                }
        }
    }
    void fetch_J_and_lambda_lists() {
        J.resize((size_t)npoints * nfreqs);
        check(mgb_get_J(ctx, 0, npoints, J.data()), "mgb_get_J");
        uint32_t cap = 0;
        check(mgb_lambda_capacity(ctx, &cap), "mgb_lambda_capacity");
        const size_t rows = (size_t)npoints * nlines();
        std::vector<double> val(rows * cap);
        std::vector<uint32_t> nr(rows * cap), cnt(rows);
        check(mgb_get_lambda_lists(ctx, 0, npoints, val.data(), nr.data(), cnt.data()), "mgb_get_lambda_lists");
        set_lambda_lists(val.data(), nr.data(), cnt.data(), cap);
    }
    void fetch_J_and_lambda() {
        const Size L = nlines();
        for (auto& ls : lineProducingSpecies) ls.banded = false;
        J.resize((size_t)npoints * nfreqs);
        std::vector<double> lam((size_t)npoints * L);
        std::vector<uint32_t> cnt((size_t)npoints * L);
        check(mgb_get_J(ctx, 0, npoints, J.data()), "mgb_get_J");
        check(mgb_get_lambda(ctx, 0, npoints, lam.data(), cnt.data()), "mgb_get_lambda");
        for (Size l = 0; l < nlspecs(); l++) { /* Lambda lists per species: lambda.Ls[p][k], nr == p */
            auto& ls = lineProducingSpecies[l];
            const Size nrad = ls.linedata.nrad;
            for (Size p = 0; p < npoints; p++)
                for (Size k = 0; k < nrad; k++) {
                    ls.lambda_Ls[(size_t)p * nrad + k]  = lam[(size_t)p * L + nrad_cum[l] + k];
                    ls.lambda_cnt[(size_t)p * nrad + k] = cnt[(size_t)p * L + nrad_cum[l] + k];
                }
        }
    }
    // ---- Model::compute_radiation_field_shortchar_order_0 (model.cpp:313-349; Solver::solve_shortchar_order_0, solver.tpp:150-235):
    //      J and the local Lambda element from the short-characteristics solver (mgb_run_shortchar_order_0) ----
    int compute_radiation_field_shortchar_order_0() {
        shortchar = true;
        try { run_radiation_field(true); } catch (...) { shortchar = false; throw; }
        shortchar = false;
        fetch_J_and_lambda();
        return 0;
    }
    // ---- Model::compute_radiation_field_feautrier_order_2_sparse (model.cpp:474-510; Solver::solve_feautrier_order_2_sparse,
    //      solver.tpp:297-363): the same solves, but only J integrated over each line (lspec.J) and Lambda leave the device ----
    int compute_radiation_field_feautrier_order_2_sparse() {
        if (parameters.n_off_diag) throw std::runtime_error("parameters->n_off_diag != 0 so cannot use sparse update.");
        run_radiation_field(true);
        const Size L = nlines();
        std::vector<double> jl((size_t)npoints * L), lam((size_t)npoints * L);
        std::vector<uint32_t> cnt((size_t)npoints * L);
        check(mgb_get_J_lines(ctx, 0, npoints, jl.data()), "mgb_get_J_lines");
        check(mgb_get_lambda(ctx, 0, npoints, lam.data(), cnt.data()), "mgb_get_lambda");
        for (Size l = 0; l < nlspecs(); l++) {
            auto& ls = lineProducingSpecies[l];
            const Size nrad = ls.linedata.nrad;
            for (Size p = 0; p < npoints; p++)
                for (Size k = 0; k < nrad; k++) {
                    ls.J[(size_t)p * nrad + k]          = jl[(size_t)p * L + nrad_cum[l] + k];
                    ls.lambda_Ls[(size_t)p * nrad + k]  = lam[(size_t)p * L + nrad_cum[l] + k];
                    ls.lambda_cnt[(size_t)p * nrad + k] = cnt[(size_t)p * L + nrad_cum[l] + k];
                }
        }
        return 0;
    }
    // ---- Model::compute_radiation_field_feautrier_order_2_anis (model.cpp:435-470; Solver::solve_feautrier_order_2_anis, solver.tpp:365-456):
    //      the same solves accumulated into lspec.J and the anisotropic tensor elements lspec.J2_0 .. J2_2_Im ----
    int compute_radiation_field_feautrier_order_2_anis() {
        if (nranks > 1) throw std::runtime_error("compute_radiation_field_feautrier_order_2_anis: one GPU only (the tensor is not part of the all-gather)");
        anisotropic = true;
        try { run_radiation_field(true); } catch (...) { anisotropic = false; throw; }
        anisotropic = false;
        const Size L = nlines();
        std::vector<double> a((size_t)6 * npoints * L);
        check(mgb_get_J_anis(ctx, 0, npoints, a.data()), "mgb_get_J_anis");
        const size_t plane = (size_t)npoints * L;
        for (Size l = 0; l < nlspecs(); l++) {
            auto& ls = lineProducingSpecies[l];
            const Size nrad = ls.linedata.nrad;
            std::vector<Real>* dst[6] = {&ls.J, &ls.J2_0, &ls.J2_1_Re, &ls.J2_1_Im, &ls.J2_2_Re, &ls.J2_2_Im};
            for (int q = 0; q < 6; q++) {
                dst[q]->assign((size_t)npoints * nrad, 0.0);
                for (Size p = 0; p < npoints; p++)
                    for (Size k = 0; k < nrad; k++) (*dst[q])[(size_t)p * nrad + k] = a[q * plane + (size_t)p * L + nrad_cum[l] + k];
            }
        }
        return 0;
    }
    // ---- Model::compute_Jeff_sparse (model.cpp:549-570) ----
    int compute_Jeff_sparse() {
        for (Size l = 0; l < nlspecs(); l++) {
            auto& lspec = lineProducingSpecies[l];
            const Size nrad = lspec.linedata.nrad;
#pragma omp parallel for schedule(static)
            for (Size p = 0; p < npoints; p++)
                for (Size k = 0; k < nrad; k++) {
                    double diff = 0.0;
                    for (Size m = 0; m < lspec.lambda_cnt[(size_t)p * nrad + k]; m++) {
                        const Size I = lspec.index(p, lspec.linedata.irad[k]);
                        diff += lspec.lambda_Ls[(size_t)p * nrad + k] * lspec.population[I];
                    }
                    lspec.Jlin[(size_t)p * nrad + k] = lspec.J[(size_t)p * nrad + k];
                    lspec.Jeff[(size_t)p * nrad + k] = lspec.Jlin[(size_t)p * nrad + k] - HH_OVER_FOUR_PI * diff;
                }
        }
        return 0;
    }
    // ---- Model::compute_Jeff (model.cpp:514-545) ----
    int compute_Jeff() {
        const Size nq = nquads();
        for (Size l = 0; l < nlspecs(); l++) {
            auto& lspec = lineProducingSpecies[l];
            const Size nrad = lspec.linedata.nrad;
#pragma omp parallel for schedule(static)
            for (Size p = 0; p < npoints; p++)
                for (Size k = 0; k < nrad; k++) {
                    Real jl = 0.0;
                    for (Size z = 0; z < nq; z++) jl += lspec.quadrature_weights[z] * J[(size_t)p * nfreqs + (size_t)(nrad_cum[l] + k) * nq + z];
                    double diff = 0.0;
                    if (lspec.banded) { /* the whole list: neighbouring points' populations enter (model.cpp:531-536) */
                        for (size_t m = lspec.ll_off[(size_t)p * nrad + k]; m < lspec.ll_off[(size_t)p * nrad + k + 1]; m++) {
                            const Size I = lspec.index(lspec.ll_nr[m], lspec.linedata.irad[k]);
                            diff += lspec.ll_Ls[m] * lspec.population[I];
                        }
                    } else
                    for (Size m = 0; m < lspec.lambda_cnt[(size_t)p * nrad + k]; m++) {
                        const Size I = lspec.index(p, lspec.linedata.irad[k]);
                        diff += lspec.lambda_Ls[(size_t)p * nrad + k] * lspec.population[I];
                    }
                    lspec.Jlin[(size_t)p * nrad + k] = jl;
                    lspec.Jeff[(size_t)p * nrad + k] = jl - HH_OVER_FOUR_PI * diff;
                    lspec.Jdif[(size_t)p * nrad + k] = HH_OVER_FOUR_PI * diff;
                }
        }
        return 0;
    }
    // ---- Lines::iteration_using_statistical_equilibrium (lines.cpp:126-137) ----
    int compute_level_populations_from_stateq() {
        for (auto& lspec : lineProducingSpecies) {
            lspec.update_using_statistical_equilibrium(abundance, nspecs, temperature_gas);
            lspec.correct_negative_populations(temperature_gas);
            lspec.check_for_convergence(parameters.pop_prec);
        }
        set_emissivity_and_opacity();
        return 0;
    }
    // ---- Lines::iteration_using_statistical_equilibrium_sparse (lines.cpp:139-150) ----
    int compute_level_populations_from_stateq_sparse() {
        for (auto& lspec : lineProducingSpecies) {
            lspec.update_using_statistical_equilibrium(abundance, nspecs, temperature_gas, true);
            lspec.correct_negative_populations(temperature_gas);
            lspec.check_for_convergence(parameters.pop_prec);
        }
        set_emissivity_and_opacity();
        return 0;
    }
    void iteration_using_Ng_acceleration(const Size order) { // lines.cpp:106-117
        for (auto& lspec : lineProducingSpecies) {
            lspec.update_using_acceleration(order);
            lspec.correct_negative_populations(temperature_gas);
            lspec.check_for_convergence(parameters.pop_prec);
        }
        set_emissivity_and_opacity();
    }
    // model.cpp:580-653
    std::tuple<bool, Size> ng_acceleration_criterion(const bool use_Ng_acceleration, const Size prior_normal_iterations) {
        const Size skip_N_its = parameters.Ng_acceleration_remove_N_its;
        if (!parameters.use_adaptive_Ng_acceleration) {
            if (!use_Ng_acceleration) return std::make_tuple(false, 0u);
            if (prior_normal_iterations == (parameters.Ng_acceleration_mem_limit + skip_N_its)) return std::make_tuple(true, parameters.Ng_acceleration_mem_limit);
            return std::make_tuple(false, 0u);
        }
        if (!use_Ng_acceleration || prior_normal_iterations < parameters.adaptive_Ng_acceleration_min_order + skip_N_its) return std::make_tuple(false, 0u);
        auto sum_fnc = [&]() {
            double s = 0.0;
            for (auto& l : lineProducingSpecies) s += parameters.adaptive_Ng_acceleration_use_max_criterion ? l.relative_change_max : l.relative_change_mean;
            return s;
        };
        const double sum_fnc_curr = sum_fnc();
        const Size nb_prev_iterations = prior_normal_iterations;
        for (auto& lspec : lineProducingSpecies) { /* Lines::trial_iteration_using_adaptive_Ng_acceleration, lines.cpp:119-124 */
            lspec.update_using_acceleration_trial(nb_prev_iterations - skip_N_its);
            lspec.check_for_convergence_trial(parameters.pop_prec);
        }
        const double sum_fnc_ng = sum_fnc();
        if ((sum_fnc_ng < sum_fnc_curr) || prior_normal_iterations == (parameters.Ng_acceleration_mem_limit + skip_N_its))
            return std::make_tuple(true, nb_prev_iterations - skip_N_its);
        return std::make_tuple(false, 0u);
    }
    // ---- Model::compute_level_populations (model.cpp:655-748) ----
    int compute_level_populations(const bool use_Ng_acceleration, const long max_niterations) { return level_population_loop(use_Ng_acceleration, max_niterations, false); }
    // ---- Model::compute_level_populations_sparse (model.cpp:756-850): the same loop over the memory-sparse radiation field ----
    int compute_level_populations_sparse(const bool use_Ng_acceleration, const long max_niterations) { return level_population_loop(use_Ng_acceleration, max_niterations, true); }
    // ---- Model::compute_level_populations_shortchar (model.cpp:859-953): the same loop over the short-characteristics radiation field ----
    int compute_level_populations_shortchar(const bool use_Ng_acceleration, const long max_niterations) {
        shortchar = true;
        int it;
        try { it = level_population_loop(use_Ng_acceleration, max_niterations, false); } catch (...) { shortchar = false; throw; }
        shortchar = false;
        return it;
    }
    int level_population_loop(const bool use_Ng_acceleration, const long max_niterations, const bool sparse) {
        if (spectralDiscretisation != SD_Lines) throw std::runtime_error("Spectral discretisation was not set for Lines!");
        /* the device twin of the level-population half solves block by block: a banded Lambda operator (one global system) takes the host path */
        if (device_level_populations && !(parameters.n_off_diag && !sparse && !shortchar)) return level_population_loop_device(use_Ng_acceleration, max_niterations, sparse);
        int iteration = 0, iteration_normal = 0;
        error_mean.clear(); error_max.clear();
        bool some_not_converged = true;
        while (some_not_converged && (iteration < max_niterations)) {
            const auto decision = ng_acceleration_criterion(use_Ng_acceleration, (Size)iteration_normal);
            if (std::get<0>(decision)) {
                iteration_using_Ng_acceleration(std::get<1>(decision));
                iteration_normal = 0;
            } else {
                iteration++;
                some_not_converged = false;
                const double t0 = now();
                if (sparse) { compute_radiation_field_feautrier_order_2_sparse(); compute_Jeff_sparse(); }
                else if (shortchar) { run_radiation_field(true); fetch_J_and_lambda(); compute_Jeff(); }
                else { compute_radiation_field_feautrier_order_2(); compute_Jeff(); }
                const double t1 = now();
                if (sparse) compute_level_populations_from_stateq_sparse(); else compute_level_populations_from_stateq();
                const double t2 = now();
                seconds_radiation_field += t1 - t0; seconds_statistical_equilibrium += t2 - t1;
                iteration_normal++;
                for (auto& l : lineProducingSpecies) {
                    error_mean.push_back(l.relative_change_mean);
                    error_max.push_back(l.relative_change_max);
                    if (l.fraction_not_converged > 1.0 - parameters.convergence_fraction) some_not_converged = true;
                }
            }
        }
        return iteration;
    }
    // ---- the same loop with the level-population half on the GPU (mgb_lp_*, magritte_b200/csrc/mgb_levelpop.cu) ----
    // Control flow is Model::compute_level_populations[_sparse] (model.cpp:655-748, 756-850) and ng_acceleration_criterion (:580-653)
    // unchanged; each step calls the device twin of the Lines:: / LineProducingSpecies:: routine the reference calls. Host copies of
    // populations, Jlin/Jeff, the line state, J (or lspec.J) and Lambda are refreshed when the loop ends.
    int level_population_loop_device(const bool use_Ng_acceleration, const long max_niterations, const bool sparse) {
        if (parameters.n_off_diag && !shortchar) throw std::runtime_error("parameters->n_off_diag != 0 so cannot use sparse update."); /* block-diagonal only */
        sync_static(true);
        sync_levelpop();
        int iteration = 0, iteration_normal = 0;
        error_mean.clear(); error_max.clear();
        bool some_not_converged = true;
        std::vector<mgb_lp_convergence> conv(nlspecs());
        auto store_conv = [&]() {
            for (Size l = 0; l < nlspecs(); l++) {
                lineProducingSpecies[l].fraction_not_converged = conv[l].fraction_not_converged;
                lineProducingSpecies[l].relative_change_mean   = conv[l].relative_change_mean;
                lineProducingSpecies[l].relative_change_max    = conv[l].relative_change_max;
            }
        };
        auto sum_fnc = [&]() {
            double s = 0.0;
            for (auto& l : lineProducingSpecies) s += parameters.adaptive_Ng_acceleration_use_max_criterion ? l.relative_change_max : l.relative_change_mean;
            return s;
        };
        auto ng_step = [&](const Size order, const bool trial) { /* update_using_acceleration[_trial] for every species */
            for (Size l = 0; l < nlspecs(); l++) {
                std::vector<double> ghi((size_t)order * order), glo((size_t)order * order);
                check(mgb_lp_ng_gram(ctx, l, order, trial ? 1 : 0, ghi.data(), glo.data()), "mgb_lp_ng_gram");
                std::vector<LD> RTR((size_t)order * order);
                for (size_t i = 0; i < RTR.size(); i++) RTR[i] = (LD)ghi[i] + (LD)glo[i];
                VecLD coef = colpiv_qr_solve(RTR, VecLD(order, 1.0L), order);
                LD sum = 0; for (LD c : coef) sum += c;
                std::vector<double> chi(order), clo(order);
                for (Size i = 0; i < order; i++) { const LD c = coef[i] / sum; chi[i] = (double)c; clo[i] = (double)(c - (LD)chi[i]); }
                check(mgb_lp_ng_combine(ctx, l, order, trial ? 1 : 0, chi.data(), clo.data()), "mgb_lp_ng_combine");
            }
        };
        while (some_not_converged && (iteration < max_niterations)) {
            bool do_ng = false; Size order = 0;
            const Size skip_N_its = parameters.Ng_acceleration_remove_N_its, prior = (Size)iteration_normal;
            if (!parameters.use_adaptive_Ng_acceleration) {
                if (use_Ng_acceleration && prior == parameters.Ng_acceleration_mem_limit + skip_N_its) { do_ng = true; order = parameters.Ng_acceleration_mem_limit; }
            } else if (use_Ng_acceleration && prior >= parameters.adaptive_Ng_acceleration_min_order + skip_N_its) {
                const double sum_fnc_curr = sum_fnc();
                ng_step(prior - skip_N_its, true);
                check(mgb_lp_trial_convergence(ctx, parameters.pop_prec, conv.data()), "mgb_lp_trial_convergence");
                store_conv();
                const double sum_fnc_ng = sum_fnc();
                if ((sum_fnc_ng < sum_fnc_curr) || prior == (parameters.Ng_acceleration_mem_limit + skip_N_its)) { do_ng = true; order = prior - skip_N_its; }
            }
            if (do_ng) {
                ng_step(order, false);
                check(mgb_lp_finish_update(ctx, parameters.pop_prec, conv.data()), "mgb_lp_finish_update");
                store_conv();
                iteration_normal = 0;
            } else {
                iteration++;
                some_not_converged = false;
                const double t0 = now();
                run_radiation_field(false);
                check(mgb_lp_compute_Jeff(ctx, sparse ? 1 : 0), "mgb_lp_compute_Jeff");
                const double t1 = now();
                check(mgb_lp_update_statistical_equilibrium(ctx, use_Ng_acceleration ? 1 : 0, parameters.pop_prec, conv.data(), nullptr), "mgb_lp_update_statistical_equilibrium");
                store_conv();
                const double t2 = now();
                seconds_radiation_field += t1 - t0; seconds_statistical_equilibrium += t2 - t1;
                iteration_normal++;
                for (auto& l : lineProducingSpecies) {
                    error_mean.push_back(l.relative_change_mean);
                    error_max.push_back(l.relative_change_max);
                    if (l.fraction_not_converged > 1.0 - parameters.convergence_fraction) some_not_converged = true;
                }
            }
        }
        download_levelpop(sparse);
        return iteration;
    }

    // ---- Model::compute_image_new(rx, ry, rz, Nxpix, Nypix) (model.cpp:999-1035) ----
    int compute_image_new(const double rx, const double ry, const double rz, const Size Nxpix, const Size Nypix) { return image_new(false, rx, ry, rz, Nxpix, Nypix); }
    // ---- Model::compute_image_optical_depth_new(raydir, Nxpix, Nypix) (model.cpp:1102-1127): Image.I holds optical depths ----
    int compute_image_optical_depth_new(const double rx, const double ry, const double rz, const Size Nxpix, const Size Nypix) { return image_new(true, rx, ry, rz, Nxpix, Nypix); }
    // ---- Model::set_column (model.cpp:1143-1154; Solver::set_column / get_column, solver.tpp:2953-2991): column [nrays][npoints] ----
    std::vector<double> density, column; // Model::density [npoints], Model::column [nrays][npoints] (model.hpp:118-119)
    int set_column() {
        if (density.size() != npoints) throw std::runtime_error("Model::density must hold one value per point.");
        if (!ctx) {
            const int rc = mgb_create(device, &ctx);
            if (rc != MGB_OK) { ctx = nullptr; throw std::runtime_error(std::string("mgb_create failed: ") + mgb_last_error(nullptr)); }
        }
        if (dirty_geometry) {
            upload_geometry();
            dirty_geometry = false; dirty_thermo = dirty_lines = dirty_freqs = true; state_uploaded = 0;
        }
        column.resize((size_t)nrays * npoints);
        check(mgb_set_column(ctx, density.data(), 0, npoints, column.data(), nullptr), "mgb_set_column");
        return 0;
    }
    int image_new(const bool optical_depth, const double rx, const double ry, const double rz, const Size Nxpix, const Size Nypix) {
        if (rx * rx + ry * ry + rz * rz == 0.0)
            throw std::runtime_error("The given ray direction vector does not point in a direction. Please use a non-zero (normed) direction vector to generate an image.");
        if (parameters.one_line_approximation)
            throw std::runtime_error(optical_depth ? "One line approximation is currently not supported for imaging." : "One line approximation is not supported for imaging.");
        sync_static(false);
        Image im;
        im.optical_depth = optical_depth;
        im.nx = Nxpix; im.ny = Nypix; im.nfreqs = nfreqs;
        const double n = 1 / std::sqrt(rx * rx + ry * ry + rz * rz);
        im.ray_direction[0] = rx * n; im.ray_direction[1] = ry * n; im.ray_direction[2] = rz * n;
        const size_t npix = (size_t)Nxpix * Nypix;
        im.ImX.resize(npix); im.ImY.resize(npix); im.I.allocate(npix * nfreqs);
        const double dir[3] = {rx, ry, rz};
        const bool per_point = spectralDiscretisation == SD_Lines;
        im.freqs = per_point ? std::vector<double>() : image_nu;
        if (state_uploaded != state_serial) { /* the imager reads the line state: upload it only when the host arrays changed since the last upload */
            std::vector<double> ptot((size_t)nlspecs() * npoints);
            for (Size l = 0; l < nlspecs(); l++) std::copy(lineProducingSpecies[l].population_tot.begin(), lineProducingSpecies[l].population_tot.end(), ptot.begin() + (size_t)l * npoints);
            check(mgb_set_line_state(ctx, emissivity.data(), opacity.data(), inverse_width.data(), ptot.data()), "mgb_set_line_state");
            state_uploaded = state_serial;
        }
        /* one process per GPU: every pixel is independent (solver.tpp:591-622), rank r images the r-th share of the pixels */
        im.pix_begin = npix * (size_t)rank / (size_t)nranks; im.pix_end = npix * ((size_t)rank + 1) / (size_t)nranks;
        check((optical_depth ? mgb_image_optical_depth : mgb_image_feautrier_order_2)(
                  ctx, approx(), (int)parameters.dimension, dir, Nxpix, Nypix, nfreqs, per_point ? nu.data() : image_nu.data(), per_point ? 1 : 0,
                  im.pix_begin, im.pix_end, im.ImX.data(), im.ImY.data(), im.I.data() + im.pix_begin * nfreqs, &last_image_stats),
              optical_depth ? "mgb_image_optical_depth" : "mgb_image_feautrier_order_2");
        if (im.pix_begin) std::fill(im.I.data(), im.I.data() + im.pix_begin * nfreqs, 0.0);
        if (im.pix_end < npix) std::fill(im.I.data() + im.pix_end * nfreqs, im.I.data() + npix * nfreqs, 0.0);
        images.push_back(std::move(im));
        return 0;
    }

  private:
    mgb_ctx* ctx = nullptr;
    bool dirty_geometry = true, dirty_thermo = true, dirty_lines = true, dirty_freqs = true, dirty_species = true, comm_dirty = false;
    std::string comm_id;
    bool have_partition = false, anisotropic = false, shortchar = false;
    unsigned long state_serial = 1, state_uploaded = 0; // host emissivity / opacity / inverse_width changed <-> last upload for the imager
    // One radiation-field pass on this rank's origins (+ the all-gather): Solver::solve_feautrier_order_2[_sparse] (solver.tpp:297-363, 458-540).
    // upload_state: the host's emissivity / opacity / inverse_width go up first (the host-twin path); otherwise the device table is current.
    void run_radiation_field(const bool upload_state) {
        sync_static(true);
        if (upload_state) {
            std::vector<double> ptot((size_t)nlspecs() * npoints);
            for (Size l = 0; l < nlspecs(); l++) std::copy(lineProducingSpecies[l].population_tot.begin(), lineProducingSpecies[l].population_tot.end(), ptot.begin() + (size_t)l * npoints);
            check(mgb_set_line_state(ctx, emissivity.data(), opacity.data(), inverse_width.data(), ptot.data()), "mgb_set_line_state");
        }
        if (nranks == 1) { o_begin = 0; o_end = npoints; }
        else if (!have_partition) {
            check(mgb_partition_balanced(ctx, nranks, rank, &o_begin, &o_end), "mgb_partition_balanced");
            have_partition = true;
        }
        const int rc = shortchar ? mgb_run_shortchar_order_0(ctx, approx(), parameters.min_dtau, o_begin, o_end, &last_stats)
                                 : mgb_run_feautrier_order_2(ctx, approx(), o_begin, o_end, &last_stats);
        if (nranks > 1) { /* agree before the collective: a rank that failed must not leave its peers inside the all-gather */
            int all = rc;
            check(mgb_comm_agree(ctx, rc, &all), "mgb_comm_agree");
            check(rc, "mgb_run_feautrier_order_2");
            if (all != MGB_OK) throw std::runtime_error("mgb_run_feautrier_order_2 failed on another rank (" + std::to_string(all) + ")");
            check(mgb_allgather_results(ctx), "mgb_allgather_results");
        } else check(rc, "mgb_run_feautrier_order_2");
    }
    static void split_ld(const VecLD& v, std::vector<double>& hi, std::vector<double>& lo) {
        hi.resize(v.size()); lo.resize(v.size());
        for (size_t i = 0; i < v.size(); i++) { hi[i] = (double)v[i]; lo[i] = (double)(v[i] - (LD)hi[i]); }
    }
    // Static level-population inputs (once) and the current host populations + line state (every loop start: the host may have set them).
    void sync_levelpop() {
        if (dirty_species) {
            check(mgb_lp_set_parameters(ctx, (double)parameters.min_line_opacity, (double)parameters.min_line_emissivity,
                                        parameters.population_inversion_fraction, parameters.min_rel_pop_for_convergence), "mgb_lp_set_parameters");
            for (Size l = 0; l < nlspecs(); l++) {
                const Linedata& ld = lineProducingSpecies[l].linedata;
                check(mgb_lp_set_species(ctx, l, ld.num, ld.nlev, ld.nrad, ld.irad.data(), ld.jrad.data(), ld.energy.data(), ld.weight.data(),
                                         ld.frequency.data(), ld.A.data(), ld.Ba.data(), ld.Bs.data()), "mgb_lp_set_species");
                for (const CollisionPartner& cp : ld.colpar) {
                    const Size ntmp = (Size)cp.tmp.size(), ncol = (Size)cp.icol.size();
                    std::vector<double> Ce((size_t)ntmp * ncol), Cd((size_t)ntmp * ncol);
                    for (Size t = 0; t < ntmp; t++) { std::copy(cp.Ce[t].begin(), cp.Ce[t].end(), Ce.begin() + (size_t)t * ncol); std::copy(cp.Cd[t].begin(), cp.Cd[t].end(), Cd.begin() + (size_t)t * ncol); }
                    const int oop = cp.orth_or_para_H2 == "o" ? 1 : (cp.orth_or_para_H2 == "p" ? 2 : 0);
                    check(mgb_lp_add_collision_partner(ctx, l, cp.num_col_partner, oop, ntmp, ncol, cp.icol.data(), cp.jcol.data(), cp.tmp.data(), Ce.data(), Cd.data()),
                          "mgb_lp_add_collision_partner");
                }
            }
            check(mgb_lp_set_abundance(ctx, nspecs, abundance.data()), "mgb_lp_set_abundance");
            dirty_species = false;
        }
        std::vector<double> hi, lo;
        for (Size l = 0; l < nlspecs(); l++) {
            split_ld(lineProducingSpecies[l].population, hi, lo);
            check(mgb_lp_set_populations(ctx, l, hi.data(), lo.data(), lineProducingSpecies[l].population_tot.data()), "mgb_lp_set_populations");
        }
        std::vector<double> ptot((size_t)nlspecs() * npoints);
        for (Size l = 0; l < nlspecs(); l++) std::copy(lineProducingSpecies[l].population_tot.begin(), lineProducingSpecies[l].population_tot.end(), ptot.begin() + (size_t)l * npoints);
        check(mgb_set_line_state(ctx, emissivity.data(), opacity.data(), inverse_width.data(), ptot.data()), "mgb_set_line_state");
    }
    void download_levelpop(const bool sparse) {
        const Size L = nlines();
        std::vector<double> hi, lo;
        for (Size l = 0; l < nlspecs(); l++) {
            auto& ls = lineProducingSpecies[l];
            hi.resize(ls.population.size()); lo.resize(ls.population.size());
            check(mgb_lp_get_populations(ctx, l, hi.data(), lo.data()), "mgb_lp_get_populations");
            for (size_t i = 0; i < hi.size(); i++) ls.population[i] = (LD)hi[i] + (LD)lo[i];
            check(mgb_lp_get_Jeff(ctx, l, ls.Jlin.data(), ls.Jeff.data()), "mgb_lp_get_Jeff");
        }
        check(mgb_lp_get_line_state(ctx, emissivity.data(), opacity.data(), nullptr), "mgb_lp_get_line_state");
        state_serial++;
        std::vector<double> jl((size_t)npoints * L), lam((size_t)npoints * L);
        std::vector<uint32_t> cnt((size_t)npoints * L);
        if (!sparse) { J.resize((size_t)npoints * nfreqs); check(mgb_get_J(ctx, 0, npoints, J.data()), "mgb_get_J"); }
        check(mgb_get_J_lines(ctx, 0, npoints, jl.data()), "mgb_get_J_lines");
        check(mgb_get_lambda(ctx, 0, npoints, lam.data(), cnt.data()), "mgb_get_lambda");
        for (Size l = 0; l < nlspecs(); l++) {
            auto& ls = lineProducingSpecies[l];
            const Size nrad = ls.linedata.nrad;
            for (Size p = 0; p < npoints; p++)
                for (Size k = 0; k < nrad; k++) {
                    if (sparse) ls.J[(size_t)p * nrad + k] = jl[(size_t)p * L + nrad_cum[l] + k];
                    ls.lambda_Ls[(size_t)p * nrad + k]  = lam[(size_t)p * L + nrad_cum[l] + k];
                    ls.lambda_cnt[(size_t)p * nrad + k] = cnt[(size_t)p * L + nrad_cum[l] + k];
                }
        }
    }
    static double now() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + 1e-9 * ts.tv_nsec; }
    void check(int rc, const char* what) {
        if (rc != MGB_OK) throw std::runtime_error(std::string(what) + " failed (" + std::to_string(rc) + "): " + mgb_last_error(ctx));
    }
    // Device context created lazily; static tables uploaded once and again only when the host arrays changed (SURVEY.md 8b "Ownership").
    void sync_static(const bool need_line_freqs) {
        if (!ctx) {
            const int rc = mgb_create(device, &ctx);
            if (rc != MGB_OK) { ctx = nullptr; throw std::runtime_error(std::string("mgb_create failed: ") + mgb_last_error(nullptr)); }
        }
        mgb_params prm;
        mgb_params_default(&prm);
        prm.min_opacity = parameters.min_opacity; prm.max_distance_opacity_contribution = parameters.max_distance_opacity_contribution;
        prm.max_interpolation_diff = parameters.max_interpolation_diff; prm.max_width_fraction = parameters.max_width_fraction;
        prm.n_off_diag = parameters.n_off_diag; prm.store_intensities = 0; prm.anisotropic = anisotropic ? 1 : 0;
        check(mgb_set_params(ctx, &prm), "mgb_set_params");
        if (dirty_geometry) {
            upload_geometry();
            dirty_geometry = false; dirty_thermo = dirty_lines = dirty_freqs = true; have_partition = false; state_uploaded = 0;
        }
        if (dirty_thermo) { check(mgb_set_thermodynamics(ctx, temperature_gas.data(), vturb2.data()), "mgb_set_thermodynamics"); dirty_thermo = false; }
        if (dirty_lines) {
            std::vector<uint32_t> nrad; std::vector<double> im, A, qw;
            for (auto& l : lineProducingSpecies) {
                nrad.push_back(l.linedata.nrad); im.push_back(l.linedata.inverse_mass);
                A.insert(A.end(), l.linedata.A.begin(), l.linedata.A.end());
                qw.insert(qw.end(), l.quadrature_weights.begin(), l.quadrature_weights.end());
            }
            check(mgb_set_lines(ctx, nlspecs(), nrad.data(), im.data(), line.data(), A.data(), nquads(), qw.data()), "mgb_set_lines");
            dirty_lines = false; dirty_freqs = true; dirty_species = true; state_uploaded = 0;
        }
        if (comm_dirty) {
            check(mgb_comm_init(ctx, nranks, rank, comm_id.data()), "mgb_comm_init");
            comm_dirty = false; have_partition = false;
        }
        if (need_line_freqs) {
            if (spectralDiscretisation != SD_Lines) throw std::runtime_error("Spectral discretisation was not set for Lines!");
            if (dirty_freqs) { check(mgb_set_frequencies(ctx, nfreqs, nu.data(), corresponding_line.data()), "mgb_set_frequencies"); dirty_freqs = false; }
        }
    }
};

} // namespace mgbh
