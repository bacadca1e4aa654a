// bindings.cpp -- pybind11 module `magritte_b200.core`: the Python face of mgbh::Model, with the method names of
// magritte.core.Model for the Feautrier order-2 path (reference src/bindings/pybindings.cpp:123-271). Arrays go in and out
// as numpy arrays in the reference's layouts through flat setters / getters; the reference's nested attribute tree
// (model.geometry.points.position.set(...), buffer-protocol containers, Io) is the Python facade magritte_b200/core_compat.py.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include "mgb_host.hpp"

namespace py = pybind11;
using namespace mgbh;
template <class T> using arr = py::array_t<T, py::array::c_style | py::array::forcecast>;

template <class T> static std::vector<T> vec(const arr<T>& a) { return std::vector<T>(a.data(), a.data() + a.size()); }
static py::array_t<double> out2(const std::vector<double>& v, size_t r, size_t c) {
    py::array_t<double> a({r, c});
    std::copy(v.begin(), v.begin() + r * c, a.mutable_data());
    return a;
}

PYBIND11_MODULE(core, m) {
    m.doc() = "magritte_b200.core: B200 drop-in for the Feautrier order-2 path of magritte.core";
    m.attr("CC") = CC; m.attr("HH") = HH; m.attr("KB") = KB; m.attr("AMU") = AMU; m.attr("T_CMB") = T_CMB;

    py::class_<Parameters>(m, "Parameters")
        .def_readwrite("n_off_diag", &Parameters::n_off_diag)
        .def_readwrite("max_width_fraction", &Parameters::max_width_fraction)
        .def_readwrite("convergence_fraction", &Parameters::convergence_fraction)
        .def_readwrite("min_rel_pop_for_convergence", &Parameters::min_rel_pop_for_convergence)
        .def_readwrite("pop_prec", &Parameters::pop_prec)
        .def_readwrite("min_opacity", &Parameters::min_opacity)
        .def_readwrite("min_dtau", &Parameters::min_dtau)
        .def_property("min_line_opacity", [](const Parameters& p) { return (double)p.min_line_opacity; }, [](Parameters& p, double v) { p.min_line_opacity = v; })
        .def_property("min_line_emissivity", [](const Parameters& p) { return (double)p.min_line_emissivity; }, [](Parameters& p, double v) { p.min_line_emissivity = v; })
        .def_readwrite("max_interpolation_diff", &Parameters::max_interpolation_diff)
        .def_readwrite("max_distance_opacity_contribution", &Parameters::max_distance_opacity_contribution)
        .def_readwrite("population_inversion_fraction", &Parameters::population_inversion_fraction)
        .def_readwrite("use_adaptive_Ng_acceleration", &Parameters::use_adaptive_Ng_acceleration)
        .def_readwrite("adaptive_Ng_acceleration_use_max_criterion", &Parameters::adaptive_Ng_acceleration_use_max_criterion)
        .def_readwrite("Ng_acceleration_mem_limit", &Parameters::Ng_acceleration_mem_limit)
        .def_readwrite("adaptive_Ng_acceleration_min_order", &Parameters::adaptive_Ng_acceleration_min_order)
        .def_readwrite("Ng_acceleration_remove_N_its", &Parameters::Ng_acceleration_remove_N_its)
        .def_readwrite("one_line_approximation", &Parameters::one_line_approximation)
        .def_readwrite("sum_opacity_emissivity_over_all_lines", &Parameters::sum_opacity_emissivity_over_all_lines)
        .def_readwrite("spherical_symmetry", &Parameters::spherical_symmetry)
        .def_readwrite("dimension", &Parameters::dimension);

    py::class_<Image>(m, "Image")
        .def_readonly("nfreqs", &Image::nfreqs)
        .def_readonly("optical_depth", &Image::optical_depth)
        .def_readonly("pix_begin", &Image::pix_begin)
        .def_readonly("pix_end", &Image::pix_end)
        .def_property_readonly("ImX", [](const Image& i) { return py::array_t<double>(i.ImX.size(), i.ImX.data()); })
        .def_property_readonly("ImY", [](const Image& i) { return py::array_t<double>(i.ImY.size(), i.ImY.data()); })
        .def_property_readonly("freqs", [](const Image& i) { return py::array_t<double>(i.freqs.size(), i.freqs.data()); })
        .def_property_readonly("I", [](const Image& i) {
            py::array_t<double> a({(size_t)i.nx * i.ny, (size_t)i.nfreqs});
            std::copy(i.I.data(), i.I.data() + i.I.size(), a.mutable_data());
            return a;
        });

    py::class_<Model>(m, "Model")
        .def(py::init<>())
        .def_readwrite("device", &Model::device)
        .def_readwrite("device_level_populations", &Model::device_level_populations)
        .def_readonly("nranks", &Model::nranks)
        .def_readonly("rank", &Model::rank)
        .def_readonly("o_begin", &Model::o_begin)
        .def_readonly("o_end", &Model::o_end)
        .def_static("nccl_unique_id", []() { return py::bytes(Model::nccl_unique_id()); })
        .def("set_communicator", [](Model& s, int nranks, int rank, py::bytes id) { s.set_communicator(nranks, rank, std::string(id)); })
        .def_property_readonly("last_image_stats", [](const Model& s) {
            py::dict d;
            d["n_pixels"] = s.last_image_stats.n_pixels; d["n_entries"] = s.last_image_stats.n_entries; d["n_solves"] = s.last_image_stats.n_solves;
            d["ms_trace"] = s.last_image_stats.ms_trace; d["ms_solve"] = s.last_image_stats.ms_solve; d["ms_total"] = s.last_image_stats.ms_total;
            d["n_launches"] = s.last_image_stats.n_launches; d["max_ray_length"] = s.last_image_stats.max_ray_length;
            return d;
        })
        .def_property_readonly("last_stats", [](const Model& s) {
            py::dict d;
            d["n_steps"] = s.last_stats.n_steps; d["n_entries"] = s.last_stats.n_entries; d["n_solves"] = s.last_stats.n_solves;
            d["n_launches"] = s.last_stats.n_launches; d["ms_solve"] = s.last_stats.ms_solve; d["ms_total"] = s.last_stats.ms_total;
            d["ms_trace_count"] = s.last_stats.ms_trace_count; d["ms_trace_fill"] = s.last_stats.ms_trace_fill; d["ms_reduce"] = s.last_stats.ms_reduce;
            d["n_solve_launches"] = s.last_stats.n_solve_launches; d["paths_cached"] = s.last_stats.paths_cached;
            return d;
        })
        .def_readonly("parameters", &Model::parameters, py::return_value_policy::reference_internal)
        .def_readonly("images", &Model::images)
        .def_readonly("error_mean", &Model::error_mean)
        .def_readonly("error_max", &Model::error_max)
        .def_readonly("seconds_radiation_field", &Model::seconds_radiation_field)
        .def_readonly("seconds_statistical_equilibrium", &Model::seconds_statistical_equilibrium)
        .def("mutable_parameters", [](Model& s) -> Parameters& { return s.parameters; }, py::return_value_policy::reference_internal)
        .def("set_geometry", [](Model& s, arr<double> pos, arr<double> vel, arr<uint32_t> nn, arr<uint32_t> nbs, arr<double> dir, arr<double> w,
                                arr<uint32_t> anti, arr<uint32_t> b2p, arr<int32_t> bc, arr<double> bt) {
            s.set_geometry((Size)nn.size(), pos.data(), vel.data(), nn.data(), nbs.data(), (size_t)nbs.size(), (Size)w.size(), dir.data(), w.data(),
                           anti.data(), (Size)b2p.size(), b2p.data(), bc.data(), bt.data());
        })
        .def("set_adaptive_directions", [](Model& s, arr<double> dir, arr<double> w) {
            if ((size_t)w.size() != (size_t)s.npoints * s.nrays || (size_t)dir.size() != 3 * (size_t)s.npoints * s.nrays)
                throw std::runtime_error("set_adaptive_directions: direction must be [npoints * nrays][3], weight [npoints * nrays]");
            s.set_adaptive_directions(dir.data(), w.data());
        })
        .def("set_thermodynamics", [](Model& s, arr<double> T, arr<double> v) { s.set_thermodynamics(T.data(), v.data()); })
        .def("set_abundance", [](Model& s, arr<double> ab) { s.set_abundance(ab.data(), (Size)ab.shape(1)); })
        .def("add_species", [](Model& s, py::dict d) {
            LineProducingSpecies ls;
            Linedata& ld = ls.linedata;
            ld.num = d["num"].cast<Size>(); ld.nlev = d["nlev"].cast<Size>(); ld.nrad = d["nrad"].cast<Size>();
            ld.inverse_mass = d["inverse_mass"].cast<double>();
            ld.irad = vec(d["irad"].cast<arr<uint32_t>>()); ld.jrad = vec(d["jrad"].cast<arr<uint32_t>>());
            ld.energy = vec(d["energy"].cast<arr<double>>()); ld.weight = vec(d["weight"].cast<arr<double>>());
            ld.frequency = vec(d["frequency"].cast<arr<double>>()); ld.A = vec(d["A"].cast<arr<double>>());
            ld.Ba = vec(d["Ba"].cast<arr<double>>()); ld.Bs = vec(d["Bs"].cast<arr<double>>());
            for (auto h : d["colpar"].cast<py::list>()) {
                py::dict c = h.cast<py::dict>();
                CollisionPartner cp;
                cp.num_col_partner = c["num_col_partner"].cast<Size>(); cp.orth_or_para_H2 = c["orth_or_para_H2"].cast<std::string>();
                cp.icol = vec(c["icol"].cast<arr<uint32_t>>()); cp.jcol = vec(c["jcol"].cast<arr<uint32_t>>());
                cp.tmp = vec(c["tmp"].cast<arr<double>>());
                auto Ce = c["Ce"].cast<arr<double>>(), Cd = c["Cd"].cast<arr<double>>();
                for (py::ssize_t t = 0; t < Ce.shape(0); t++) {
                    cp.Ce.emplace_back(Ce.data() + t * Ce.shape(1), Ce.data() + (t + 1) * Ce.shape(1));
                    cp.Cd.emplace_back(Cd.data() + t * Cd.shape(1), Cd.data() + (t + 1) * Cd.shape(1));
                }
                ld.colpar.push_back(std::move(cp));
            }
            ls.quadrature_roots = vec(d["quadrature_roots"].cast<arr<double>>());
            ls.quadrature_weights = vec(d["quadrature_weights"].cast<arr<double>>());
            s.add_species(std::move(ls));
        })
        .def("compute_inverse_line_widths", &Model::compute_inverse_line_widths)
        .def("compute_spectral_discretisation", py::overload_cast<>(&Model::compute_spectral_discretisation))
        .def("compute_spectral_discretisation", py::overload_cast<Real, Real>(&Model::compute_spectral_discretisation))
        .def("compute_spectral_discretisation", py::overload_cast<Real, Real, Size>(&Model::compute_spectral_discretisation))
        .def("compute_LTE_level_populations", &Model::compute_LTE_level_populations)
        .def("compute_radiation_field_feautrier_order_2", &Model::compute_radiation_field_feautrier_order_2)
        .def("compute_Jeff", &Model::compute_Jeff)
        .def("compute_level_populations_from_stateq", &Model::compute_level_populations_from_stateq)
        .def("compute_level_populations", &Model::compute_level_populations)
        .def("compute_radiation_field_feautrier_order_2_sparse", &Model::compute_radiation_field_feautrier_order_2_sparse)
        .def("compute_radiation_field_shortchar_order_0", &Model::compute_radiation_field_shortchar_order_0)
        .def("compute_level_populations_shortchar", &Model::compute_level_populations_shortchar)
        .def("compute_Jeff_sparse", &Model::compute_Jeff_sparse)
        .def("compute_radiation_field_feautrier_order_2_anis", &Model::compute_radiation_field_feautrier_order_2_anis)
        .def("get_J_anis", [](const Model& s, Size l) { /* (J, J2_0, J2_1_Re, J2_1_Im, J2_2_Re, J2_2_Im) of species l, each [npoints][nrad] */
            const auto& ls = s.lineProducingSpecies.at(l);
            return py::make_tuple(out2(ls.J, s.npoints, ls.linedata.nrad), out2(ls.J2_0, s.npoints, ls.linedata.nrad), out2(ls.J2_1_Re, s.npoints, ls.linedata.nrad),
                                  out2(ls.J2_1_Im, s.npoints, ls.linedata.nrad), out2(ls.J2_2_Re, s.npoints, ls.linedata.nrad), out2(ls.J2_2_Im, s.npoints, ls.linedata.nrad));
        })
        .def("compute_level_populations_from_stateq_sparse", &Model::compute_level_populations_from_stateq_sparse)
        .def("compute_level_populations_sparse", &Model::compute_level_populations_sparse)
        .def("compute_image_new", &Model::compute_image_new)
        .def("compute_image_optical_depth_new", &Model::compute_image_optical_depth_new)
        .def("set_column", &Model::set_column)
        .def_property("density", [](const Model& s) { return py::array_t<double>(s.density.size(), s.density.data()); },
                      [](Model& s, arr<double> d) { s.density = vec(d); })
        .def_property_readonly("column", [](const Model& s) { return s.column.empty() ? out2(s.column, 0, s.npoints) : out2(s.column, s.nrays, s.npoints); })
        // test hook: radiation.J and the diagonal Lambda lists as a radiation-field call leaves them (host-only tests of compute_Jeff / stateq)
        .def("set_radiation_field", [](Model& s, arr<double> J, arr<double> lam, arr<uint32_t> cnt, py::object J_lines) {
            s.J = vec(J);
            const Size L = s.nlines();
            if (!J_lines.is_none()) { /* lspec.J of the memory-sparse solver, [npoints][nlines] */
                auto jl = J_lines.cast<arr<double>>();
                for (Size l = 0; l < s.nlspecs(); l++) {
                    auto& ls = s.lineProducingSpecies[l];
                    const Size nrad = ls.linedata.nrad;
                    for (Size p = 0; p < s.npoints; p++)
                        for (Size k = 0; k < nrad; k++) ls.J[(size_t)p * nrad + k] = jl.data()[(size_t)p * L + s.nrad_cum[l] + k];
                }
            }
            for (Size l = 0; l < s.nlspecs(); l++) {
                auto& ls = s.lineProducingSpecies[l];
                const Size nrad = ls.linedata.nrad;
                for (Size p = 0; p < s.npoints; p++)
                    for (Size k = 0; k < nrad; k++) {
                        ls.lambda_Ls[(size_t)p * nrad + k]  = lam.data()[(size_t)p * L + s.nrad_cum[l] + k];
                        ls.lambda_cnt[(size_t)p * nrad + k] = cnt.data()[(size_t)p * L + s.nrad_cum[l] + k];
                    }
            }
        }, py::arg("J"), py::arg("lam"), py::arg("cnt"), py::arg("J_lines") = py::none())
        // test hook, n_off_diag > 0: radiation.J and the banded Lambda lists [npoints][nlines][cap] (insertion order) of a radiation-field call
        .def("set_radiation_field_lists", [](Model& s, arr<double> J, arr<double> val, arr<uint32_t> nr, arr<uint32_t> cnt) {
            s.J = vec(J);
            const size_t rows = (size_t)s.npoints * s.nlines();
            if ((size_t)cnt.size() != rows || val.size() != nr.size() || rows == 0 || (size_t)val.size() % rows) throw std::runtime_error("set_radiation_field_lists: shapes");
            s.set_lambda_lists(val.data(), nr.data(), cnt.data(), (size_t)val.size() / rows);
        }, py::arg("J"), py::arg("val"), py::arg("nr"), py::arg("cnt"))
        .def("stateq_solver_info", [](const Model& s, size_t l) { /* (Krylov iterations, final residual) of the last banded statistical-equilibrium solve */
            return py::make_tuple(s.lineProducingSpecies.at(l).stateq_iterations, (double)s.lineProducingSpecies.at(l).stateq_residual);
        })
        // results (copies, reference layouts)
        .def("get_J", [](const Model& s) { return out2(s.J, s.npoints, s.nfreqs); })
        .def("get_nu", [](const Model& s) { return out2(s.nu, s.npoints, s.nfreqs); })
        .def("get_nu_rows", [](const Model& s, size_t o0, size_t o1) { /* rows [o0, o1) of radiation.frequencies.nu (a 1M-point model's nu is 7 GB) */
            if (o1 > s.npoints || o0 > o1) throw std::runtime_error("get_nu_rows: bad range");
            py::array_t<double> a({o1 - o0, (size_t)s.nfreqs});
            std::copy(s.nu.begin() + o0 * s.nfreqs, s.nu.begin() + o1 * s.nfreqs, a.mutable_data());
            return a;
        })
        .def("get_J_rows", [](const Model& s, size_t o0, size_t o1) {
            if (o1 > s.npoints || o0 > o1 || s.J.size() < (size_t)s.npoints * s.nfreqs) throw std::runtime_error("get_J_rows: bad range or no J");
            py::array_t<double> a({o1 - o0, (size_t)s.nfreqs});
            std::copy(s.J.begin() + o0 * s.nfreqs, s.J.begin() + o1 * s.nfreqs, a.mutable_data());
            return a;
        })
        .def("get_emissivity", [](const Model& s) { return out2(s.emissivity, s.npoints, s.nlines()); })
        .def("get_opacity", [](const Model& s) { return out2(s.opacity, s.npoints, s.nlines()); })
        .def("get_inverse_width", [](const Model& s) { return out2(s.inverse_width, s.npoints, s.nlines()); })
        .def("get_population", [](const Model& s, Size l) { /* lineProducingSpecies[l].population, long double -> double */
            const auto& ls = s.lineProducingSpecies.at(l);
            py::array_t<double> a({(size_t)s.npoints, (size_t)ls.linedata.nlev});
            for (size_t i = 0; i < ls.population.size(); i++) a.mutable_data()[i] = (double)ls.population[i];
            return a;
        })
        .def("get_population_hilo", [](const Model& s, Size l) { /* the long double populations as hi + lo doubles */
            const auto& ls = s.lineProducingSpecies.at(l);
            py::array_t<double> hi({(size_t)s.npoints, (size_t)ls.linedata.nlev}), lo({(size_t)s.npoints, (size_t)ls.linedata.nlev});
            for (size_t i = 0; i < ls.population.size(); i++) {
                hi.mutable_data()[i] = (double)ls.population[i];
                lo.mutable_data()[i] = (double)(ls.population[i] - (LD)hi.data()[i]);
            }
            return py::make_tuple(hi, lo);
        })
        .def("set_population", [](Model& s, Size l, arr<double> hi, arr<double> lo, arr<double> tot) { /* checkpoint / resume (lineProducingSpecies.cpp:49-71) */
            auto& ls = s.lineProducingSpecies.at(l);
            if ((size_t)hi.size() != ls.population.size() || (size_t)lo.size() != ls.population.size() || (size_t)tot.size() != ls.population_tot.size())
                throw std::runtime_error("set_population: wrong size");
            for (size_t i = 0; i < ls.population.size(); i++) ls.population[i] = (LD)hi.data()[i] + (LD)lo.data()[i];
            std::copy(tot.data(), tot.data() + tot.size(), ls.population_tot.begin());
            ls.populations.clear(); ls.residuals.clear();
            ls.populations.push_back(ls.population);
        })
        .def("refresh_line_state_from_populations", [](Model& s) { s.set_emissivity_and_opacity(); })
        .def("get_lambda_cnt", [](const Model& s, Size l) {
            const auto& ls = s.lineProducingSpecies.at(l);
            py::array_t<uint32_t> a({(size_t)s.npoints, (size_t)ls.linedata.nrad});
            std::copy(ls.lambda_cnt.begin(), ls.lambda_cnt.end(), a.mutable_data());
            return a;
        })
        .def("get_population_tot", [](const Model& s, Size l) { const auto& v = s.lineProducingSpecies.at(l).population_tot; return py::array_t<double>(v.size(), v.data()); })
        .def("get_Jeff", [](const Model& s, Size l) { const auto& ls = s.lineProducingSpecies.at(l); return out2(ls.Jeff, s.npoints, ls.linedata.nrad); })
        .def("get_J_species", [](const Model& s, Size l) { const auto& ls = s.lineProducingSpecies.at(l); return out2(ls.J, s.npoints, ls.linedata.nrad); })
        .def("get_Jlin", [](const Model& s, Size l) { const auto& ls = s.lineProducingSpecies.at(l); return out2(ls.Jlin, s.npoints, ls.linedata.nrad); })
        .def("get_lambda", [](const Model& s, Size l) { const auto& ls = s.lineProducingSpecies.at(l); return out2(ls.lambda_Ls, s.npoints, ls.linedata.nrad); })
        .def("convergence", [](const Model& s, Size l) { const auto& ls = s.lineProducingSpecies.at(l);
                                                         return py::make_tuple(ls.fraction_not_converged, ls.relative_change_mean, ls.relative_change_max); });
}
