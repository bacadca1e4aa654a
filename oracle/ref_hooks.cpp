// TEST INFRASTRUCTURE ONLY -- parity-oracle instrumentation, never shipped or timed.
//
// Our own translation unit, compiled against the UNMODIFIED reference headers where they
// lie under /root/reference (see oracle/build_ref.sh). The reference keeps the per-ray
// arrays nr_/nr_interp_/interp_factor_/shift_/dZ_ in thread-private scratch of `Solver`
// (reference: src/solver/solver.hpp:20-44) and never exposes them; the "ray paths
// bit-exact" parity check needs them, so this module calls the reference's public
// Solver::setup / Solver::trace_ray_for_line (src/solver/solver.tpp:1-13, 773-800) for one
// (origin, ray pair, line) and hands the arrays back to Python (as raw little-endian bytes; oracle/refcore.py wraps them). It is linked against the
// same object files as oracle/_ref/magritte/core.so, so it runs the reference's own code.
#include "../configure.hpp"
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include "model/model.hpp"
#include "solver/solver.hpp"

namespace py = pybind11;

struct TraceHook {
    Solver solver;
    Model* model;

    TraceHook(Model& m) : model(&m) { solver.setup<CoMoving, false>(m); }

    Size length() const { return solver.length; }
    Size centre() const { return solver.centre; }

    py::dict trace(const Size o, const Size rr, const Size l) {
        Model& m      = *model;
        const Size ar = m.geometry.rays.get_antipod_index(rr);
        const Size c  = solver.centre;
        solver.nr_()[c]    = o;
        solver.shift_()[c] = 1.0;
        // Same call sequence as reference src/solver/solver.tpp:484-496
        const Size first = solver.trace_ray_for_line<CoMoving, false>(m, l, o, rr, -1, c - 1, c - 1) + 1;
        const Size last  = solver.trace_ray_for_line<CoMoving, false>(m, l, o, ar, +1, c + 1, c) - 1;
        const Size n     = last + 1 - first;
        std::vector<uint32_t> nr(n), nri(n);
        std::vector<double> shift(n), fac(n), dZ(n > 0 ? n - 1 : 0);
        for (Size i = 0; i < n; i++) {
            nr[i]    = solver.nr_()[first + i];
            shift[i] = solver.shift_()[first + i];
            // the centre entry of nr_interp_/interp_factor_ is never written by the reference
            nri[i] = (first + i == c) ? o : solver.nr_interp_()[first + i];
            fac[i] = (first + i == c) ? 0.0 : solver.interp_factor_()[first + i];
            if (i + 1 < n) dZ[i] = solver.dZ_()[first + i];
        }
        py::dict d;
        d["first"]         = first;
        d["last"]          = last;
        d["centre"]        = c;
        d["nr"]            = py::bytes(reinterpret_cast<const char*>(nr.data()), nr.size() * 4);
        d["nr_interp"]     = py::bytes(reinterpret_cast<const char*>(nri.data()), nri.size() * 4);
        d["interp_factor"] = py::bytes(reinterpret_cast<const char*>(fac.data()), fac.size() * 8);
        d["shift"]         = py::bytes(reinterpret_cast<const char*>(shift.data()), shift.size() * 8);
        d["dZ"]            = py::bytes(reinterpret_cast<const char*>(dZ.data()), dZ.size() * 8);
        return d;
    }
};

py::bytes dshift_max(Model& m) {
    m.set_dshift_max();
    return py::bytes(reinterpret_cast<const char*>(m.dshift_max.vec.data()), m.dshift_max.vec.size() * 8);
}

PYBIND11_MODULE(ref_hooks, module) {
    module.doc() = "Parity-oracle hooks into the unmodified reference solver (test only).";
    py::class_<TraceHook>(module, "TraceHook")
        .def(py::init<Model&>(), py::keep_alive<1, 2>())
        .def("trace", &TraceHook::trace)
        .def("length", &TraceHook::length)
        .def("centre", &TraceHook::centre);
    module.def("dshift_max", &dshift_max);
}
