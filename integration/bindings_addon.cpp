// integration/bindings_addon.cpp -- the pybind11 side of the boundary as compiled code: what a maintainer adds next to the class
// definitions of cpp/bindings.cpp:85-202. `cpp_GpuMieKinGas` takes the constructor arguments of `cpp_MieKinGas` (cpp/bindings.cpp:171-179)
// and exposes the reference's own property methods unchanged -- they now run on collision integrals fed from the device -- plus the two
// additive batched entry points of SURVEY 8(b).3, `omega_table` and `transport_batch`.
//
// Built as its own extension module (integration/_build/libkgshim_py*.so, integration/Makefile) from the unmodified reference objects,
// gpu_miekingas.cpp and this file; the EoS is handed over as arrays (set_eos_tables -> GenericEoS, cpp/eos_interface.h:57-119), so the
// module needs neither thermopack nor the -DPYLIB variant of the reference classes.
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>
#include <pybind11/numpy.h>
#include <pybind11/eigen.h>
#include "gpu_miekingas.h"

namespace py = pybind11;

namespace {
struct ArrayEoS {                    // table-backed EoS with the interface GenericEoS asks for (cpp/eos_interface.h:57-119)
    int VAPPH{2};
    int nc{0};
    std::vector<double> dmudn_tab, cp_id, dvdn_tab;
    double p{0.}, vm{0.};
    std::vector<std::vector<double>> dmudn(double, double, const std::vector<double>) const {
        std::vector<std::vector<double>> out(nc, std::vector<double>(nc));
        for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) out[i][j] = dmudn_tab[i * nc + j];
        return out;
    }
    double Cp_ideal(double, int c) const { return cp_id.at(c - 1); }
    double Cp_real(double, double, const std::vector<double>) const { throw std::runtime_error("ArrayEoS: Cp_real"); }
    double pressure_tv(double, double, const std::vector<double>) const { return p; }
    double specific_volume(double, double, const std::vector<double>, int) const { return vm; }
    std::vector<double> dvdn(double, double, const std::vector<double>, int) const { return dvdn_tab; }
};
std::vector<double> flat(const py::array_t<double, py::array::c_style | py::array::forcecast>& a) {
    return std::vector<double>(a.data(), a.data() + a.size());
}
}

PYBIND11_MODULE(libkgshim_py, m) {
    m.doc() = "GpuMieKinGas: the reference's MieKinGas with its collision integrals and transfer lengths computed by libkgcuda.so";
    py::class_<GpuMieKinGas>(m, "cpp_GpuMieKinGas")
        .def(py::init<vector1d, vector2d, vector2d, vector2d, vector2d, bool, bool>())                 // cpp/bindings.cpp:171-179
        // the reference's own methods (cpp/bindings.cpp:85-135), untouched
        .def("viscosity", [](GpuMieKinGas& s, double T, double Vm, vector1d x, int N) { return s.viscosity(T, Vm, x, N); })
        .def("viscosity_tp", [](GpuMieKinGas& s, double T, double p, vector1d x, int N) { return s.viscosity_tp(T, p, x, N); })
        .def("thermal_conductivity", [](GpuMieKinGas& s, double T, double Vm, vector1d x, int N) { return s.thermal_conductivity(T, Vm, x, N); })
        .def("thermal_conductivity_tp", [](GpuMieKinGas& s, double T, double p, vector1d x, int N) { return s.thermal_conductivity_tp(T, p, x, N); })
        .def("interdiffusion", [](GpuMieKinGas& s, double T, double Vm, vector1d x, int N, int frame, int dep, int solvent, bool compress) {
                 return s.interdiffusion(T, Vm, x, N, frame, dep, solvent, compress); },
             py::arg("T"), py::arg("Vm"), py::arg("x"), py::arg("N") = 2, py::arg("frame_of_reference") = (int) FrameOfReference::CoN,
             py::arg("dependent_idx") = -1, py::arg("solvent_idx") = -1, py::arg("do_compress") = true)
        .def("thermal_diffusion_coeff", [](GpuMieKinGas& s, double T, double Vm, vector1d x, int N, int frame, int dep, int solvent) {
                 return s.thermal_diffusion_coeff(T, Vm, x, N, frame, dep, solvent); },
             py::arg("T"), py::arg("Vm"), py::arg("x"), py::arg("N") = 2, py::arg("frame_of_reference") = (int) FrameOfReference::CoN,
             py::arg("dependent_idx") = -1, py::arg("solvent_idx") = -1)
        .def("get_mtl", [](GpuMieKinGas& s, double rho, double T, vector1d x) { return s.get_mtl(rho, T, x); })
        .def("get_etl", [](GpuMieKinGas& s, double rho, double T, vector1d x) { return s.get_etl(rho, T, x); })
        .def("get_rdf", [](GpuMieKinGas& s, double rho, double T, vector1d x) { return s.get_rdf(rho, T, x); })
        .def("omega", [](GpuMieKinGas& s, int i, int j, int l, int r, double T) { return s.omega(i, j, l, r, T); })
        .def("set_omega_correlation_active", &GpuMieKinGas::set_omega_correlation_active)
        .def("set_eos_tables", [](GpuMieKinGas& s, int nc_eos, vector1d dmudn, vector1d cp_id, double p, double vm, vector1d dvdn) {
                 ArrayEoS e; e.nc = nc_eos; e.dmudn_tab = dmudn; e.cp_id = cp_id; e.dvdn_tab = dvdn; e.p = p; e.vm = vm;
                 s.set_eos(GenericEoS(std::move(e))); },
             "EoS-derived inputs of the next property calls as arrays: dmudn [nc][nc], Cp_ideal [nc], pressure, molar volume (what "
             "specific_volume(T, p, x) returns for the *_tp calls), dvdn [nc]")
        // additive batched entry points
        .def("omega_table", [](GpuMieKinGas& s, vector1d T, std::vector<int> l, std::vector<int> r) {
                 const std::vector<double> om = s.omega_table(T, l, r);
                 const py::ssize_t nT = (py::ssize_t) T.size(), nlr = (py::ssize_t) l.size(), np = (py::ssize_t) (om.size() / (T.size() * l.size()));
                 py::array_t<double> out({nT, np, nlr});
                 std::copy(om.begin(), om.end(), out.mutable_data());
                 return out; },
             "Omega^(l,r)_ij [nT][npairs][nlr] by quadrature on the device; the values are also stored in the KineticGasCache")
        .def("transport_batch", [](GpuMieKinGas& s, py::array_t<double, py::array::c_style | py::array::forcecast> T,
                                   py::array_t<double, py::array::c_style | py::array::forcecast> Vm,
                                   py::array_t<double, py::array::c_style | py::array::forcecast> x, int N,
                                   py::array_t<double, py::array::c_style | py::array::forcecast> dmudn,
                                   py::array_t<double, py::array::c_style | py::array::forcecast> cp_ideal) {
                 const py::ssize_t n = T.size();
                 const GpuMieKinGas::BatchResult r = s.transport_batch(flat(T), flat(Vm), flat(x), N, flat(dmudn), flat(cp_ideal));
                 py::dict out;
                 auto put = [&](const char* key, const std::vector<double>& v, std::vector<py::ssize_t> shape) {
                     if (v.empty()) return;
                     py::array_t<double> a(shape);
                     std::copy(v.begin(), v.end(), a.mutable_data());
                     out[key] = a;
                 };
                 put("viscosity", r.viscosity, {n});
                 put("conductivity", r.conductivity, {n});
                 put("D", r.D, {n, (py::ssize_t) r.nD, (py::ssize_t) r.nD});
                 put("DT", r.DT, {n, r.nD + 1});
                 return out; },
             py::arg("T"), py::arg("Vm"), py::arg("x"), py::arg("N") = 2, py::arg("dmudn") = py::array_t<double>(), py::arg("cp_ideal") = py::array_t<double>(),
             "viscosity, thermal conductivity, interdiffusion (CoN, independent) and thermal diffusion coefficients of n state points in one "
             "device pass; dmudn [n][nce][nce] and cp_ideal [n][nce] are the EoS-derived inputs (empty: the properties that need them are skipped)")
        .def("counters", [](GpuMieKinGas& s) {
                 py::dict d;
                 d["cpu_theta_integrand_calls"] = s.cpu_theta_integrand_calls(); d["device_feeds"] = s.device_feeds(); d["omega_values_stored"] = s.omega_values_stored();
                 return d; });
}
