// integration/gpu_saft_eos.h -- the device SAFT-VR Mie EoS (libkgcuda.so, kg_eos_*) behind the reference's EoS interface.
//
// cpp/eos_interface.h:57-135: a KineticGas object holds a std::unique_ptr<GenericEoS>; any class with methods of equivalent
// signature (dmudn, Cp_ideal, Cp_real, pressure_tv, specific_volume, dvdn, member VAPPH) can be wrapped:
//     kin.set_eos(GenericEoS(GpuSaftEoS(nc, m, sigma, eps_div_k, la, lr, cp_type, cp_par, mw)));
// This class takes thermopack's place (ThermoWrapper, cpp/eos_interface.h:117-132): same numbers to 1e-12 (tests/test_gpu_eos.py),
// computed on the GPU. The handle is shared: GenericEoS moves its argument.
#pragma once
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>
#include "../include/kgcuda.h"

class GpuSaftEoS {
public:
    int VAPPH{2};
    GpuSaftEoS(int nc, const std::vector<double>& m, const std::vector<double>& sigma, const std::vector<double>& eps_div_k,
               const std::vector<double>& la, const std::vector<double>& lr, const std::vector<int32_t>& cp_type,
               const std::vector<double>& cp_par, const std::vector<double>& mw, int device = 0) : nc(nc) {
        kg_eos* e = nullptr;
        if (kg_eos_saftvrmie_create(device, nc, m.data(), sigma.data(), eps_div_k.data(), la.data(), lr.data(), cp_type.empty() ? nullptr : cp_type.data(),
                                    cp_par.empty() ? nullptr : cp_par.data(), mw.empty() ? nullptr : mw.data(), &e))
            throw std::runtime_error(std::string("GpuSaftEoS: ") + kg_eos_last_error(nullptr));
        h = std::shared_ptr<kg_eos>(e, kg_eos_destroy);
    }
    std::vector<std::vector<double>> dmudn(double t, double V, const std::vector<double> n) const {
        std::vector<double> flat((size_t) nc * nc);
        chk(kg_eos_tv_batch(h.get(), 1, &t, &V, n.data(), flat.data(), nullptr, nullptr, nullptr));
        std::vector<std::vector<double>> out(nc, std::vector<double>(nc));
        for (int i = 0; i < nc; i++) for (int j = 0; j < nc; j++) out[i][j] = flat[(size_t) i * nc + j];
        return out;
    }
    double Cp_ideal(double t, int c) const {            // c is 1-based, as thermopack's idealenthalpysingle (cpp/eos_interface.h:123)
        std::vector<double> cp(nc), x(nc, 1.0 / nc);
        const double V = 1.0;
        chk(kg_eos_tv_batch(h.get(), 1, &t, &V, x.data(), nullptr, cp.data(), nullptr, nullptr));
        return cp[c - 1];
    }
    double Cp_real(double, double, const std::vector<double>) const { throw std::runtime_error("GpuSaftEoS: Cp_real is not implemented"); }
    double pressure_tv(double t, double V, const std::vector<double> n) const {
        double p = 0;
        chk(kg_eos_tv_batch(h.get(), 1, &t, &V, n.data(), nullptr, nullptr, &p, nullptr));
        return p;
    }
    double specific_volume(double t, double p, const std::vector<double> n, int phase) const {
        if (phase != VAPPH) throw std::runtime_error("GpuSaftEoS: vapour-like root only");
        double V = 0; int32_t st = 0;
        chk(kg_eos_specific_volume_batch(h.get(), 1, &t, &p, n.data(), &V, &st));
        if (st) throw std::runtime_error("GpuSaftEoS: density solver did not converge");
        return V;
    }
    std::vector<double> dvdn(double t, double p, const std::vector<double> n, int phase) const {
        const double V = specific_volume(t, p, n, phase);
        std::vector<double> out(nc);
        chk(kg_eos_tv_batch(h.get(), 1, &t, &V, n.data(), nullptr, nullptr, nullptr, out.data()));
        return out;
    }
private:
    int nc;
    std::shared_ptr<kg_eos> h;
    void chk(int rc) const { if (rc) throw std::runtime_error(std::string("GpuSaftEoS: ") + kg_eos_last_error(h.get())); }
};
