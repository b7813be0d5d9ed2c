// Stub of thermopack's C++ wrapper (TEST INFRASTRUCTURE ONLY, used to build oracle/_ref).
// The in-tree cppThermopack wrapper is older than the reference expects and the Fortran library cannot
// be built here (no Fortran compiler), so the reference's EoS plumbing is compiled against this stub.
// The oracle driver injects EoS-derived inputs (dmu/dn, Cp_ideal, ...) through GenericEoS from tables
// instead (cpp/eos_interface.h:57-119), so none of these methods is ever reached.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

struct KgStubProperty {
    double value() const { throw std::runtime_error("thermopack stub"); }
    double dt() const { throw std::runtime_error("thermopack stub"); }
    double dv() const { throw std::runtime_error("thermopack stub"); }
    double dp() const { throw std::runtime_error("thermopack stub"); }
    std::vector<double> dn() const { throw std::runtime_error("thermopack stub"); }
    operator double() const { throw std::runtime_error("thermopack stub"); }
};
struct KgStubVectorProperty {
    std::vector<std::vector<double>> dn() const { throw std::runtime_error("thermopack stub"); }
    std::vector<double> dt() const { throw std::runtime_error("thermopack stub"); }
};

class Thermo {
public:
    int VAPPH{2};
    Thermo() = default;
    Thermo(const std::string&) {}
    virtual ~Thermo() = default;
    KgStubVectorProperty chemical_potential_tv(double, double, const std::vector<double>&, bool = false, bool = false, bool = false) const { throw std::runtime_error("thermopack stub: chemical_potential_tv"); }
    KgStubProperty idealenthalpysingle(double, int, bool = false) const { throw std::runtime_error("thermopack stub: idealenthalpysingle"); }
    KgStubProperty enthalpy_tvp(double, double, const std::vector<double>&, bool = false, bool = false, bool = false) const { throw std::runtime_error("thermopack stub: enthalpy_tvp"); }
    double pressure_tv(double, double, const std::vector<double>&) const { throw std::runtime_error("thermopack stub: pressure_tv"); }
    KgStubProperty specific_volume(double, double, const std::vector<double>&, int, bool = false, bool = false, bool = false) const { throw std::runtime_error("thermopack stub: specific_volume"); }
};
