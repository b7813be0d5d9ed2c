#pragma once
#include "thermo.h"
class Saftvrqmie : public Thermo { public:
    Saftvrqmie(const std::string& c, int = 1) : Thermo(c) {}
    void set_pure_fluid_param(int, double, double, double, double, double) {}
};
