#pragma once
#include "thermo.h"
class Ideal : public Thermo { public: Ideal() = default; Ideal(const std::string& c) : Thermo(c) {} };
