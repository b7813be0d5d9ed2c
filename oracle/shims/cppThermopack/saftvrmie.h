#pragma once
#include "thermo.h"
class Saftvrmie : public Thermo { public: Saftvrmie(const std::string& c) : Thermo(c) {} };
