#pragma once
#include "thermo.h"
class LJs_uv : public Thermo { public: LJs_uv(const std::string& = "Default", double = 1.0) {} };
