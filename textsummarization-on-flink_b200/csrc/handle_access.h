// Library-internal access to a pgen_t from translation units other than capi.cu.
#pragma once
#include <string>

#include "../../include/pgen.h"
#include "engine.h"

pgn::Engine* pgen_internal_engine(pgen_t* h);
bool pgen_internal_coverage(pgen_t* h);
int pgen_internal_fail(pgen_t* h, int code, const std::string& msg);
