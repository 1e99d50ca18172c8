// compat/load.hh — stands where src/load.hh:1-12 stands.
#pragma once
#include "cpu.hh"
#include "../load.hh"

inline MatrixXd cpu_load_matrix(const char* filename) { return load_matrix(filename); }   // src/load.cc:35-66
