// compat/cpu.hh — lets the reference's OWN translation units (src/main.cc, src/GPU/main.cc, src/bench.cc) compile
// unmodified against the B200 path: put this directory on the include path in place of the reference's src/ and
// link libicp_b200.so instead of cpu.cc / GPU/gpu.cc / GPU/compute.cu.  Stands where src/cpu.hh:1-120 stands.
//
// The reference spells its matrices Eigen::MatrixXd.  Eigen is not vendored by the reference (its build fetches it)
// and is absent here, so the names it uses are mapped onto icpb::Matrix (same column-major memory image) and the two
// Eigen expressions the benchmark harness evaluates on its own are supplied by a stand-in (ICP_B200_EIGEN_STANDIN
// in ../icp.hh).  None of this is on the path: every CPU:: / GPU:: entry point below is one call into the library.
#pragma once
#define ICP_B200_EIGEN_STANDIN 1
#include <cmath>
#include <tuple>

#include "../icp.hh"

namespace Eigen {
using MatrixXd = icpb::Matrix;
}
using Eigen::MatrixXd;
