// compat/GPU/gpu.hh — stands where src/GPU/gpu.hh:1-116 stands (GPU::Matrix, GPU::ICP, gpu_*_params and the five
// operator wrappers are all declared by ../icp.hh).
#pragma once
#include "../cpu.hh"
