// compat/header.hh — stands where src/GPU/header.hh:1-5 stands.
#pragma once
#include <iostream>

#include "load.hh"
#include "GPU/gpu.hh"
