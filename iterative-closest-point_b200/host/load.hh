// load.hh — point-cloud text I/O with the reference's formats (src/load.hh:9-12).
#pragma once
#include "icp.hh"

GPU::Matrix load_matrix(const char* filename);         // src/load.cc:3-33
void write_matrix(const GPU::Matrix& matrix);          // src/load.cc:68-97 -> "output.txt"
void write_matrix_to(const GPU::Matrix& matrix, const char* filename);
