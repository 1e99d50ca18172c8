// loadcheck.cc — test helper: load_matrix(in) -> write_matrix_to(out).  Exercises host/load.cc (the drop-in for
// src/load.cc:3-33,83-97) without touching a GPU, so the CPU test-suite can compare it with the oracle's loader.
#include "load.hh"

int main(int argc, char* argv[])
{
    if (argc < 3) return 64;
    write_matrix_to(load_matrix(argv[1]), argv[2]);
    return 0;
}
