// main.cc — the `icp [ref] [scene] [nb_max_iter]` driver (reference: src/main.cc:5-25, src/GPU/main.cc:3-21)
// on top of the B200 path.  Same arguments, same stdout/stderr text, same output.txt, same exit codes.
// Extra, optional: environment ICP_ALLOW_UNEQUAL=1 skips the equal-size check (needed for the bunny
// pair bun000/bun045, which the reference driver refuses: src/cpu.cc:44-47).
#include <cstdlib>
#include <iostream>

#include "load.hh"

int main(int argc, char* argv[])
{
    if (argc < 4) {
        std::cout << "Usage: ./icp [path_to_ref_cloud] [path_to_transform_cloud] [nb_iter]\n";
        return -1;
    }
    const int max_iter = std::atoi(argv[3]);
    auto matrix_ref = load_matrix(argv[1]);
    auto matrix_scene = load_matrix(argv[2]);

    const char* allow = std::getenv("ICP_ALLOW_UNEQUAL");
    CPU::ICP icp(matrix_ref, matrix_scene, max_iter);
    icp.find_corresponding(!(allow && allow[0] == '1'));
    write_matrix(icp.new_p);
    return 0;
}
