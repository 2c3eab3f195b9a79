// obstacle — the reference's obstacle driver (apps/obstacle/obstacle.cpp:224-296) on the B200 path: -k degree (0 or 1),
// -N cells per side, domain [-1,1]^2.
#include <unistd.h>

#include <cstdio>
#include <iostream>

#include "proton_host.hpp"
namespace pb = proton_b200;

int main(int argc, char** argv) {
    size_t degree = 1;
    pb::mesh_init_params<double> mip; mip.Nx = 5; mip.Ny = 5; mip.min_x = -1; mip.min_y = -1;
    int ch;
    while ((ch = getopt(argc, argv, "k:N:")) != -1) {
        switch (ch) {
        case 'k': degree = atoi(optarg); if (degree != 0 && degree != 1) { std::cout << "Degree can be 0 or 1. Falling back to 1" << std::endl; degree = 1; } break;
        case 'N': mip.Nx = atoi(optarg); mip.Ny = atoi(optarg); break;
        default: std::cout << "wrong arguments" << std::endl; return 1;
        }
    }
    try {
        pb::cartesian_mesh<double> msh(mip, pb::cartesian_mesh<double>::quad);
        size_t iters = 0;
        double err = pb::run_hho_obstacle(msh, degree, &iters);
        printf("Error: %.6g\n", err);
    } catch (const std::exception& ex) { std::cerr << ex.what() << std::endl; return 2; }
    return 0;
}
