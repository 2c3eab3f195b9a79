// cuthho_square — the reference's fictitious-domain cutHHO driver (apps/cuthho/cuthho_square.cpp:1940-2068,
// run_cuthho_fictdom :806-1080) on the B200 path.  Same getopt interface: -k degree -M Nx -N Ny -r refine steps -f.
// -i runs the two-material interface problem (run_cuthho_interface, :1625-1846) the same way.
#include <unistd.h>

#include <cmath>
#include <cstdio>
#include <iostream>

#include "proton_host.hpp"
namespace pb = proton_b200;

int main(int argc, char** argv) {
    size_t degree = 0, int_refsteps = 4; bool solve_interface = false, solve_fictdom = false;
    pb::mesh_init_params<double> mip; mip.Nx = 5; mip.Ny = 5;
    int ch;
    while ((ch = getopt(argc, argv, "k:M:N:r:ifDAd")) != -1) {
        switch (ch) {
        case 'k': degree = atoi(optarg); break;
        case 'M': mip.Nx = atoi(optarg); break;
        case 'N': mip.Ny = atoi(optarg); break;
        case 'r': int_refsteps = atoi(optarg); break;
        case 'i': solve_interface = true; break;
        case 'f': solve_fictdom = true; break;
        case 'D': case 'A': case 'd': break;
        default: std::cout << "wrong arguments" << std::endl; return 1;
        }
    }
    if (solve_interface) {
        try {       // run_cuthho_interface (cuthho_square.cpp:1625-1846): as in main(), the interface run comes first
            pb::cartesian_mesh<double> msh(mip, pb::cartesian_mesh<double>::cuthho_poly);
            pb::circle_level_set<double> level_set_function(0.35, 0.5, 0.5);
            msh.cut_preprocess(level_set_function, (int)int_refsteps);
            auto assembler = pb::make_assembler(msh, pb::hho_degree_info(degree + 1, degree));
            assembler.assemble_all(PB_STAB_CUT, PB_PROBLEM_INTERFACE, PB_FUN_SINSIN, 0, 5.0);
            std::cout << "System unknowns (condensed): " << assembler.system_size() << std::endl;
            pb::cg_params<double> cgp; cgp.apply_preconditioner = true; cgp.convergence_threshold = 1e-13; cgp.divergence_threshold = 1e8;
            cgp.max_iter = 20 * assembler.system_size(); cgp.histfile = "cuthho_cg_hist.dat";       // :1736-1743
            std::vector<double> sol; size_t iters = 0;
            auto er = pb::conjugated_gradient(assembler, sol, cgp, &iters);
            std::cout << "CG iterations: " << iters << (er == pb::cg_exit_reason::CONVERGED ? "" : " (NOT CONVERGED)") << std::endl;
            auto e = assembler.error_norms();
            printf("Energy-norm absolute error:           %.10e\n", std::sqrt(e[0]));                  // :1837
            assembler.write_solution("interface_uT.dat");                                                // uT_gp, :1795-1796, 1839-1842
        } catch (const std::exception& ex) { std::cerr << ex.what() << std::endl; return 2; }
    }
    if (!solve_fictdom) return 0;
    try {
        pb::cartesian_mesh<double> msh(mip, pb::cartesian_mesh<double>::cuthho_poly);
        pb::circle_level_set<double> level_set_function(0.35, 0.5, 0.5);
        msh.cut_preprocess(level_set_function, (int)int_refsteps);
        auto assembler = pb::make_assembler(msh, pb::hho_degree_info(degree + 1, degree));
        assembler.assemble_all(PB_STAB_CUT, PB_PROBLEM_FICTDOM, PB_FUN_SINSIN, 0, 5.0);
        std::cout << "System unknowns (condensed): " << assembler.system_size() << std::endl;
        pb::cg_params<double> cgp; cgp.apply_preconditioner = true; cgp.convergence_threshold = 1e-13; cgp.max_iter = 20 * assembler.system_size();
        cgp.histfile = "cuthho_cg_hist.dat";      // as the reference's CG variant of this solve (cuthho_square.cpp:922-928)
        std::vector<double> sol; size_t iters = 0;
        auto er = pb::conjugated_gradient(assembler, sol, cgp, &iters);
        std::cout << "CG iterations: " << iters << (er == pb::cg_exit_reason::CONVERGED ? "" : " (NOT CONVERGED)") << std::endl;
        auto e = assembler.error_norms();
        printf("h = %.6e\nH1-error : %.10e\n", std::sqrt(2.0) / mip.Nx, std::sqrt(e[0]));
    } catch (const std::exception& ex) { std::cerr << ex.what() << std::endl; return 2; }
    return 0;
}
