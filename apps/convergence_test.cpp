// convergence_test — the reference's uncut HHO Poisson sweep (apps/convergence_test/convergence_test.cpp) on the
// B200 path.  Same parameters (deg_min, deg_max, min_N, steps, precond, direct, stab_hho; convergence_test.cpp:52-70,
// 336-361), read from a flat `name = value` file (the reference reads them through Lua/sol2, which is not available
// here: only number / true / false values and `--` comments are understood).  `direct = true` has no counterpart on
// the device path (no sparse LU): the solve is always the reference's conjugated_gradient at 1e-12.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>

#include "proton_host.hpp"
namespace pb = proton_b200;

struct convergence_test_params {
    size_t deg_min = 0, deg_max = 3, min_N = 4, steps = 5; bool preconditioner = true, direct = false, stab_hho = true;
    bool write_files = true;        // hho_history_*, cg_history_*, solution_N_*_k_*.dat as the reference writes them (`write_files = false` skips them)
    bool block_precond = false;     // extension: block-Jacobi instead of the diagonal preconditioner
};

static bool read_config(const char* fn, convergence_test_params& ctp) {
    std::ifstream ifs(fn);
    if (!ifs) return false;
    std::string line;
    while (std::getline(ifs, line)) {
        auto c = line.find("--"); if (c != std::string::npos) line = line.substr(0, c);
        auto e = line.find('='); if (e == std::string::npos) continue;
        std::string k = line.substr(0, e), v = line.substr(e + 1);
        auto trim = [](std::string& s) { size_t a = s.find_first_not_of(" \t;"), b = s.find_last_not_of(" \t;\r"); s = a == std::string::npos ? "" : s.substr(a, b - a + 1); };
        trim(k); trim(v);
        bool bv = v == "true"; double nv = (v == "true" || v == "false") ? 0 : atof(v.c_str());
        if (k == "deg_min") ctp.deg_min = (size_t)nv; else if (k == "deg_max") ctp.deg_max = (size_t)nv; else if (k == "min_N") ctp.min_N = (size_t)nv;
        else if (k == "steps") ctp.steps = (size_t)nv; else if (k == "precond") ctp.preconditioner = bv; else if (k == "direct") ctp.direct = bv;
        else if (k == "stab_hho") ctp.stab_hho = bv; else if (k == "write_files") ctp.write_files = bv; else if (k == "block_precond") ctp.block_precond = bv;
    }
    return true;
}

int main(int argc, char** argv) {
    convergence_test_params ctp;
    if (argc >= 2 && !read_config(argv[1], ctp)) { std::cout << "Problems opening configuration file" << std::endl; return 1; }
    if (ctp.direct)     // convergence_test.cpp:222-228 solves with Eigen::SparseLU when `direct = true`; there is no sparse LU on the device path
        std::cout << "Note: `direct = true` is ignored: the solve is conjugated_gradient at 1e-12 (no sparse direct solver on the GPU path)" << std::endl;
    try {
        for (size_t k = ctp.deg_min; k <= ctp.deg_max; k++) {
            std::cout << "Testing degree " << k << ". Expected rates: L2 " << k + 2 << ", energy " << k + 1 << std::endl;
            double prev_l2 = 0, prev_en = 0, prev_h = 0;
            std::ofstream hho_conv_ofs;           // convergence_test.cpp:155-161
            if (ctp.write_files) { std::stringstream ss; ss << (ctp.preconditioner ? "hho_history_precond_" : "hho_history_") << k << ".txt"; hho_conv_ofs.open(ss.str()); }
            for (size_t i = 0, N = ctp.min_N; i < ctp.steps; i++, N *= 2) {
                pb::mesh_init_params<double> mip; mip.Nx = N; mip.Ny = N;
                pb::cartesian_mesh<double> msh(mip, pb::cartesian_mesh<double>::quad);
                auto assembler = pb::make_assembler(msh, pb::hho_degree_info(k + 1, k));
                assembler.assemble_all(ctp.stab_hho ? PB_STAB_FANCY : PB_STAB_NAIVE, PB_PROBLEM_HHO, PB_FUN_SINSIN);
                pb::cg_params<double> cgp; cgp.apply_preconditioner = ctp.preconditioner; cgp.convergence_threshold = 1e-12;
                cgp.max_iter = 3 * assembler.system_size();
                // the reference keeps its default guard (stop when the residual grows 100x); on the condensed face system the
                // 2-norm residual of Jacobi-PCG overshoots in the first iterations (x160 at k=3, N=512), so the guard is widened
                cgp.divergence_threshold = 1e8;
                cgp.block_preconditioner = ctp.block_precond;
                if (ctp.write_files) {            // convergence_test.cpp:232-242
                    std::stringstream ss; ss << (ctp.preconditioner ? "cg_history_precond_" : "cg_history_") << N << "_" << k << ".txt"; cgp.histfile = ss.str();
                }
                std::vector<double> sol; size_t iters = 0;
                if (pb::conjugated_gradient(assembler, sol, cgp, &iters) != pb::cg_exit_reason::CONVERGED) std::cout << "Warning! Solver didn't converge..." << std::endl;
                auto e = assembler.error_norms();
                if (ctp.write_files) {            // convergence_test.cpp:250-252, 280-286, 311-312
                    std::stringstream ss; ss << "solution_N_" << N << "_k_" << k << ".dat"; assembler.write_solution(ss.str());
                    hho_conv_ofs << std::sqrt(2.0) / N << " " << e[0] << " " << e[1] << std::endl;
                }
                const double h = std::sqrt(2.0) / N, l2 = std::sqrt(e[0]), en = std::sqrt(e[2]);     // mesh_h = diameter of cell 0
                printf("  N = %5zu  h = %.4e  L2 %.10e", N, h, l2);
                if (i > 0) printf(" (%.2f)", std::log(prev_l2 / l2) / std::log(prev_h / h)); else printf("       ");
                printf("  energy %.10e", en);
                if (i > 0) printf(" (%.2f)", std::log(prev_en / en) / std::log(prev_h / h));
                printf("  CG its %zu\n", iters);
                prev_l2 = l2; prev_en = en; prev_h = h;
            }
        }
    } catch (const std::exception& ex) { std::cerr << ex.what() << std::endl; return 2; }
    return 0;
}
