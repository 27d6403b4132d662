// cg.cuh -- preconditioned CG engine shared by nsb_cg, the heat stepper and the SIMPLE time stepper (cg.cu, simple.cu).
#pragma once

#include "common.cuh"

namespace nsb {

struct CgState {
    double wd, wdn, alpha, beta, err0, tol_abs;
    double red[2];
    int it, maxsteps, stop, status, record;
};

// work vectors of one CG solve: rhs goes into q, the solution comes out of u (u and s have room for the operator's halo tail)
struct CgBuffers {
    DevBuf<double> u, d, w, s, q, hist, alphas, betas;
    DevBuf<CgState> st;
};

// a captured batch of iterations, reusable as long as the buffers, the operator and the preconditioner stay the same
struct CgGraph {
    cudaGraphExec_t exec = nullptr;
    int64_t launches = 0;
    int chunk = 0;
    CgGraph() = default;
    CgGraph(const CgGraph&) = delete;
    CgGraph& operator=(const CgGraph&) = delete;
    void reset() {
        if (exec) cudaGraphExecDestroy(exec);
        exec = nullptr;
        launches = 0;
        chunk = 0;
    }
    ~CgGraph() { reset(); }
};

int cg_alloc(Context& c, CgBuffers& b, nsb_mat_s* A, int maxsteps, bool lanczos);
// CG on the device-resident right-hand side in b.q -> b.u (NGSolve CGSolver's stopping rule sqrt<Cr,r> <= tol * initial)
int cg_solve_device(Context& c, nsb_mat_s* A, nsb_pre_s* pre, CgBuffers& b, bool initialize, int maxsteps, double tol, CgState* hs,
                    float* ms, CgGraph* cache = nullptr);

}  // namespace nsb
