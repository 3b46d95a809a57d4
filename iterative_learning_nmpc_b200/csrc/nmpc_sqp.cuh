// nmpc_sqp.cuh -- SQP / SQP_RTI driver kernels: NLP residuals, termination, shift warm start.
//
// Device-side counterpart of acados' ocp_nlp_sqp / ocp_nlp_sqp_rti outer loop
// and of QuadrupedAcadosSolver.warm_start_solver (mpc_controller/utils/solver.py:290-342).
// No host round trip happens inside a solve: every problem carries a `done` flag that the
// linearise / QP kernels test on entry.
#pragma once
#include "nmpc_common.cuh"

#define RES_THREADS 128

__device__ __forceinline__ void atomic_max_nonneg(unsigned long long* addr, double v)
{
    // NaN is reported as +inf so that it survives the max
    if (!(v == v)) v = __longlong_as_double(0x7ff0000000000000LL);
    atomicMax(addr, (unsigned long long)__double_as_longlong(v));
}

__global__ void nmpc_fill_kernel(double* p, int n, double v)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// start of a solve: clear flags
__global__ void nmpc_sqp_begin_kernel(int B, int rti, int32_t* status, int32_t* done, int32_t* sqp_iter,
                                      int32_t* qp_iter, int32_t* qp_status, int32_t* qp_ncond,
                                      unsigned long long* res_acc)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    status[b] = rti ? NMPC_STATUS_SUCCESS : NMPC_STATUS_MAXITER;
    done[b] = 0; sqp_iter[b] = 0; qp_iter[b] = 0; qp_status[b] = 0; qp_ncond[b] = 0;
    for (int i = 0; i < 4; i++) res_acc[(size_t)b * 4 + i] = 0ULL;
}

// termination test of SQP iteration `it` (acados: all four residual norms below nlp_tol)
__global__ void nmpc_sqp_decide_kernel(int b0, int Bc, int N, int it, int max_iter, double tol,
                                       unsigned long long* res_acc, const double* cost_stage,
                                       double* res, double* cost, int32_t* status, int32_t* done)
{
    const int bl = blockIdx.x * blockDim.x + threadIdx.x;
    if (bl >= Bc) return;
    const int b = b0 + bl;
    if (done[b]) return;
    double r[4];
    bool bad = false, conv = true;
    for (int i = 0; i < 4; i++) {
        r[i] = __longlong_as_double((long long)res_acc[(size_t)b * 4 + i]);
        res_acc[(size_t)b * 4 + i] = 0ULL;
        res[(size_t)b * 4 + i] = r[i];
        if (isinf(r[i]) || !(r[i] == r[i])) bad = true;
        if (!(r[i] < tol)) conv = false;
    }
    double c = 0.0;
    for (int k = 0; k <= N; k++) c += cost_stage[(size_t)bl * (N + 1) + k];
    cost[b] = c;
    if (bad) { status[b] = NMPC_STATUS_NAN; done[b] = 1; }
    else if (conv) { status[b] = NMPC_STATUS_SUCCESS; done[b] = 1; }
    else if (it >= max_iter) { status[b] = NMPC_STATUS_MAXITER; done[b] = 1; }
}

// after a QP: a failed QP (HPIPM status other than success / max-iter) ends the problem
__global__ void nmpc_sqp_post_qp_kernel(int b0, int Bc, const int32_t* qp_status, int32_t* status, int32_t* done)
{
    const int bl = blockIdx.x * blockDim.x + threadIdx.x;
    if (bl >= Bc) return;
    const int b = b0 + bl;
    if (done[b]) return;
    if (qp_status[b] != 0 && qp_status[b] != 1) { status[b] = NMPC_STATUS_QP_FAILURE; done[b] = 1; }
}

// RTI cost bookkeeping (no residual test in SQP_RTI)
__global__ void nmpc_cost_sum_kernel(int b0, int Bc, int N, const double* cost_stage, double* cost)
{
    const int bl = blockIdx.x * blockDim.x + threadIdx.x;
    if (bl >= Bc) return;
    double c = 0.0;
    for (int k = 0; k <= N; k++) c += cost_stage[(size_t)bl * (N + 1) + k];
    cost[b0 + bl] = c;
}

// shift warm start (solver.py:290-342): one CTA per problem, data staged through shared memory
__global__ void __launch_bounds__(128, 4) nmpc_shift_kernel(int N, const int32_t* start_node, int repeat_last,
                                                         double* x, double* u, double* pi, double* lam,
                                                         double* dtn, double* lamt, double dt_nodes)
{
    const int b = blockIdx.x, tid = threadIdx.x;
    const int s = start_node[b];
    if (s <= 0 || s > N) return;
    const int nws = N - s;
    double* xb = x + (size_t)b * (N + 1) * NX;
    double* ub = u + (size_t)b * N * NU;
    double* pb = pi + (size_t)b * N * NX;
    double* lb = lam + (size_t)b * (N + 1) * 2 * NG;
    // forward-moving copies: process node by node so a source is never overwritten before use
    for (int k = 1; k <= nws; k++) {
        for (int i = tid; i < NX; i += 128) xb[k * NX + i] = xb[(k + s) * NX + i];
        __syncthreads();
    }
    for (int k = 0; k < nws; k++) {
        for (int i = tid; i < NU; i += 128) ub[k * NU + i] = ub[(k + s) * NU + i];
        for (int i = tid; i < NX; i += 128) pb[k * NX + i] = pb[(k + s) * NX + i];
        __syncthreads();
    }
    for (int k = 0; k <= nws; k++) {
        for (int i = tid; i < 2 * NG; i += 128) lb[k * 2 * NG + i] = lb[(k + s) * 2 * NG + i];
        __syncthreads();
    }
    if (dtn && tid == 0) {
        // enable_time_opt (solver.py:324-336): dt of the warm-started nodes from the previous solution, nominal on the tail
        double* db = dtn + (size_t)b * N;
        double* lt = lamt + (size_t)b * N * 2;
        for (int k = 0; k < nws; k++) { db[k] = db[k + s]; lt[2 * k] = lt[2 * (k + s)]; lt[2 * k + 1] = lt[2 * (k + s) + 1]; }
        for (int k = nws; k < N; k++) db[k] = dt_nodes;
        if (nws < N) { lt[2 * nws] = 0.0; lt[2 * nws + 1] = 0.0; }      // shifted in from the terminal node, which has no dt
    }
    // force tail
    for (int e = tid; e < (N - nws) * 12; e += 128) {
        const int k = nws + e / 12, i = e % 12;
        ub[k * NU + IFRC + i] = repeat_last ? ub[(N - 1) * NU + IFRC + i] : 0.0;
    }
    __syncthreads();
    if (repeat_last && nws < N) {
        for (int e = tid; e < (N + 1 - nws) * NX; e += 128) {
            const int k = nws + e / NX, i = e % NX;
            if (k < N) xb[k * NX + i] = xb[N * NX + i];
        }
        for (int e = tid; e < (N - nws) * NQ; e += 128) {
            const int k = nws + e / NQ, i = e % NQ;
            if (k < N - 1) ub[k * NU + i] = ub[(N - 1) * NU + i];
        }
    }
}

// Packed result rows for the gather to the dataset-writer rank: per problem
// [x (N+1)*42 | u N*30 | dt_node N | status] = 1868 doubles at N = 25 -- the 1867 numbers parse_sol returns
// (solver.py:403-429: q, v, h per node, a, f, dt per interval) plus the solver status.
__global__ void __launch_bounds__(128) nmpc_pack_solution_kernel(int N, double dt, const double* __restrict__ dtn,
                                                                 const double* __restrict__ x,
                                                                 const double* __restrict__ u,
                                                                 const int32_t* __restrict__ status,
                                                                 double* __restrict__ out)
{
    const int b = blockIdx.x, tid = threadIdx.x;
    const int nx = (N + 1) * NX, nu = N * NU, row = nx + nu + N + 1;
    double* o = out + (size_t)b * row;
    const double* xb = x + (size_t)b * nx;
    const double* ub = u + (size_t)b * nu;
    for (int i = tid; i < nx; i += 128) o[i] = xb[i];
    for (int i = tid; i < nu; i += 128) o[nx + i] = ub[i];
    for (int i = tid; i < N; i += 128) o[nx + nu + i] = dtn ? dtn[(size_t)b * N + i] : dt;
    if (tid == 0) o[row - 1] = (double)status[b];
}
