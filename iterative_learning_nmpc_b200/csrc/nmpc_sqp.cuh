// nmpc_sqp.cuh -- SQP / SQP_RTI driver kernels: NLP residuals, termination, shift warm start.
//
// Device-side counterpart of acados' ocp_nlp_sqp / ocp_nlp_sqp_rti outer loop
// and of QuadrupedAcadosSolver.warm_start_solver (mpc_controller/utils/solver.py:290-342).
// No host round trip happens inside a solve: every problem carries a `done` flag that the
// linearise / QP kernels test on entry.
#pragma once
#include "nmpc_common.cuh"

#define RES_THREADS 128

struct ResArgs {
    const double* G; const double* M; const double* C; const double* D;
    const double* x; const double* x0; const double* pi; const double* lam;
    unsigned long long* res_acc;   // [B][4] bit patterns of non-negative doubles (atomicMax)
    const int32_t* done;
    int N, b0;
};

__device__ __forceinline__ void atomic_max_nonneg(unsigned long long* addr, double v)
{
    // NaN is reported as +inf so that it survives the max
    if (!(v == v)) v = __longlong_as_double(0x7ff0000000000000LL);
    atomicMax(addr, (unsigned long long)__double_as_longlong(v));
}

// inf-norms of the KKT residuals of the NLP at the current linearisation: one CTA per node
__global__ void __launch_bounds__(RES_THREADS, 4) nmpc_residual_kernel(ResArgs a)
{
    const int k = blockIdx.x, bl = blockIdx.y, b = a.b0 + bl, tid = threadIdx.x;
    if (a.done && a.done[b]) return;
    const int N = a.N;
    const bool term = (k == N);
    __shared__ double s_pik[NX], s_pin[NX], s_dl[NG], s_red[4][RES_THREADS / 32];
    const double* Gp = term ? nullptr : a.G + ((size_t)bl * N + k) * G_SZ;
    const double* Mp = a.M + ((size_t)bl * (N + 1) + k) * M_SZ;
    const double* Cp = a.C + ((size_t)bl * (N + 1) + k) * C_SZ;
    const double* Dp = a.D + ((size_t)bl * (N + 1) + k) * D_SZ;
    const double* lam = a.lam + ((size_t)b * (N + 1) + k) * 2 * NG;
    double r_s = 0, r_e = 0, r_i = 0, r_c = 0;
    for (int i = tid; i < NX; i += RES_THREADS) {
        s_pik[i] = k > 0 ? a.pi[((size_t)b * N + k - 1) * NX + i] : 0.0;
        s_pin[i] = term ? 0.0 : a.pi[((size_t)b * N + k) * NX + i];
        if (k == 0) r_e = nanmax(r_e, fabs(a.x[(size_t)b * (N + 1) * NX + i] - a.x0[(size_t)b * NX + i]));
        if (!term) r_e = nanmax(r_e, fabs(Gp[i * G_LD + NZ]));
    }
    for (int j = tid; j < NG; j += RES_THREADS) {
        const double ml = Dp[2 * NG + j], mu = Dp[3 * NG + j];
        const double c = -Dp[j];
        double dl = 0.0;
        if (ml > 0.5) { dl += lam[j]; r_i = nanmax(r_i, -c); r_c = nanmax(r_c, fabs(lam[j] * c)); }
        if (mu > 0.5) { dl -= lam[NG + j]; r_i = nanmax(r_i, c); r_c = nanmax(r_c, fabs(lam[NG + j] * c)); }
        s_dl[j] = dl;
    }
    __syncthreads();
    for (int c = tid; c < NZ; c += RES_THREADS) {
        if ((k == 0 && c >= NU) || (term && c < NU)) continue;
        double s = Mp[NZ * M_LD + c];
        if (!term) for (int l = 0; l < NX; l++) s += Gp[l * G_LD + c] * s_pin[l];
        if (k > 0 && c >= NU) s -= s_pik[c - NU];
        for (int j = 0; j < NG; j++) s -= Cp[j * C_LD + c] * s_dl[j];
        r_s = nanmax(r_s, fabs(s));
    }
    r_s = warp_max(r_s); r_e = warp_max(r_e); r_i = warp_max(r_i); r_c = warp_max(r_c);
    if ((tid & 31) == 0) { s_red[0][tid >> 5] = r_s; s_red[1][tid >> 5] = r_e; s_red[2][tid >> 5] = r_i; s_red[3][tid >> 5] = r_c; }
    __syncthreads();
    if (tid < 4) {
        double v = 0.0;
        for (int w = 0; w < RES_THREADS / 32; w++) v = nanmax(v, s_red[tid][w]);
        atomic_max_nonneg(a.res_acc + (size_t)b * 4 + tid, v);
    }
}

// start of a solve: clear flags
__global__ void nmpc_sqp_begin_kernel(int B, int rti, int32_t* status, int32_t* done, int32_t* sqp_iter,
                                      int32_t* qp_iter, int32_t* qp_status, unsigned long long* res_acc)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    status[b] = rti ? NMPC_STATUS_SUCCESS : NMPC_STATUS_MAXITER;
    done[b] = 0; sqp_iter[b] = 0; qp_iter[b] = 0; qp_status[b] = 0;
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
                                                         double* x, double* u, double* pi, double* lam)
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
