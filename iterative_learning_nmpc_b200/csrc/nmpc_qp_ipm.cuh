// nmpc_qp_ipm.cuh -- OCP-QP interior point on the device, fp64.
//
// Replaces HPIPM's d_ocp_qp_ipm (mode `speed`) + BLASFEO as used by acados for the reference's
// solver: Mehrotra predictor-corrector on the OCP-QP of one SQP step,
// backward square-root Riccati factorisation over the stage blocks, forward sweep, a second
// (vector-only) sweep pair for the corrector on the same factorisation, split primal/dual step.
// The algorithm is the one restated in oracle/nmpc_oracle.c::qp_ipm (the checker).
//
// B200 mapping.  An IPM iteration is split by data dependence:
//   * everything that is independent across the N+1 stages (residuals, Gamma/gamma, the barrier
//     Hessian H + C' Gamma C, slack/multiplier directions, step-length candidates, the variable
//     update) runs as STAGE-PARALLEL kernels, one CTA per (stage, problem): 26 x B CTAs stream
//     the stage blocks with bulk TMA copies and keep all 148 SMs busy;
//   * only the Riccati recursion is sequential in the stage index.  It runs one CTA per OCP with
//     160 threads and ~54 KB of shared memory, i.e. FOUR OCPs per SM whose barrier stalls overlap.
//     The stage Hessian lives in REGISTERS (3x6 tiles) from the rank-42 update AL'AL through a
//     fused right-looking Cholesky; only the current 6-column panel goes through shared memory,
//     and the factor is kept panel-packed (2880 doubles instead of 73 x 73): one contiguous bulk
//     store, half the HBM bytes, and L_{k+1}^{xx} simply stays where it is for the next stage.
// Reductions use fixed trees or order-independent atomics (max / min / integer add), so a
// problem's iterates do not depend on its position in the batch.
#pragma once
#include "nmpc_common.cuh"

// ---------------------------------------------------------------------------------------------
// panel-packed lower storage of a 73 x 72 matrix (72 x 72 lower triangle + appended row 72):
// twelve 6-column panels; panel jb holds rows 6*jb .. 72, row-major with 6 doubles per row.
// ---------------------------------------------------------------------------------------------
#define PB 6
#define NPANEL 12
__host__ __device__ constexpr int P_OFF(int jb) { return jb * (456 - 18 * jb); }
#define LP_SZ 2880                       // P_OFF(12)
#define LP_UPANELS 1830                  // P_OFF(5): panels holding Luu, Lxu, lu
#define LX_SZ (LP_SZ - LP_UPANELS)       // 1050: panels holding Lxx, lx
__host__ __device__ __forceinline__ int lp_idx(int i, int j)
{
    const int jb = j / PB;
    return P_OFF(jb) + (i - PB * jb) * PB + (j - PB * jb);
}

// per-stage slots of the per-problem vector workspace (doubles)
#define V_RG 0
#define V_RB (V_RG + NZ)
#define V_RD (V_RB + NX)
#define V_RM (V_RD + 2 * NG)
#define V_QT (V_RM + 2 * NG)
#define V_DZ (V_QT + NZ)
#define V_DPI (V_DZ + NZ)
#define V_DT (V_DPI + NX)
#define V_DLAM (V_DT + 2 * NG)
#define V_P (V_DLAM + 2 * NG)
#define V_LROW (V_P + NX)                // l = L^-1 m of the corrector solve
#define V_STAGE 752                      // 750 rounded up to a multiple of 8

#define QP_THREADS 160                   // 156 tile threads + 4 threads owning row 72
#define QP_WARPS (QP_THREADS / 32)
#define N_TILES 156                      // 3 x 6 tiles of the lower triangle of a 72 x 72 matrix

#define POS_INF_BITS 0x7ff0000000000000ULL
#define ONE_BITS 0x3ff0000000000000ULL

struct QpArgs {
    const double* G; const double* M; const double* C; const double* D;   // chunk-local stage blocks
    double* Ht;                     // [Bc][N+1][LP_SZ]  H + C' Gamma C, panel-packed
    double* L;                      // [Bc][N+1][LP_SZ]  Riccati factors, panel-packed
    double* vec;                    // [Bc][(N+1)*V_STAGE]
    double* qz; double* qpi; double* qlam; double* qt;   // QP iterate, global problem index
    unsigned long long* res_acc;    // [B][4] running inf-norms (bit patterns; NaN stored as +inf)
    unsigned long long* alpha;      // [B][2] step-length candidates (bit patterns of doubles in (0,1])
    int32_t* nc;                    // [B]    number of active inequality sides
    double* mu_part;                // [B][N+1] per-stage sum of lam*t
    double* mu;                     // [B]
    double* sigma_mu;               // [B]
    int32_t* qp_done;               // [B]
    int32_t* qp_iter; int32_t* qp_status; double* qp_res;
    const double* dx0_ext;          // [B][NX] or null -> x0 - x[0]
    const double* x0;
    double* x; double* u; double* pi; double* lam;       // SQP iterate (updated by qp_apply)
    const int32_t* sqp_done; int32_t* sqp_iter;
    int N, b0, max_iter, warm, it, last;
    double tol;
};

__device__ __forceinline__ void atomic_max_norm(unsigned long long* addr, double v)
{
    if (!(v == v)) v = __longlong_as_double((long long)POS_INF_BITS);   // NaN survives the max as +inf
    atomicMax(addr, (unsigned long long)__double_as_longlong(v));
}

// block-wide reductions over QP_THREADS threads with a fixed tree (deterministic)
__device__ __forceinline__ double block_max(double v, double* red, int tid)
{
    v = warp_max(v);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    double r = red[0];
#pragma unroll
    for (int w = 1; w < QP_WARPS; w++) r = nanmax(r, red[w]);
    return r;
}
__device__ __forceinline__ double block_sum(double v, double* red, int tid)
{
    v = warp_sum(v);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    double r = red[0];
#pragma unroll
    for (int w = 1; w < QP_WARPS; w++) r += red[w];
    return r;
}

// tile (3 rows x 6 columns) owned by thread t < N_TILES: column block cb in 0..11, row block rb in
// 2*cb..23.  Tiles are numbered column block by column block, so that in the right-looking
// factorisation whole warps run out of trailing work together (warp 0 after panel 1, ...).
__device__ __forceinline__ void tile_of(int t, int& rb, int& cb)
{
    int c = 0;
    while ((c + 1) * (24 - c) <= t) c++;          // first tile of column block c is c * (25 - c)
    cb = c;
    rb = 2 * c + (t - c * (25 - c));
}

__device__ __forceinline__ void load6(const double* p, double* v)
{
    const double2 b0 = *reinterpret_cast<const double2*>(p);
    const double2 b1 = *reinterpret_cast<const double2*>(p + 2);
    const double2 b2 = *reinterpret_cast<const double2*>(p + 4);
    v[0] = b0.x; v[1] = b0.y; v[2] = b1.x; v[3] = b1.y; v[4] = b2.x; v[5] = b2.y;
}

// =============================================================================================
// K0: start of a QP.  One CTA per problem.
// =============================================================================================
__global__ void __launch_bounds__(128) nmpc_qp_init_kernel(QpArgs a)
{
    const int bl = blockIdx.x, b = a.b0 + bl, tid = threadIdx.x, N = a.N;
    const bool skip = a.sqp_done && a.sqp_done[b];
    if (tid == 0) {
        a.qp_done[b] = skip ? 1 : 0;
        if (!skip) { a.qp_iter[b] = 0; a.qp_status[b] = 1; }     // finished problems keep their statistics
        a.nc[b] = 0;
        a.alpha[2 * b] = ONE_BITS; a.alpha[2 * b + 1] = ONE_BITS;
        for (int i = 0; i < 4; i++) a.res_acc[4 * b + i] = 0ULL;
    }
    if (skip) return;
    double* qz = a.qz + (size_t)b * (N + 1) * NZ;
    double* qpi = a.qpi + (size_t)b * N * NX;
    if (a.warm == 0) {
        for (int i = tid; i < (N + 1) * NZ; i += 128) qz[i] = 0.0;
        for (int i = tid; i < N * NX; i += 128) qpi[i] = 0.0;
    }
    __syncthreads();
    if (tid < NX)
        qz[NU + tid] = a.dx0_ext ? a.dx0_ext[(size_t)b * NX + tid]
                                 : a.x0[(size_t)b * NX + tid] - a.x[(size_t)b * (N + 1) * NX + tid];
    if (tid < NU) qz[(size_t)N * NZ + tid] = 0.0;
}

// =============================================================================================
// K1: stage-parallel residuals, Gamma/gamma, barrier Hessian.  One CTA per (stage, problem).
//     shared: C (42x72) | M (73x73+1: full symmetric H + gradient row) | vectors
// =============================================================================================
#define K1_C 0
#define K1_M (K1_C + C_SZ)
#define K1_V (K1_M + M_SZ)
#define K1V_Z 0
#define K1V_ZXN (K1V_Z + NZ)
#define K1V_PIK (K1V_ZXN + NX)
#define K1V_PIN (K1V_PIK + NX)
#define K1V_LAM (K1V_PIN + NX)
#define K1V_T (K1V_LAM + 2 * NG)
#define K1V_D (K1V_T + 2 * NG)           // lg | ug | ml | mu
#define K1V_GAM (K1V_D + 4 * NG)
#define K1V_GAMV (K1V_GAM + NG)
#define K1V_DL (K1V_GAMV + NG)
#define K1V_CZ (K1V_DL + NG)
#define K1V_RD (K1V_CZ + NG)
#define K1V_RM (K1V_RD + 2 * NG)
#define K1V_RED (K1V_RM + 2 * NG)
#define K1V_END (K1V_RED + 8)
#define K1_SMEM_BYTES ((K1_V + K1V_END) * 8)

__global__ void __launch_bounds__(QP_THREADS, 3) nmpc_qp_stage_kernel(QpArgs a)
{
    extern __shared__ __align__(16) double smem[];
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int k = blockIdx.x, bl = blockIdx.y, b = a.b0 + bl, N = a.N;
    if (a.qp_done[b]) return;
    const bool term = (k == N);
    double* sC = smem + K1_C;
    double* sM = smem + K1_M;
    double* sv = smem + K1_V;
    double* red = sv + K1V_RED;

    const double* Gp = term ? nullptr : a.G + ((size_t)bl * N + k) * G_SZ;
    const double* Mp = a.M + ((size_t)bl * (N + 1) + k) * M_SZ;
    const double* Cp = a.C + ((size_t)bl * (N + 1) + k) * C_SZ;
    const double* Dp = a.D + ((size_t)bl * (N + 1) + k) * D_SZ;
    double* vk = a.vec + ((size_t)bl * (N + 1) + k) * V_STAGE;
    const double* qz = a.qz + (size_t)b * (N + 1) * NZ;
    const double* qpi = a.qpi + (size_t)b * N * NX;
    double* qlam = a.qlam + ((size_t)b * (N + 1) + k) * 2 * NG;
    double* qt = a.qt + ((size_t)b * (N + 1) + k) * 2 * NG;

    if (tid == 0) {
        mbar_init(&bar, 1);
        fence_barrier_init();
        mbar_arrive_expect_tx(&bar, (C_SZ + M_SZ) * 8);
        bulk_g2s(sC, Cp, C_SZ * 8, &bar);
        bulk_g2s(sM, Mp, M_SZ * 8, &bar);
    }
    for (int i = tid; i < NZ; i += QP_THREADS) sv[K1V_Z + i] = qz[(size_t)k * NZ + i];
    for (int i = tid; i < NX; i += QP_THREADS) {
        sv[K1V_ZXN + i] = term ? 0.0 : qz[(size_t)(k + 1) * NZ + NU + i];
        sv[K1V_PIK + i] = (k > 0) ? qpi[(size_t)(k - 1) * NX + i] : 0.0;
        sv[K1V_PIN + i] = term ? 0.0 : qpi[(size_t)k * NX + i];
    }
    for (int i = tid; i < 4 * NG; i += QP_THREADS) sv[K1V_D + i] = Dp[i];
    for (int i = tid; i < 2 * NG; i += QP_THREADS) { sv[K1V_LAM + i] = qlam[i]; sv[K1V_T + i] = qt[i]; }
    __syncthreads();
    mbar_wait(&bar, 0);

    // ---- C z (warp per row) ----
    for (int j = warp; j < NG; j += QP_WARPS) {
        double s = 0.0;
        for (int c = lane; c < NZ; c += 32) s += sC[j * C_LD + c] * sv[K1V_Z + c];
        s = warp_sum(s);
        if (lane == 0) sv[K1V_CZ + j] = s;
    }
    __syncthreads();
    double m_g = 0.0, m_b = 0.0, m_d = 0.0, m_m = 0.0, acc_mu = 0.0;
    int cnt = 0;
    if (tid < 2 * NG) {
        const int e = tid / NG, j = tid % NG;
        const double msk = sv[K1V_D + 2 * NG + tid];
        const double cz = sv[K1V_CZ + j];
        const double slack = e ? (sv[K1V_D + NG + j] - cz) : (cz - sv[K1V_D + j]);
        double t = sv[K1V_T + tid], lam = sv[K1V_LAM + tid];
        if (a.it == 0 && (a.warm < 2 || !(t > 0.0) || !(lam > 0.0))) {
            if (msk > 0.5) { t = fmax(slack, IPM_THR0); lam = IPM_MU0 / t; } else { t = 1.0; lam = 0.0; }
            sv[K1V_T + tid] = t; sv[K1V_LAM + tid] = lam;
            qt[tid] = t; qlam[tid] = lam;
        }
        const double rd = msk > 0.5 ? (slack - t) : 0.0;
        const double rm = msk > 0.5 ? lam * t : 0.0;
        sv[K1V_RD + tid] = rd; sv[K1V_RM + tid] = rm;
        vk[V_RD + tid] = rd; vk[V_RM + tid] = rm;
        acc_mu = rm; cnt = (msk > 0.5) ? 1 : 0;
        m_d = fabs(rd); m_m = fabs(rm);
    }
    __syncthreads();
    if (tid < NG) {
        const int j = tid;
        const double ml = sv[K1V_D + 2 * NG + j], mu_ = sv[K1V_D + 3 * NG + j];
        const double ll = sv[K1V_LAM + j], lu = sv[K1V_LAM + NG + j];
        const double tl = sv[K1V_T + j], tu = sv[K1V_T + NG + j];
        double Gam = 0.0, gam = 0.0, dl = 0.0;
        if (ml > 0.5) { Gam += ll / tl; gam += (sv[K1V_RM + j] + ll * sv[K1V_RD + j]) / tl; dl += ll; }
        if (mu_ > 0.5) { Gam += lu / tu; gam -= (sv[K1V_RM + NG + j] + lu * sv[K1V_RD + NG + j]) / tu; dl -= lu; }
        sv[K1V_GAM + j] = Gam; sv[K1V_GAMV + j] = gam; sv[K1V_DL + j] = dl;
    }
    __syncthreads();
    // ---- stationarity residual and q tilde: thread per column (threads 0..71);
    //      dynamics residual rb = G [z;1] - zx+ : warps 3,4, one row per warp pass ----
    if (tid < NZ) {
        const int c = tid;
        double s = sM[NZ * M_LD + c];
        for (int d = 0; d < NZ; d++) s += sM[c * M_LD + d] * sv[K1V_Z + d];
        if (c >= NU && k > 0) s -= sv[K1V_PIK + c - NU];
        double qp = 0.0;
        for (int j = 0; j < NG; j++) {
            const double cj = sC[j * C_LD + c];
            s -= cj * sv[K1V_DL + j];
            qp += cj * sv[K1V_GAMV + j];
        }
        if (!term) for (int l = 0; l < NX; l++) s += Gp[l * G_LD + c] * sv[K1V_PIN + l];
        if ((k == 0 && c >= NU) || (term && c < NU)) s = 0.0;
        vk[V_RG + c] = s; vk[V_QT + c] = qp + s;
        m_g = fabs(s);
    }
    if (!term && warp >= 3) {
        for (int l = warp - 3; l < NX; l += QP_WARPS - 3) {
            double s = 0.0;
            for (int c = lane; c < NZ; c += 32) s += Gp[l * G_LD + c] * sv[K1V_Z + c];
            s = warp_sum(s);
            if (lane == 0) {
                const double rb = s + Gp[l * G_LD + NZ] - sv[K1V_ZXN + l];
                vk[V_RB + l] = rb;
                m_b = nanmax(m_b, fabs(rb));
            }
        }
    }
    // ---- norms, mu ----
    m_g = block_max(m_g, red, tid);
    m_b = block_max(m_b, red, tid);
    m_d = block_max(m_d, red, tid);
    m_m = block_max(m_m, red, tid);
    const double mu_sum = block_sum(acc_mu, red, tid);
    if (a.it == 0) {
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        if (lane == 0 && cnt) atomicAdd(a.nc + b, cnt);
    }
    if (tid == 0) {
        atomic_max_norm(a.res_acc + 4 * b + 0, m_g);
        atomic_max_norm(a.res_acc + 4 * b + 1, m_b);
        atomic_max_norm(a.res_acc + 4 * b + 2, m_d);
        atomic_max_norm(a.res_acc + 4 * b + 3, m_m);
        a.mu_part[(size_t)b * (N + 1) + k] = mu_sum;
    }
    if (a.last) return;

    // ---- barrier Hessian tiles: H + C' Gamma C, written panel-packed ----
    if (tid < N_TILES) {
        int rb, cb;
        tile_of(tid, rb, cb);
        const int r0 = 3 * rb, c0 = 6 * cb;
        double acc[3][6];
#pragma unroll
        for (int r = 0; r < 3; r++)
#pragma unroll
            for (int c = 0; c < 6; c++) acc[r][c] = sM[(r0 + r) * M_LD + c0 + c];
        for (int j = 0; j < NG; j++) {
            const double g = sv[K1V_GAM + j];
            if (g == 0.0) continue;
            const double* row = sC + j * C_LD;
            const double a0 = row[r0] * g, a1 = row[r0 + 1] * g, a2 = row[r0 + 2] * g;
            double bv[6];
            load6(row + c0, bv);
#pragma unroll
            for (int c = 0; c < 6; c++) { acc[0][c] += a0 * bv[c]; acc[1][c] += a1 * bv[c]; acc[2][c] += a2 * bv[c]; }
        }
        double* out = a.Ht + ((size_t)bl * (N + 1) + k) * LP_SZ + P_OFF(cb) + (r0 - c0) * PB;
#pragma unroll
        for (int r = 0; r < 3; r++) {
            double2* o2 = reinterpret_cast<double2*>(out + r * PB);
            o2[0] = make_double2(acc[r][0], acc[r][1]);
            o2[1] = make_double2(acc[r][2], acc[r][3]);
            o2[2] = make_double2(acc[r][4], acc[r][5]);
        }
    }
}

// =============================================================================================
// K2: exit tests of IPM iteration `it`.  One thread per problem.
// =============================================================================================
__global__ void nmpc_qp_decide_kernel(QpArgs a, int Bc)
{
    const int bl = blockIdx.x * blockDim.x + threadIdx.x;
    if (bl >= Bc) return;
    const int b = a.b0 + bl, N = a.N;
    if (a.qp_done[b]) return;
    double r[4];
    bool bad = false, conv = true;
    for (int i = 0; i < 4; i++) {
        r[i] = __longlong_as_double((long long)a.res_acc[4 * b + i]);
        a.res_acc[4 * b + i] = 0ULL;
        a.qp_res[(size_t)b * 4 + i] = r[i];
        if (isinf(r[i]) || !(r[i] == r[i])) bad = true;
        if (!(r[i] < a.tol)) conv = false;
    }
    double s = 0.0;
    for (int k = 0; k <= N; k++) s += a.mu_part[(size_t)b * (N + 1) + k];
    const int nc = a.nc[b];
    a.mu[b] = nc > 0 ? s / nc : 0.0;
    a.qp_iter[b] = a.it;
    a.alpha[2 * b] = ONE_BITS; a.alpha[2 * b + 1] = ONE_BITS;   // consumed by the update of this iteration
    if (bad) { a.qp_status[b] = 3; a.qp_done[b] = 1; }
    else if (conv) { a.qp_status[b] = 0; a.qp_done[b] = 1; }
    else if (a.it >= a.max_iter) { a.qp_status[b] = 1; a.qp_done[b] = 1; }
}

// =============================================================================================
// sequential kernels: shared-memory carve-up (doubles)
// =============================================================================================
#define S_LP 0                              // current factor, panel-packed (LP_SZ)
#define S_U (S_LP + LP_SZ)                  // K3: G -> AL (42 x 74);  K7: Lxx panels of stage k+1 (LX_SZ)
#define SV_QT 0                             // q tilde / m                 (NZ)
#define SV_RB (SV_QT + NZ)                  //                             (NX)
#define SV_W (SV_RB + NX)                   //                             (NX)
#define SV_TMP (SV_W + NX)                  //                             (NZ)
#define SV_PN (SV_TMP + NZ)                 // p_{k+1}                     (NX)
#define SV_DX (SV_PN + NX)
#define SV_DXN (SV_DX + NX)
#define SV_Z (SV_DXN + NX)                  // [du; dx]                    (NZ)
#define SV_INV (SV_Z + NZ)                  // reciprocal diagonal of L    (NZ)
#define SV_AL13 (SV_INV + NZ)               // rows 39..41 of AL before the in-place write (3 x 72)
#define SV_END (SV_AL13 + 3 * NZ)
#define K3_SMEM_BYTES ((S_U + G_SZ + SV_END) * 8)
#define K7_SMEM_BYTES ((S_U + LX_SZ + SV_END) * 8)

// w = L+ (L+' rb) + p+ with L+ = Lxx (42 x 42 lower) read through `lx(l, i)`; result in sv[SV_W]
template <typename LX>
__device__ __forceinline__ void value_fn_times_rb(const LX& lx, double* sv, int tid)
{
    if (tid < NX) {
        double s = 0.0;
        for (int l = tid; l < NX; l++) s += lx(l, tid) * sv[SV_RB + l];
        sv[SV_TMP + tid] = s;
    }
    __syncthreads();
    if (tid < NX) {
        double s = sv[SV_PN + tid];
        for (int l = 0; l <= tid; l++) s += lx(tid, l) * sv[SV_TMP + l];
        sv[SV_W + tid] = s;
    }
    __syncthreads();
}

// One stage of the forward sweep.  sLp holds L_k (at least the u panels; all panels if DPI).
// in: sv[SV_DX] = dx_k, sv[SV_RB], (DPI) sv[SV_PN] = p_k;  out: dz_k to vk, sv[SV_DX] = dx_{k+1}
template <bool DPI>
__device__ __forceinline__ void forward_stage(const double* sLp, double* sv, const double* Gp, double* vk,
                                              double* dpi_prev, bool term, int tid)
{
    const int warp = tid >> 5, lane = tid & 31;
    if (DPI && dpi_prev) {          // dpi_{k-1} = Lxx (Lxx' dx) + p_k
        if (tid < NX) {
            double s = 0.0;
            for (int l = tid; l < NX; l++) s += sLp[lp_idx(NU + l, NU + tid)] * sv[SV_DX + l];
            sv[SV_TMP + tid] = s;
        }
        __syncthreads();
        if (tid < NX) {
            double s = sv[SV_PN + tid];
            for (int l = 0; l <= tid; l++) s += sLp[lp_idx(NU + tid, NU + l)] * sv[SV_TMP + l];
            dpi_prev[tid] = s;
        }
    }
    if (!term) {
        if (warp == 0) {            // du = -Luu^-T (lu + Lxu' dx)
            double rhs = 0.0;
            if (lane < NU) {
                rhs = sLp[lp_idx(NZ, lane)];
                for (int l = 0; l < NX; l++) rhs += sLp[lp_idx(NU + l, lane)] * sv[SV_DX + l];
            }
            double du = 0.0;
            for (int i = NU - 1; i >= 0; i--) {
                const double d = sLp[lp_idx(i, i)];
                const double inv = d != 0.0 ? 1.0 / d : 0.0;
                const double dui = -__shfl_sync(0xffffffffu, rhs, i) * inv;
                if (lane == i) du = dui;
                if (lane < i) rhs += sLp[lp_idx(i, lane)] * dui;
            }
            if (lane < NU) sv[SV_Z + lane] = du;
        }
    } else if (tid < NU) sv[SV_Z + tid] = 0.0;
    if (tid >= 32 && tid < 32 + NX) sv[SV_Z + NU + tid - 32] = sv[SV_DX + tid - 32];
    __syncthreads();
    for (int i = tid; i < NZ; i += QP_THREADS) vk[V_DZ + i] = sv[SV_Z + i];
    if (!term) {                    // dx+ = G [du; dx] + rb : G streamed from L2/HBM; all loads of a warp's
                                    // nine rows are issued before the first reduction (memory-level parallelism)
        const double z0 = sv[SV_Z + lane], z1 = sv[SV_Z + 32 + lane];
        const double z2 = lane < NZ - 64 ? sv[SV_Z + 64 + lane] : 0.0;
        double s[9];
#pragma unroll
        for (int i = 0; i < 9; i++) {
            const int l = warp + QP_WARPS * i;
            s[i] = 0.0;
            if (l < NX) {
                const double* row = Gp + l * G_LD;
                s[i] = row[lane] * z0 + row[32 + lane] * z1;
                if (lane < NZ - 64) s[i] += row[64 + lane] * z2;
            }
        }
#pragma unroll
        for (int i = 0; i < 9; i++) {
            const int l = warp + QP_WARPS * i;
            const double r = warp_sum(s[i]);
            if (lane == 0 && l < NX) sv[SV_DXN + l] = r + sv[SV_RB + l];
        }
    }
    __syncthreads();
    if (!term && tid < NX) sv[SV_DX + tid] = sv[SV_DXN + tid];
    __syncthreads();
}

// Phase-cycle instrumentation of K3 (debug builds with -DNMPC_TIMING only): CTA 0, thread 0
// accumulates clock64() deltas per phase into g_phase_cycles.
#ifdef NMPC_TIMING
__device__ unsigned long long g_phase_cycles[16];
// register accumulators (static indices), flushed once at the end of the kernel
#define TMARK(ph) do { const long long t_ = clock64(); t_acc[ph] += (unsigned long long)(t_ - t_prev); t_prev = t_; } while (0)
#else
#define TMARK(ph) do { } while (0)
#endif

// =============================================================================================
// K3: backward square-root Riccati factorisation + forward sweep (affine direction).
//     One CTA per OCP; 4 CTAs per SM.
// =============================================================================================
__global__ void __launch_bounds__(QP_THREADS, 4) nmpc_qp_factor_kernel(QpArgs a)
{
    extern __shared__ __align__(16) double smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ unsigned s_gmask[24];
    const int tid = threadIdx.x;
    const int bl = blockIdx.x, b = a.b0 + bl, N = a.N;
    if (a.qp_done[b]) return;
    double* sLp = smem + S_LP;
    double* sU = smem + S_U;
    double* sv = smem + S_U + G_SZ;

    const double* Gb = a.G + (size_t)bl * N * G_SZ;
    const double* Hb = a.Ht + (size_t)bl * (N + 1) * LP_SZ;
    double* Lb = a.L + (size_t)bl * (N + 1) * LP_SZ;
    double* vb = a.vec + (size_t)bl * (N + 1) * V_STAGE;

    uint32_t phase = 0;
#ifdef NMPC_TIMING
    unsigned long long t_acc[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    long long t_prev = clock64();
#endif
    if (tid == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    for (int i = tid; i < LP_SZ; i += QP_THREADS) sLp[i] = 0.0;

    // this thread's 3x6 tile of the stage Hessian
    const bool is_tile = tid < N_TILES;
    int rb = 0, cb = 0;
    if (is_tile) tile_of(tid, rb, cb);
    const int rcb0 = 3 * (tid - N_TILES);          // threads 156..159: column blocks of the short AL tiles
    const int r0 = 3 * rb, c0 = 6 * cb;
    const int doff = r0 - c0;                      // diagonal tiles: column of the diagonal in row 0 (0 or 3)
    const bool diag_tile = is_tile && (rb / 2 == cb);
    __syncthreads();

    for (int k = N; k >= 0; k--) {
        const bool term = (k == N);
        const double* vk = vb + (size_t)k * V_STAGE;
        const double* Hk = Hb + (size_t)k * LP_SZ;
        const int jb0 = term ? 5 : 0;
        double acc[3][6];

        for (int i = tid; i < NZ; i += QP_THREADS) sv[SV_QT + i] = vk[V_QT + i];
        if (!term) for (int i = tid; i < NX; i += QP_THREADS) sv[SV_RB + i] = vk[V_RB + i];
        __syncthreads();
        TMARK(0);

        if (!term) {
            // w = L+ (L+' rb) + p+ ; L+ = Lxx of stage k+1, still in the x panels of sLp
            value_fn_times_rb([&](int l, int i) { return sLp[lp_idx(NU + l, NU + i)]; }, sv, tid);
            TMARK(1);
            mbar_wait(&bar, phase); phase ^= 1;          // G_k has landed in sU
            TMARK(2);
            // m = q tilde + G' w
            if (tid < NZ) {
                double s = sv[SV_QT + tid];
                for (int l = 0; l < NX; l++) s += sU[l * G_LD + tid] * sv[SV_W + l];
                sv[SV_QT + tid] = s;
            }
            // rows of G that are non-zero inside each 6-column block (the dynamics are sparse: double
            // integrator rows + six momentum rows); generic in the data, dense G just sets every bit
            if (tid < 12 * 2) {
                const int cc = tid >> 1, half = tid & 1;
                unsigned mbits = 0u;
                for (int l = 21 * half; l < 21 * half + 21; l++) {
                    double gv[6];
                    load6(sU + l * G_LD + 6 * cc, gv);
                    const bool nz = (gv[0] != 0.0) | (gv[1] != 0.0) | (gv[2] != 0.0) | (gv[3] != 0.0) | (gv[4] != 0.0) | (gv[5] != 0.0);
                    if (nz) mbits |= 1u << (l - 21 * half);
                }
                s_gmask[tid] = mbits;
            }
            __syncthreads();
            TMARK(3);
            // AL = L+' G (42 x 72): 3 x 6 register tiles, rows 3*ib.., cols 6*cc..; the 156 tiles of
            // ib < 13 go to the tile threads, the 12 short tiles of ib = 13 to the last four threads
            {
                double al[3][6];
                const int npass = is_tile ? 1 : 3;
                for (int pass = 0; pass < npass; pass++) {
                    const int ib = is_tile ? tid / 12 : 13;
                    const int cc = is_tile ? tid % 12 : rcb0 + pass;
                    const int i0 = 3 * ib, cc0 = 6 * cc;
#pragma unroll
                    for (int r = 0; r < 3; r++)
#pragma unroll
                        for (int c = 0; c < 6; c++) al[r][c] = 0.0;
                    const int jbi = (NU + i0) / PB, ci = (NU + i0) % PB;     // panel / column of L+(:, i0)
                    const double* lcol = sLp + P_OFF(jbi) + ci;
                    unsigned long long rows = ((unsigned long long)s_gmask[2 * cc + 1] << 21) | s_gmask[2 * cc];
                    rows &= ~((1ULL << i0) - 1ULL);
                    while (rows) {
                        const int l = __ffsll((long long)rows) - 1;
                        rows &= rows - 1ULL;
                        // L+(l, i0..i0+2): entries above the diagonal are stored as zeros
                        const double* lp = lcol + (NU + l - PB * jbi) * PB;
                        const double l0 = lp[0], l1 = lp[1], l2 = lp[2];
                        double gv[6];
                        load6(sU + l * G_LD + cc0, gv);
#pragma unroll
                        for (int c = 0; c < 6; c++) { al[0][c] += l0 * gv[c]; al[1][c] += l1 * gv[c]; al[2][c] += l2 * gv[c]; }
                    }
                    if (!is_tile) {
#pragma unroll
                        for (int r = 0; r < 3; r++)
#pragma unroll
                            for (int c = 0; c < 6; c++) sv[SV_AL13 + r * NZ + cc0 + c] = al[r][c];
                    }
                }
                __syncthreads();          // every read of G is done: write AL in place
                TMARK(4);
                if (is_tile) {
                    const int i0 = 3 * (tid / 12), cc0 = 6 * (tid % 12);
#pragma unroll
                    for (int r = 0; r < 3; r++)
#pragma unroll
                        for (int c = 0; c < 6; c++) sU[(i0 + r) * G_LD + cc0 + c] = al[r][c];
                } else {
                    for (int e = tid - N_TILES; e < 3 * NZ; e += QP_THREADS - N_TILES)
                        sU[(39 + e / NZ) * G_LD + e % NZ] = sv[SV_AL13 + e];
                }
            }
            __syncthreads();
            TMARK(5);
        }
        // stage Hessian tile from HBM straight into registers (18 contiguous doubles), + AL' AL
        if (is_tile) {
            const double* hp = Hk + P_OFF(cb) + (r0 - c0) * PB;
#pragma unroll
            for (int r = 0; r < 3; r++) load6(hp + r * PB, acc[r]);
            if (!term) {
                for (int i = 0; i < NX; i++) {
                    const double* row = sU + i * G_LD;
                    const double a0 = row[r0], a1 = row[r0 + 1], a2 = row[r0 + 2];
                    double bv[6];
                    load6(row + c0, bv);
#pragma unroll
                    for (int c = 0; c < 6; c++) { acc[0][c] += a0 * bv[c]; acc[1][c] += a1 * bv[c]; acc[2][c] += a2 * bv[c]; }
                }
                if (diag_tile) {                       // diagonal regularisation
#pragma unroll
                    for (int r = 0; r < 3; r++)
#pragma unroll
                        for (int c = 0; c < 6; c++) if (c == doff + r) acc[r][c] += IPM_REG_PRIM;
                }
            }
        }
        __syncthreads();            // all reads of AL (sU) and of L+ (sLp) are done
        TMARK(6);
        if (tid == 0) {
            bulk_store_wait_read();               // the previous factor has left sLp
            if (k > 0) {                          // sU is free: prefetch G of the next stage
                mbar_arrive_expect_tx(&bar, G_SZ * 8);
                bulk_g2s(sU, Gb + (size_t)(k - 1) * G_SZ, G_SZ * 8, &bar);
            }
        }
        __syncthreads();
        TMARK(7);

        // ---- fused right-looking Cholesky on the register tiles; panels go through sLp ----
        for (int jb = jb0; jb < NPANEL; jb++) {
            const int j0 = PB * jb;
            double* pan = sLp + P_OFF(jb);                      // rows j0..72, 6 doubles each
            if (is_tile) {
                if (cb == jb) {
#pragma unroll
                    for (int r = 0; r < 3; r++)
#pragma unroll
                        for (int c = 0; c < 6; c++) pan[(r0 + r - j0) * PB + c] = acc[r][c];
                }
            }
            // row 72 of the panel: the linear term m (kept in shared memory, updated below by the
            // same threads 0..5 that copy it here)
            if (tid < PB) pan[(NZ - j0) * PB + tid] = sv[SV_QT + j0 + tid];
            __syncthreads();
            TMARK(8);
            // panel factorisation: thread t handles row j0 + t; the 6x6 diagonal block is factored
            // redundantly in registers by every row thread (no extra barrier).  Non-positive pivots
            // zero the column (BLASFEO potrf rule).  The diagonal-block rows are written back after
            // the barrier below: nobody reads them during the trailing update.
            const int nrow = NZ + 1 - j0;
            double drow[PB] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
            if (tid < nrow) {
                double Dm[PB][PB];       // strictly-lower entries of the block factor; Dm[a][a] = 1/L[a][a]
#pragma unroll
                for (int aa = 0; aa < PB; aa++) {
#pragma unroll
                    for (int c = 0; c <= aa; c++) {
                        double v = pan[aa * PB + c];
#pragma unroll
                        for (int e = 0; e < c; e++) v -= Dm[aa][e] * Dm[c][e];
                        if (c == aa) {
                            // one rsqrt per pivot (the serial chain of the factorisation): 1/L_aa, L_aa = v/L_aa
                            double d = 0.0, inv = 0.0;
                            if (v > 0.0) { inv = rsqrt(v); d = v * inv; }
                            Dm[aa][aa] = inv;
                            if (tid == aa) drow[c] = d;
                        } else {
                            const double lv = v * Dm[c][c];
                            Dm[aa][c] = lv;
                            if (tid == aa) drow[c] = lv;
                        }
                    }
                    if (tid == aa) {
#pragma unroll
                        for (int c = aa + 1; c < PB; c++) drow[c] = 0.0;
                    }
                }
                TMARK(14);
                if (tid >= PB) {
                    double rowv[PB];
                    load6(pan + tid * PB, rowv);
#pragma unroll
                    for (int jj = 0; jj < PB; jj++) {
                        double v = rowv[jj];
#pragma unroll
                        for (int e = 0; e < jj; e++) v -= rowv[e] * Dm[jj][e];
                        rowv[jj] = v * Dm[jj][jj];
                    }
#pragma unroll
                    for (int jj = 0; jj < PB; jj++) pan[tid * PB + jj] = rowv[jj];
                }
            }
            TMARK(9);
            __syncthreads();
            TMARK(10);
            if (tid < PB) {
#pragma unroll
                for (int c = 0; c < PB; c++) pan[tid * PB + c] = drow[c];
            }
            // trailing update of the register tiles right of this panel
            if (is_tile) {
                if (cb > jb) {
                    const double* pa = pan + (r0 - j0) * PB;
                    const double* pb = pan + (c0 - j0) * PB;
#pragma unroll
                    for (int kk = 0; kk < PB; kk++) {
                        const double a0 = pa[kk], a1 = pa[PB + kk], a2 = pa[2 * PB + kk];
#pragma unroll
                        for (int c = 0; c < 6; c++) {
                            const double bv = pb[c * PB + kk];
                            acc[0][c] -= a0 * bv; acc[1][c] -= a1 * bv; acc[2][c] -= a2 * bv;
                        }
                    }
                }
            }
            // trailing update of row 72: m_c -= l . L(c, j0..j0+5), thread t owns column j0 + 6 + t
            if (tid < NZ - j0 - PB) {
                const double* pa = pan + (NZ - j0) * PB;
                const double* pb = pan + (PB + tid) * PB;
                double av[6], bv[6];
                load6(pa, av);
                load6(pb, bv);
                double sm_ = sv[SV_QT + j0 + PB + tid];
#pragma unroll
                for (int kk = 0; kk < PB; kk++) sm_ -= av[kk] * bv[kk];
                sv[SV_QT + j0 + PB + tid] = sm_;
            }
            TMARK(11);
        }
        // ---- factor -> HBM (one bulk store); p_k = Lxx lx stays in shared memory for stage k-1 ----
        fence_proxy_async();
        __syncthreads();
        if (tid == 0) bulk_s2g(Lb + (size_t)k * LP_SZ, sLp, LP_SZ * 8);
        if (tid < NX) {
            double s = 0.0;
            for (int l = 0; l <= tid; l++) s += sLp[lp_idx(NU + tid, NU + l)] * sLp[lp_idx(NZ, NU + l)];
            sv[SV_PN + tid] = s;
        }
        __syncthreads();
        TMARK(12);
    }

    // ================= forward sweep (affine direction; multipliers dpi not needed) =================
    if (tid == 0) bulk_store_wait_all();
    if (tid < NX) sv[SV_DX + tid] = 0.0;
    __syncthreads();
    for (int k = 0; k <= N; k++) {
        const bool term = (k == N);
        double* vk = vb + (size_t)k * V_STAGE;
        if (!term) {
            if (k > 0 && tid == 0) {          // stage 0's factor is still in sLp
                mbar_arrive_expect_tx(&bar, LP_UPANELS * 8);
                bulk_g2s(sLp, Lb + (size_t)k * LP_SZ, LP_UPANELS * 8, &bar);
            }
            for (int i = tid; i < NX; i += QP_THREADS) sv[SV_RB + i] = vk[V_RB + i];
            __syncthreads();
            if (k > 0) { mbar_wait(&bar, phase); phase ^= 1; }
        }
        forward_stage<false>(sLp, sv, term ? nullptr : Gb + (size_t)k * G_SZ, vk, nullptr, term, tid);
    }
    TMARK(13);
#ifdef NMPC_TIMING
    if (bl == 0 && tid == 0) {
#pragma unroll
        for (int i = 0; i < 16; i++) g_phase_cycles[i] += t_acc[i];
    }
#endif
}

// =============================================================================================
// K4 / K8: slack and multiplier directions from dz, step-length candidates.  One CTA per stage.
// =============================================================================================
__global__ void __launch_bounds__(128) nmpc_qp_step_kernel(QpArgs a)
{
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int k = blockIdx.x, bl = blockIdx.y, b = a.b0 + bl, N = a.N;
    if (a.qp_done[b]) return;
    __shared__ double s_dz[NZ], s_cz[NG], s_red[2][4];
    const double* Cp = a.C + ((size_t)bl * (N + 1) + k) * C_SZ;
    const double* Dp = a.D + ((size_t)bl * (N + 1) + k) * D_SZ;
    double* vk = a.vec + ((size_t)bl * (N + 1) + k) * V_STAGE;
    const double* qlam = a.qlam + ((size_t)b * (N + 1) + k) * 2 * NG;
    const double* qt = a.qt + ((size_t)b * (N + 1) + k) * 2 * NG;
    for (int i = tid; i < NZ; i += 128) s_dz[i] = vk[V_DZ + i];
    __syncthreads();
    for (int j = warp; j < NG; j += 4) {
        double s = 0.0;
        for (int c = lane; c < NZ; c += 32) s += Cp[j * C_LD + c] * s_dz[c];
        s = warp_sum(s);
        if (lane == 0) s_cz[j] = s;
    }
    __syncthreads();
    double a_p = 1.0, a_d = 1.0;
    if (tid < 2 * NG) {
        const int e = tid / NG, j = tid % NG;
        const double msk = Dp[2 * NG + tid];
        double dtv = 0.0, dl = 0.0;
        if (msk > 0.5) {
            const double lam = qlam[tid], t = qt[tid];
            const double cdz = s_cz[j];
            dtv = (e ? -cdz : cdz) + vk[V_RD + tid];
            dl = -(vk[V_RM + tid] + lam * dtv) / t;
            if (dtv < 0.0) a_p = fmin(a_p, -t / dtv);
            if (dl < 0.0) a_d = fmin(a_d, -lam / dl);
        }
        vk[V_DT + tid] = dtv; vk[V_DLAM + tid] = dl;
    }
    a_p = warp_min(a_p); a_d = warp_min(a_d);
    if (lane == 0) { s_red[0][warp] = a_p; s_red[1][warp] = a_d; }
    __syncthreads();
    if (tid < 2) {
        const double v = fmin(fmin(s_red[tid][0], s_red[tid][1]), fmin(s_red[tid][2], s_red[tid][3]));
        // candidates are in (0, 1]: positive doubles order like their bit patterns
        if (v < 1.0) atomicMin(a.alpha + 2 * b + tid, (unsigned long long)__double_as_longlong(v));
    }
}

// =============================================================================================
// K5: centring parameter.  One CTA per problem: mu_aff with a fixed reduction tree.
// =============================================================================================
__global__ void __launch_bounds__(QP_THREADS) nmpc_qp_sigma_kernel(QpArgs a)
{
    const int tid = threadIdx.x, bl = blockIdx.x, b = a.b0 + bl, N = a.N;
    if (a.qp_done[b]) return;
    __shared__ double red[8];
    const double a_p = __longlong_as_double((long long)a.alpha[2 * b]);
    const double a_d = __longlong_as_double((long long)a.alpha[2 * b + 1]);
    const double alpha_aff = fmin(a_p, a_d);
    const double* vb = a.vec + (size_t)bl * (N + 1) * V_STAGE;
    const double* Db = a.D + (size_t)bl * (N + 1) * D_SZ;
    const double* qlam = a.qlam + (size_t)b * (N + 1) * 2 * NG;
    const double* qt = a.qt + (size_t)b * (N + 1) * 2 * NG;
    double s = 0.0;
    for (int i = tid; i < (N + 1) * 2 * NG; i += QP_THREADS) {
        const int k = i / (2 * NG), id = i % (2 * NG);
        if (Db[(size_t)k * D_SZ + 2 * NG + id] > 0.5) {
            const double* vk = vb + (size_t)k * V_STAGE;
            s += (qlam[i] + alpha_aff * vk[V_DLAM + id]) * (qt[i] + alpha_aff * vk[V_DT + id]);
        }
    }
    s = block_sum(s, red, tid);       // also orders the reads of alpha before the reset below
    if (tid == 0) {
        const int nc = a.nc[b];
        const double mu = a.mu[b];
        const double mu_aff = nc > 0 ? s / nc : 0.0;
        double sg = mu > 0.0 ? mu_aff / mu : 0.0;
        sg = sg * sg * sg;
        a.sigma_mu[b] = sg * mu;
        a.alpha[2 * b] = ONE_BITS; a.alpha[2 * b + 1] = ONE_BITS;      // reset for the corrector step
    }
}

// =============================================================================================
// K6: corrector right-hand side  rm <- lam t + dt_aff dlam_aff - sigma mu,  q tilde = rg + C' gamma
// =============================================================================================
__global__ void __launch_bounds__(96) nmpc_qp_corr_rhs_kernel(QpArgs a)
{
    const int tid = threadIdx.x;
    const int k = blockIdx.x, bl = blockIdx.y, b = a.b0 + bl, N = a.N;
    if (a.qp_done[b]) return;
    __shared__ double s_g[2 * NG], s_gam[NG];
    const double* Cp = a.C + ((size_t)bl * (N + 1) + k) * C_SZ;
    const double* Dp = a.D + ((size_t)bl * (N + 1) + k) * D_SZ;
    double* vk = a.vec + ((size_t)bl * (N + 1) + k) * V_STAGE;
    const double* qlam = a.qlam + ((size_t)b * (N + 1) + k) * 2 * NG;
    const double* qt = a.qt + ((size_t)b * (N + 1) + k) * 2 * NG;
    const double sigma_mu = a.sigma_mu[b];
    if (tid < 2 * NG) {
        double v = 0.0;
        if (Dp[2 * NG + tid] > 0.5) {
            const double lam = qlam[tid], t = qt[tid];
            const double rmc = lam * t + vk[V_DT + tid] * vk[V_DLAM + tid] - sigma_mu;
            vk[V_RM + tid] = rmc;
            v = (rmc + lam * vk[V_RD + tid]) / t;
        }
        s_g[tid] = v;
    }
    __syncthreads();
    if (tid < NG) s_gam[tid] = s_g[tid] - s_g[NG + tid];
    __syncthreads();
    if (tid < NZ) {
        double s = vk[V_RG + tid];
        for (int j = 0; j < NG; j++) s += Cp[j * C_LD + tid] * s_gam[j];
        vk[V_QT + tid] = s;
    }
}

// =============================================================================================
// K7: corrector solve on the stored factorisation: backward vector sweep + forward sweep with
//     the multiplier directions.  One CTA per OCP, ~37 KB of shared memory -> 5 CTAs per SM.
// =============================================================================================
__global__ void __launch_bounds__(QP_THREADS, 5) nmpc_qp_corr_solve_kernel(QpArgs a)
{
    extern __shared__ __align__(16) double smem[];
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x;
    const int bl = blockIdx.x, b = a.b0 + bl, N = a.N;
    if (a.qp_done[b]) return;
    double* sLp = smem + S_LP;
    double* sLx = smem + S_U;              // Lxx panels of stage k+1 (LX_SZ)
    double* sv = smem + S_U + LX_SZ;

    const double* Gb = a.G + (size_t)bl * N * G_SZ;
    const double* Lb = a.L + (size_t)bl * (N + 1) * LP_SZ;
    double* vb = a.vec + (size_t)bl * (N + 1) * V_STAGE;

    uint32_t phase = 0;
    if (tid == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    __syncthreads();

    for (int k = N; k >= 0; k--) {
        const bool term = (k == N);
        double* vk = vb + (size_t)k * V_STAGE;
        const double* Gp = term ? nullptr : Gb + (size_t)k * G_SZ;
        if (tid == 0) {
            mbar_arrive_expect_tx(&bar, LP_SZ * 8);
            bulk_g2s(sLp, Lb + (size_t)k * LP_SZ, LP_SZ * 8, &bar);
        }
        for (int i = tid; i < NZ; i += QP_THREADS) sv[SV_QT + i] = vk[V_QT + i];
        if (!term) for (int i = tid; i < NX; i += QP_THREADS) sv[SV_RB + i] = vk[V_RB + i];
        __syncthreads();
        if (!term) {
            value_fn_times_rb([&](int l, int i) { return sLx[lp_idx(NU + l, NU + i) - LP_UPANELS]; }, sv, tid);
            if (tid < NZ) {       // m = q tilde + G' w
                double s = sv[SV_QT + tid];
                for (int l = 0; l < NX; l++) s += Gp[l * G_LD + tid] * sv[SV_W + l];
                sv[SV_QT + tid] = s;
            }
        }
        mbar_wait(&bar, phase); phase ^= 1;
        if (tid < NZ) { const double d = sLp[lp_idx(tid, tid)]; sv[SV_INV + tid] = d != 0.0 ? 1.0 / d : 0.0; }
        __syncthreads();
        // l = L_k^-1 m : warp 0, column-oriented forward substitution, rows lane, lane+32, lane+64
        if (tid < 32) {
            const int o = term ? NU : 0;
            double m0 = sv[SV_QT + tid], m1 = sv[SV_QT + 32 + tid];
            double m2 = (tid + 64 < NZ) ? sv[SV_QT + 64 + tid] : 0.0;
            for (int j = o; j < NZ; j++) {
                const int slot = j >> 5, ln = j & 31;
                const double mine = slot == 0 ? m0 : (slot == 1 ? m1 : m2);
                const double xj = __shfl_sync(0xffffffffu, mine, ln) * sv[SV_INV + j];
                if (tid == ln) { if (slot == 0) m0 = xj; else if (slot == 1) m1 = xj; else m2 = xj; }
                if (tid > j) m0 -= sLp[lp_idx(tid, j)] * xj;
                if (tid + 32 > j) m1 -= sLp[lp_idx(tid + 32, j)] * xj;
                if (tid + 64 > j && tid + 64 < NZ) m2 -= sLp[lp_idx(tid + 64, j)] * xj;
            }
            sLp[lp_idx(NZ, tid)] = m0; vk[V_LROW + tid] = m0;
            sLp[lp_idx(NZ, tid + 32)] = m1; vk[V_LROW + tid + 32] = m1;
            if (tid + 64 < NZ) { sLp[lp_idx(NZ, tid + 64)] = m2; vk[V_LROW + tid + 64] = m2; }
        }
        __syncthreads();
        for (int i = tid; i < LX_SZ; i += QP_THREADS) sLx[i] = sLp[LP_UPANELS + i];
        if (tid < NX) {
            double s = 0.0;
            for (int l = 0; l <= tid; l++) s += sLp[lp_idx(NU + tid, NU + l)] * sLp[lp_idx(NZ, NU + l)];
            sv[SV_PN + tid] = s; vk[V_P + tid] = s;
        }
        __syncthreads();
    }

    // ================= forward sweep with multiplier directions =================
    if (tid < NX) sv[SV_DX + tid] = 0.0;
    __syncthreads();
    for (int k = 0; k <= N; k++) {
        const bool term = (k == N);
        double* vk = vb + (size_t)k * V_STAGE;
        if (k > 0 && tid == 0) {
            mbar_arrive_expect_tx(&bar, LP_SZ * 8);
            bulk_g2s(sLp, Lb + (size_t)k * LP_SZ, LP_SZ * 8, &bar);
        }
        for (int i = tid; i < NX; i += QP_THREADS) {
            sv[SV_RB + i] = term ? 0.0 : vk[V_RB + i];
            sv[SV_PN + i] = vk[V_P + i];
        }
        if (k > 0) {
            mbar_wait(&bar, phase); phase ^= 1;
            // the stored factor carries the affine l row: patch in the corrector's
            for (int i = tid; i < NZ; i += QP_THREADS) sLp[lp_idx(NZ, i)] = vk[V_LROW + i];
        }
        __syncthreads();
        forward_stage<true>(sLp, sv, term ? nullptr : Gb + (size_t)k * G_SZ, vk,
                            k > 0 ? vb + (size_t)(k - 1) * V_STAGE + V_DPI : nullptr, term, tid);
    }
}

// =============================================================================================
// K9: apply the step (split primal / dual step lengths), or stop on a vanishing step.
// =============================================================================================
__global__ void __launch_bounds__(128) nmpc_qp_update_kernel(QpArgs a)
{
    const int tid = threadIdx.x;
    const int k = blockIdx.x, bl = blockIdx.y, b = a.b0 + bl, N = a.N;
    if (a.qp_done[b]) return;
    double alpha_p = __longlong_as_double((long long)a.alpha[2 * b]);
    double alpha_d = __longlong_as_double((long long)a.alpha[2 * b + 1]);
    alpha_p *= (1.0 - alpha_p) * 0.99 + alpha_p * 0.9999999;
    alpha_d *= (1.0 - alpha_d) * 0.99 + alpha_d * 0.9999999;
    if (alpha_p < IPM_ALPHA_MIN && alpha_d < IPM_ALPHA_MIN) {
        // every CTA of this problem takes the same decision; stage 0 records it
        if (k == 0 && tid == 0) { a.qp_status[b] = 2; a.qp_done[b] = 1; }
        return;
    }
    const double* vk = a.vec + ((size_t)bl * (N + 1) + k) * V_STAGE;
    const double* Dp = a.D + ((size_t)bl * (N + 1) + k) * D_SZ;
    double* qz = a.qz + ((size_t)b * (N + 1) + k) * NZ;
    double* qlam = a.qlam + ((size_t)b * (N + 1) + k) * 2 * NG;
    double* qt = a.qt + ((size_t)b * (N + 1) + k) * 2 * NG;
    if (tid < NZ) qz[tid] += alpha_p * vk[V_DZ + tid];
    if (k < N && tid < NX) a.qpi[((size_t)b * N + k) * NX + tid] += alpha_d * vk[V_DPI + tid];
    if (tid < 2 * NG && Dp[2 * NG + tid] > 0.5) {
        qt[tid] = fmax(qt[tid] + alpha_p * vk[V_DT + tid], IPM_T_MIN);
        qlam[tid] = fmax(qlam[tid] + alpha_d * vk[V_DLAM + tid], IPM_LAM_MIN);
    }
}

// =============================================================================================
// end of the QP: SQP full step (x, u) += dz, multipliers replaced by the QP's
// =============================================================================================
__global__ void __launch_bounds__(128) nmpc_qp_apply_kernel(QpArgs a)
{
    const int tid = threadIdx.x, bl = blockIdx.x, b = a.b0 + bl, N = a.N;
    if (a.sqp_done && a.sqp_done[b]) return;
    if (tid == 0 && a.sqp_iter) a.sqp_iter[b] += 1;
    const int status = a.qp_status[b];
    if (!(status == 0 || status == 1)) return;
    const double* qz = a.qz + (size_t)b * (N + 1) * NZ;
    const double* qpi = a.qpi + (size_t)b * N * NX;
    const double* qlam = a.qlam + (size_t)b * (N + 1) * 2 * NG;
    const double* Db = a.D + (size_t)bl * (N + 1) * D_SZ;
    double* x = a.x + (size_t)b * (N + 1) * NX;
    double* u = a.u + (size_t)b * N * NU;
    double* pi = a.pi + (size_t)b * N * NX;
    double* lam = a.lam + (size_t)b * (N + 1) * 2 * NG;
    for (int i = tid; i < (N + 1) * NX; i += 128) x[i] += qz[(size_t)(i / NX) * NZ + NU + i % NX];
    for (int i = tid; i < N * NU; i += 128) u[i] += qz[(size_t)(i / NU) * NZ + i % NU];
    for (int i = tid; i < N * NX; i += 128) pi[i] = qpi[i];
    for (int i = tid; i < (N + 1) * 2 * NG; i += 128) {
        const int k = i / (2 * NG), id = i % (2 * NG);
        lam[i] = Db[(size_t)k * D_SZ + 2 * NG + id] > 0.5 ? qlam[i] : 0.0;
    }
}
