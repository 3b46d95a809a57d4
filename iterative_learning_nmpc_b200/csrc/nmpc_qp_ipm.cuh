// nmpc_qp_ipm.cuh -- OCP-QP interior-point kernel: one CTA per OCP, fp64, all IPM iterations
// and all Riccati sweeps inside one launch.
//
// Replaces HPIPM's d_ocp_qp_ipm (mode `speed`) + BLASFEO as used by acados for the reference's
// solver: Mehrotra predictor-corrector on the OCP-QP of one SQP
// step, backward square-root Riccati factorisation over the stage blocks, forward sweep, a second
// (vector-only) sweep pair for the corrector on the same factorisation, split primal/dual step.
// The algorithm is the one restated in oracle/nmpc_oracle.c::qp_ipm (the checker); the exit test
// is evaluated after the backward sweep that produced the residuals (same decision, the sweep's
// factor is simply discarded on exit).
//
// B200 mapping: stage blocks G (24.9 KB), M (42.6 KB), C (24.2 KB) are staged into shared memory
// with 1-D bulk TMA copies (cp.async.bulk -> UBLKCP) completing on an mbarrier; factors go back
// to HBM with bulk stores.  L_{k+1}^{xx} stays resident in shared memory between stages.
// 91 KB of shared memory per CTA -> 2 CTAs per SM, 256 threads each; fp64 FMA pipe work is
// register-tiled 4x4 (rank-k updates) / 3x4 (triangular product).
#pragma once
#include "nmpc_common.cuh"

// per-stage slots of the per-problem vector workspace (doubles)
#define V_RG 0
#define V_RB (V_RG + NZ)
#define V_RD (V_RB + NX)
#define V_RM (V_RD + 2 * NG)
#define V_DZ (V_RM + 2 * NG)
#define V_DPI (V_DZ + NZ)
#define V_DT (V_DPI + NX)
#define V_DLAM (V_DT + 2 * NG)
#define V_P (V_DLAM + 2 * NG)
#define V_STAGE (V_P + NX)          // 606

struct QpArgs {
    const double* G; const double* M; const double* C; const double* D;   // chunk-local blocks
    double* L;                      // [Bc][N+1][L_SZ]
    double* vec;                    // [Bc][(N+1)*V_STAGE]
    double* qz; double* qpi; double* qlam; double* qt;   // stored QP solution, global index
    const double* dx0_ext;          // [B][NX] or null -> x0 - x[0]
    const double* x0;
    double* x; double* u; double* pi; double* lam;       // SQP iterate (updated when apply)
    int32_t* qp_iter; int32_t* qp_status; double* qp_res;
    const int32_t* done; int32_t* sqp_iter;
    int N, b0, max_iter, warm, apply;
    double tol;
};

// shared-memory carve-up (doubles)
#define S_L 0
#define S_U (S_L + NX * NX)             // union: C (phase 1) / G (phase 2)
#define S_M (S_U + G_SZ)
#define S_V (S_M + M_SZ)
//   vectors
#define SV_Z 0
#define SV_ZXN (SV_Z + NZ)
#define SV_PIK (SV_ZXN + NX)
#define SV_PIN (SV_PIK + NX)
#define SV_LAM (SV_PIN + NX)
#define SV_T (SV_LAM + 2 * NG)
#define SV_D (SV_T + 2 * NG)            // lg | ug | ml | mu
#define SV_GAM (SV_D + 4 * NG)          // Gamma
#define SV_GAMV (SV_GAM + NG)           // gamma
#define SV_DL (SV_GAMV + NG)            // masked lam_l - lam_u
#define SV_CZ (SV_DL + NG)
#define SV_RG (SV_CZ + NG)
#define SV_RB (SV_RG + NZ)
#define SV_QT (SV_RB + NX)              // q tilde / m  (NZ+1)
#define SV_W (SV_QT + NZ + 2)
#define SV_PN (SV_W + NX)               // p_{k+1}
#define SV_RD (SV_PN + NX)
#define SV_RM (SV_RD + 2 * NG)
#define SV_TMP (SV_RM + 2 * NG)         // NZ scratch
#define SV_DX (SV_TMP + NZ)             // forward sweep state
#define SV_DXN (SV_DX + NX)
#define SV_BLK (SV_DXN + NX)            // 8x8 diagonal block
#define SV_RED (SV_BLK + 64)            // 32 reduction slots
#define SV_END (SV_RED + 32)
#define QP_SMEM_BYTES ((S_V + SV_END) * 8 + 16)

__device__ __forceinline__ double block_reduce(double v, int op, double* red, int tid)
{
    // op: 0 max, 1 min, 2 sum.  Result broadcast to all threads.
    if (op == 0) v = warp_max(v); else if (op == 1) v = warp_min(v); else v = warp_sum(v);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    double r = red[0];
#pragma unroll
    for (int w = 1; w < QP_THREADS / 32; w++) {
        const double o = red[w];
        if (op == 0) r = nanmax(r, o); else if (op == 1) r = fmin(r, o); else r += o;
    }
    return r;
}

// s = sum_c A[r*ld + c] * x[c] for row r = tid/4 (4 threads per row, quad-reduced).  Must be
// called by ALL threads of the CTA (full-warp shuffles); rows >= nrows return 0.
__device__ __forceinline__ double rowdot4(const double* A, int ld, const double* x, int nrows, int ncols, int tid)
{
    const int r = tid >> 2, q = tid & 3;
    double s = 0.0;
    if (r < nrows) {
        const double* row = A + r * ld;
        for (int c = q; c < ncols; c += 4) s += row[c] * x[c];
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    return s;
}

// Blocked left-looking Cholesky of sM[o..72) x [o..72) (ld 73) plus the appended row 72
// (l = L^-1 m).  One thread per row; PB x PB diagonal blocks (PB = 6 divides both 72 and 42) are
// factored redundantly in registers, so a block column costs two barriers.  Non-positive pivots
// zero the column (BLASFEO potrf rule).
#define PB 6
__device__ __forceinline__ void potrf_block(double* sM, int o, double* blk, int tid)
{
    for (int j0 = o; j0 < NZ; j0 += PB) {
        const int i = j0 + tid;
        const bool active = (i <= NZ);
        double acc[PB];
        if (active) {
#pragma unroll
            for (int jj = 0; jj < PB; jj++) acc[jj] = sM[i * M_LD + j0 + jj];
            for (int c = o; c < j0; c++) {
                const double li = sM[i * M_LD + c];
#pragma unroll
                for (int jj = 0; jj < PB; jj++) acc[jj] -= li * sM[(j0 + jj) * M_LD + c];
            }
            if (tid < PB) {
#pragma unroll
                for (int jj = 0; jj < PB; jj++) blk[tid * PB + jj] = acc[jj];
            }
        }
        __syncthreads();
        if (active) {
            double Dm[PB][PB];   // strictly-lower entries of the block factor; Dm[a][a] = 1/L[a][a]
#pragma unroll
            for (int a = 0; a < PB; a++) {
#pragma unroll
                for (int c = 0; c <= a; c++) {
                    double v = blk[a * PB + c];
#pragma unroll
                    for (int e = 0; e < c; e++) v -= Dm[a][e] * Dm[c][e];
                    if (c == a) {
                        double d = 0.0, inv = 0.0;
                        if (v > 0.0) { d = sqrt(v); inv = 1.0 / d; }
                        Dm[a][a] = inv;
                        if (tid == a) sM[i * M_LD + j0 + a] = d;
                    } else {
                        const double lv = v * Dm[c][c];
                        Dm[a][c] = lv;
                        if (tid == a) sM[i * M_LD + j0 + c] = lv;
                    }
                }
                if (tid == a) {
#pragma unroll
                    for (int c = a + 1; c < PB; c++) sM[i * M_LD + j0 + c] = 0.0;
                }
            }
            if (tid >= PB) {
#pragma unroll
                for (int jj = 0; jj < PB; jj++) {
                    double v = acc[jj];
#pragma unroll
                    for (int e = 0; e < jj; e++) v -= acc[e] * Dm[jj][e];
                    acc[jj] = v * Dm[jj][jj];
                    sM[i * M_LD + j0 + jj] = acc[jj];
                }
            }
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(QP_THREADS, 2) nmpc_qp_ipm_kernel(QpArgs a)
{
    extern __shared__ __align__(16) double smem[];
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x;
    const int bl = blockIdx.x, b = a.b0 + bl;
    if (a.done && a.done[b]) return;
    const int N = a.N;

    double* sL = smem + S_L;
    double* sU = smem + S_U;
    double* sM = smem + S_M;
    double* sv = smem + S_V;
    double* red = sv + SV_RED;

    const double* Gb = a.G + (size_t)bl * N * G_SZ;
    const double* Mb = a.M + (size_t)bl * (N + 1) * M_SZ;
    const double* Cb = a.C + (size_t)bl * (N + 1) * C_SZ;
    const double* Db = a.D + (size_t)bl * (N + 1) * D_SZ;
    double* Lb = a.L + (size_t)bl * (N + 1) * L_SZ;
    double* vb = a.vec + (size_t)bl * (N + 1) * V_STAGE;
    double* qz = a.qz + (size_t)b * (N + 1) * NZ;
    double* qpi = a.qpi + (size_t)b * N * NX;
    double* qlam = a.qlam + (size_t)b * (N + 1) * 2 * NG;
    double* qt = a.qt + (size_t)b * (N + 1) * 2 * NG;

    uint32_t phase = 0;
    if (tid == 0) { mbar_init(&bar, 1); fence_barrier_init(); }

    // syrk tile of this thread: 4x4 tile (ti, tj), tj <= ti < 18, 171 tiles
    int ti = 0, tj = 0;
    const bool has_tile = tid < 171;
    if (has_tile) { int rem = tid; while (rem > ti) { rem -= ti + 1; ti++; } tj = rem; }

    // ---- initial point ----
    if (a.warm == 0) {
        for (int i = tid; i < (N + 1) * NZ; i += QP_THREADS) qz[i] = 0.0;
        for (int i = tid; i < N * NX; i += QP_THREADS) qpi[i] = 0.0;
    }
    __syncthreads();
    if (tid < NX) {
        const double d0 = a.dx0_ext ? a.dx0_ext[(size_t)b * NX + tid]
                                    : a.x0[(size_t)b * NX + tid] - a.x[(size_t)b * (N + 1) * NX + tid];
        qz[NU + tid] = d0;
    }
    if (tid < NU) qz[(size_t)N * NZ + tid] = 0.0;
    __syncthreads();

    int status = 1, it = 0;
    double nc = 0.0, res_g = 0, res_b = 0, res_d = 0, res_m = 0, mu = 0;

    for (;; it++) {
        const bool last = (it >= a.max_iter);
        double acc_mu = 0.0, acc_nc = 0.0, m_g = 0.0, m_b = 0.0, m_d = 0.0, m_m = 0.0;

        // =============================== PASS A: backward ===============================
        for (int k = N; k >= 0; k--) {
            const bool term = (k == N);
            double* vk = vb + (size_t)k * V_STAGE;
            // stage loads: C -> sU, M -> sM (bulk), vectors by threads
            if (tid == 0) {
                mbar_arrive_expect_tx(&bar, (C_SZ + M_SZ) * 8);
                bulk_g2s(sU, Cb + (size_t)k * C_SZ, C_SZ * 8, &bar);
                bulk_g2s(sM, Mb + (size_t)k * M_SZ, M_SZ * 8, &bar);
            }
            for (int i = tid; i < NZ; i += QP_THREADS) sv[SV_Z + i] = qz[(size_t)k * NZ + i];
            for (int i = tid; i < NX; i += QP_THREADS) {
                sv[SV_ZXN + i] = term ? 0.0 : qz[(size_t)(k + 1) * NZ + NU + i];
                sv[SV_PIK + i] = (k > 0) ? qpi[(size_t)(k - 1) * NX + i] : 0.0;
                sv[SV_PIN + i] = term ? 0.0 : qpi[(size_t)k * NX + i];
            }
            for (int i = tid; i < 4 * NG; i += QP_THREADS) sv[SV_D + i] = Db[(size_t)k * D_SZ + i];
            if (!(it == 0 && a.warm < 2)) {
                for (int i = tid; i < 2 * NG; i += QP_THREADS) {
                    sv[SV_LAM + i] = qlam[(size_t)k * 2 * NG + i];
                    sv[SV_T + i] = qt[(size_t)k * 2 * NG + i];
                }
            }
            __syncthreads();
            mbar_wait(&bar, phase); phase ^= 1;

            // ---- phase 1: constraint part ----
            {
                const double s = rowdot4(sU, C_LD, sv + SV_Z, NG, NZ, tid);
                if ((tid & 3) == 0 && (tid >> 2) < NG) sv[SV_CZ + (tid >> 2)] = s;
            }
            __syncthreads();
            if (tid < 2 * NG) {
                const int e = tid / NG, j = tid % NG;
                const double msk = sv[SV_D + 2 * NG + tid];
                const double cz = sv[SV_CZ + j];
                const double slack = e ? (sv[SV_D + NG + j] - cz) : (cz - sv[SV_D + j]);
                double t, lam;
                if (it == 0 && (a.warm < 2 || !(sv[SV_T + tid] > 0.0) || !(sv[SV_LAM + tid] > 0.0))) {
                    if (msk > 0.5) { t = fmax(slack, IPM_THR0); lam = IPM_MU0 / t; } else { t = 1.0; lam = 0.0; }
                    sv[SV_T + tid] = t; sv[SV_LAM + tid] = lam;
                    qt[(size_t)k * 2 * NG + tid] = t; qlam[(size_t)k * 2 * NG + tid] = lam;
                } else { t = sv[SV_T + tid]; lam = sv[SV_LAM + tid]; }
                const double rd = msk > 0.5 ? (slack - t) : 0.0;
                const double rm = msk > 0.5 ? lam * t : 0.0;
                sv[SV_RD + tid] = rd; sv[SV_RM + tid] = rm;
                vk[V_RD + tid] = rd; vk[V_RM + tid] = rm;
                acc_mu += rm; acc_nc += (msk > 0.5) ? 1.0 : 0.0;
                m_d = nanmax(m_d, fabs(rd)); m_m = nanmax(m_m, fabs(rm));
            }
            __syncthreads();
            if (tid < NG) {
                const int j = tid;
                const double ml = sv[SV_D + 2 * NG + j], mu_ = sv[SV_D + 3 * NG + j];
                const double ll = sv[SV_LAM + j], lu = sv[SV_LAM + NG + j];
                const double tl = sv[SV_T + j], tu = sv[SV_T + NG + j];
                double Gam = 0.0, gam = 0.0, dl = 0.0;
                if (ml > 0.5) { Gam += ll / tl; gam += (sv[SV_RM + j] + ll * sv[SV_RD + j]) / tl; dl += ll; }
                if (mu_ > 0.5) { Gam += lu / tu; gam -= (sv[SV_RM + NG + j] + lu * sv[SV_RD + NG + j]) / tu; dl -= lu; }
                sv[SV_GAM + j] = Gam; sv[SV_GAMV + j] = gam; sv[SV_DL + j] = dl;
            }
            __syncthreads();
            // rg (without G^T pi+) and C^T gamma, thread per column
            if (tid < NZ) {
                const int c = tid;
                double s = sM[NZ * M_LD + c];
                for (int d = 0; d <= c; d++) s += sM[c * M_LD + d] * sv[SV_Z + d];
                for (int d = c + 1; d < NZ; d++) s += sM[d * M_LD + c] * sv[SV_Z + d];
                if (c >= NU && k > 0) s -= sv[SV_PIK + c - NU];
                double qp = 0.0;
                for (int j = 0; j < NG; j++) {
                    const double cj = sU[j * C_LD + c];
                    s -= cj * sv[SV_DL + j];
                    qp += cj * sv[SV_GAMV + j];
                }
                sv[SV_RG + c] = s; sv[SV_QT + c] = qp;
            }
            // Hessian accumulators: H + C^T Gamma C
            double accm[4][4];
            if (has_tile && !last) {
#pragma unroll
                for (int r = 0; r < 4; r++)
#pragma unroll
                    for (int s = 0; s < 4; s++) accm[r][s] = sM[(4 * ti + r) * M_LD + 4 * tj + s];
                for (int j = 0; j < NG; j++) {
                    const double g = sv[SV_GAM + j];
                    if (g == 0.0) continue;
                    const double2 a0 = *reinterpret_cast<const double2*>(sU + j * C_LD + 4 * ti);
                    const double2 a1 = *reinterpret_cast<const double2*>(sU + j * C_LD + 4 * ti + 2);
                    const double2 b0 = *reinterpret_cast<const double2*>(sU + j * C_LD + 4 * tj);
                    const double2 b1 = *reinterpret_cast<const double2*>(sU + j * C_LD + 4 * tj + 2);
                    const double av[4] = {a0.x * g, a0.y * g, a1.x * g, a1.y * g};
                    const double bv[4] = {b0.x, b0.y, b1.x, b1.y};
#pragma unroll
                    for (int r = 0; r < 4; r++)
#pragma unroll
                        for (int s = 0; s < 4; s++) accm[r][s] += av[r] * bv[s];
                }
            }
            __syncthreads();   // all reads of C (sU) and of the gradient row done

            // ---- phase 2: dynamics part ----
            if (!term) {
                if (tid == 0) {
                    mbar_arrive_expect_tx(&bar, G_SZ * 8);
                    bulk_g2s(sU, Gb + (size_t)k * G_SZ, G_SZ * 8, &bar);
                }
                mbar_wait(&bar, phase); phase ^= 1;
                // rg += G^T pi+ ; rb = G [z;1] - zx+
                if (tid < NZ) {
                    double s = sv[SV_RG + tid];
                    for (int l = 0; l < NX; l++) s += sU[l * G_LD + tid] * sv[SV_PIN + l];
                    sv[SV_RG + tid] = s;
                }
                {
                    const double s = rowdot4(sU, G_LD, sv + SV_Z, NX, NZ, tid);
                    const int l = tid >> 2;
                    if ((tid & 3) == 0 && l < NX) {
                        const double rb = s + sU[l * G_LD + NZ] - sv[SV_ZXN + l];
                        sv[SV_RB + l] = rb; vk[V_RB + l] = rb;
                        m_b = nanmax(m_b, fabs(rb));
                    }
                }
            }
            __syncthreads();
            if (tid < NZ) {
                double s = sv[SV_RG + tid];
                if ((k == 0 && tid >= NU) || (term && tid < NU)) s = 0.0;
                sv[SV_RG + tid] = s; vk[V_RG + tid] = s;
                m_g = nanmax(m_g, fabs(s));
                sv[SV_QT + tid] += s;                 // q tilde = rg + C^T gamma
            }
            if (last) { __syncthreads(); continue; }   // residual-only pass

            if (!term) {
                // w = L+ (L+^T rb) + p+
                if (tid < NX) {
                    double s = 0.0;
                    for (int l = tid; l < NX; l++) s += sL[l * NX + tid] * sv[SV_RB + l];
                    sv[SV_TMP + tid] = s;
                }
                __syncthreads();
                if (tid < NX) {
                    double s = sv[SV_PN + tid];
                    for (int l = 0; l <= tid; l++) s += sL[tid * NX + l] * sv[SV_TMP + l];
                    sv[SV_W + tid] = s;
                }
                __syncthreads();
                // m = q tilde + G^T w  -> row 72 of sM
                if (tid < NZ) {
                    double s = sv[SV_QT + tid];
                    for (int l = 0; l < NX; l++) s += sU[l * G_LD + tid] * sv[SV_W + l];
                    sM[NZ * M_LD + tid] = s;
                }
                // AL = L+^T G, in place (3x4 register tiles, 14 x 18 threads)
                double al[3][4];
                const bool has_t = tid < 252;
                const int i0 = (tid / 18) * 3, c0 = (tid % 18) * 4;
                if (has_t) {
#pragma unroll
                    for (int r = 0; r < 3; r++)
#pragma unroll
                        for (int s = 0; s < 4; s++) al[r][s] = 0.0;
                    for (int l = i0; l < NX; l++) {
                        const double l0 = sL[l * NX + i0], l1 = sL[l * NX + i0 + 1], l2 = sL[l * NX + i0 + 2];
                        const double2 g0 = *reinterpret_cast<const double2*>(sU + l * G_LD + c0);
                        const double2 g1 = *reinterpret_cast<const double2*>(sU + l * G_LD + c0 + 2);
                        const double gv[4] = {g0.x, g0.y, g1.x, g1.y};
#pragma unroll
                        for (int s = 0; s < 4; s++) { al[0][s] += l0 * gv[s]; al[1][s] += l1 * gv[s]; al[2][s] += l2 * gv[s]; }
                    }
                }
                __syncthreads();
                if (has_t) {
#pragma unroll
                    for (int r = 0; r < 3; r++)
#pragma unroll
                        for (int s = 0; s < 4; s++) sU[(i0 + r) * G_LD + c0 + s] = al[r][s];
                }
                __syncthreads();
                if (has_tile) {
                    for (int i = 0; i < NX; i++) {
                        const double2 a0 = *reinterpret_cast<const double2*>(sU + i * G_LD + 4 * ti);
                        const double2 a1 = *reinterpret_cast<const double2*>(sU + i * G_LD + 4 * ti + 2);
                        const double2 b0 = *reinterpret_cast<const double2*>(sU + i * G_LD + 4 * tj);
                        const double2 b1 = *reinterpret_cast<const double2*>(sU + i * G_LD + 4 * tj + 2);
                        const double av[4] = {a0.x, a0.y, a1.x, a1.y};
                        const double bv[4] = {b0.x, b0.y, b1.x, b1.y};
#pragma unroll
                        for (int r = 0; r < 4; r++)
#pragma unroll
                            for (int s = 0; s < 4; s++) accm[r][s] += av[r] * bv[s];
                    }
                }
            } else {
                if (tid < NZ) sM[NZ * M_LD + tid] = sv[SV_QT + tid];
            }
            if (has_tile) {
#pragma unroll
                for (int r = 0; r < 4; r++)
#pragma unroll
                    for (int s = 0; s < 4; s++) {
                        const int rr = 4 * ti + r, cc = 4 * tj + s;
                        if (cc <= rr) sM[rr * M_LD + cc] = accm[r][s] + ((rr == cc) ? IPM_REG_PRIM : 0.0);
                    }
            }
            __syncthreads();
            // prefetch next stage's C while the factorisation runs? (sU is free from here)
            potrf_block(sM, term ? NU : 0, sv + SV_BLK, tid);

            // factor block -> HBM (bulk store), Lxx -> sL, p_k = Lxx lx
            fence_proxy_async();
            __syncthreads();
            if (tid == 0) bulk_s2g(Lb + (size_t)k * L_SZ, sM, L_SZ * 8);
            for (int e = tid; e < NX * NX; e += QP_THREADS) {
                const int i = e / NX, j = e % NX;
                sL[e] = (j <= i) ? sM[(NU + i) * M_LD + NU + j] : 0.0;
            }
            if (tid < NX) {
                double s = 0.0;
                for (int l = 0; l <= tid; l++) s += sM[(NU + tid) * M_LD + NU + l] * sM[NZ * M_LD + NU + l];
                sv[SV_PN + tid] = s; vk[V_P + tid] = s;
            }
            if (tid == 0) bulk_store_wait_read();
            __syncthreads();
        }
        // ---- residual norms, exit tests ----
        res_g = block_reduce(m_g, 0, red, tid);
        res_b = block_reduce(m_b, 0, red, tid);
        res_d = block_reduce(m_d, 0, red, tid);
        res_m = block_reduce(m_m, 0, red, tid);
        const double mu_sum = block_reduce(acc_mu, 2, red, tid);
        if (it == 0) nc = block_reduce(acc_nc, 2, red, tid);
        mu = nc > 0.0 ? mu_sum / nc : 0.0;
        if (!(res_g == res_g) || !(res_b == res_b) || !(res_d == res_d) || !(res_m == res_m)) { status = 3; break; }
        if (res_g < a.tol && res_b < a.tol && res_d < a.tol && res_m < a.tol) { status = 0; break; }
        if (last) { status = 1; break; }

        double sigma_mu = 0.0, alpha_p = 1.0, alpha_d = 1.0;
        // ======================= PASS B (affine) / C / D (corrector) =======================
        for (int corr = 0; corr < 2; corr++) {
            if (corr == 1) {
                // ---------------- PASS C: backward vector sweep with the corrector rhs ----------------
                for (int k = N; k >= 0; k--) {
                    const bool term = (k == N);
                    double* vk = vb + (size_t)k * V_STAGE;
                    if (tid == 0) {
                        mbar_arrive_expect_tx(&bar, (C_SZ + L_SZ) * 8);
                        bulk_g2s(sU, Cb + (size_t)k * C_SZ, C_SZ * 8, &bar);
                        bulk_g2s(sM, Lb + (size_t)k * L_SZ, L_SZ * 8, &bar);
                    }
                    if (tid < 2 * NG) {
                        const double msk = Db[(size_t)k * D_SZ + 2 * NG + tid];
                        const double lam = qlam[(size_t)k * 2 * NG + tid], t = qt[(size_t)k * 2 * NG + tid];
                        double v = 0.0;
                        if (msk > 0.5) {
                            const double rmc = lam * t + vk[V_DT + tid] * vk[V_DLAM + tid] - sigma_mu;
                            v = (rmc + lam * vk[V_RD + tid]) / t;
                        }
                        sv[SV_RM + tid] = v;     // per-side gamma
                    }
                    for (int i = tid; i < NX; i += QP_THREADS) sv[SV_RB + i] = term ? 0.0 : vk[V_RB + i];
                    __syncthreads();
                    if (tid < NG) sv[SV_GAMV + tid] = sv[SV_RM + tid] - sv[SV_RM + NG + tid];
                    __syncthreads();
                    mbar_wait(&bar, phase); phase ^= 1;
                    if (tid < NZ) {
                        double s = vk[V_RG + tid];
                        for (int j = 0; j < NG; j++) s += sU[j * C_LD + tid] * sv[SV_GAMV + j];
                        sv[SV_QT + tid] = s;
                    }
                    __syncthreads();
                    if (!term) {
                        if (tid == 0) {
                            mbar_arrive_expect_tx(&bar, G_SZ * 8);
                            bulk_g2s(sU, Gb + (size_t)k * G_SZ, G_SZ * 8, &bar);
                        }
                        if (tid < NX) {
                            double s = 0.0;
                            for (int l = tid; l < NX; l++) s += sL[l * NX + tid] * sv[SV_RB + l];
                            sv[SV_TMP + tid] = s;
                        }
                        __syncthreads();
                        if (tid < NX) {
                            double s = sv[SV_PN + tid];
                            for (int l = 0; l <= tid; l++) s += sL[tid * NX + l] * sv[SV_TMP + l];
                            sv[SV_W + tid] = s;
                        }
                        mbar_wait(&bar, phase); phase ^= 1;
                        __syncthreads();
                        if (tid < NZ) {
                            double s = sv[SV_QT + tid];
                            for (int l = 0; l < NX; l++) s += sU[l * G_LD + tid] * sv[SV_W + l];
                            sv[SV_QT + tid] = s;
                        }
                        __syncthreads();
                    }
                    // l = L_k^-1 m (warp 0, column-oriented forward substitution)
                    if (tid < 32) {
                        const int o = term ? NU : 0;
                        double m0 = sv[SV_QT + tid], m1 = sv[SV_QT + 32 + tid];
                        double m2 = (tid + 64 < NZ) ? sv[SV_QT + 64 + tid] : 0.0;
                        for (int j = o; j < NZ; j++) {
                            const double d = sM[j * M_LD + j];
                            const double inv = d != 0.0 ? 1.0 / d : 0.0;
                            const int slot = j >> 5, ln = j & 31;
                            const double mine = slot == 0 ? m0 : (slot == 1 ? m1 : m2);
                            const double xj = __shfl_sync(0xffffffffu, mine, ln) * inv;
                            if (tid == ln) { if (slot == 0) m0 = xj; else if (slot == 1) m1 = xj; else m2 = xj; }
                            if (tid > j) m0 -= sM[tid * M_LD + j] * xj;
                            if (tid + 32 > j) m1 -= sM[(tid + 32) * M_LD + j] * xj;
                            if (tid + 64 > j && tid + 64 < NZ) m2 -= sM[(tid + 64) * M_LD + j] * xj;
                        }
                        sM[NZ * M_LD + tid] = m0; sM[NZ * M_LD + 32 + tid] = m1;
                        if (tid + 64 < NZ) sM[NZ * M_LD + 64 + tid] = m2;
                        double* Lrow = Lb + (size_t)k * L_SZ + NZ * M_LD;
                        Lrow[tid] = m0; Lrow[32 + tid] = m1;
                        if (tid + 64 < NZ) Lrow[64 + tid] = m2;
                    }
                    __syncthreads();
                    for (int e = tid; e < NX * NX; e += QP_THREADS) {
                        const int i = e / NX, j = e % NX;
                        sL[e] = (j <= i) ? sM[(NU + i) * M_LD + NU + j] : 0.0;
                    }
                    if (tid < NX) {
                        double s = 0.0;
                        for (int l = 0; l <= tid; l++) s += sM[(NU + tid) * M_LD + NU + l] * sM[NZ * M_LD + NU + l];
                        sv[SV_PN + tid] = s; vk[V_P + tid] = s;
                    }
                    __syncthreads();
                }
            }
            // ---------------- PASS B / D: forward sweep ----------------
            double a_p = 1.0, a_d = 1.0;
            if (tid < NX) sv[SV_DX + tid] = 0.0;
            __syncthreads();
            for (int k = 0; k <= N; k++) {
                const bool term = (k == N);
                double* vk = vb + (size_t)k * V_STAGE;
                if (tid == 0) {
                    mbar_arrive_expect_tx(&bar, (L_SZ + (term ? 0 : G_SZ)) * 8);
                    bulk_g2s(sM, Lb + (size_t)k * L_SZ, L_SZ * 8, &bar);
                    if (!term) bulk_g2s(sU, Gb + (size_t)k * G_SZ, G_SZ * 8, &bar);
                }
                for (int i = tid; i < NX; i += QP_THREADS) {
                    sv[SV_RB + i] = term ? 0.0 : vk[V_RB + i];
                    sv[SV_PN + i] = vk[V_P + i];
                }
                __syncthreads();
                mbar_wait(&bar, phase); phase ^= 1;
                // dpi_k = Lxx (Lxx^T dx) + p_k
                if (corr == 1 && k > 0) {
                    if (tid < NX) {
                        double s = 0.0;
                        for (int l = tid; l < NX; l++) s += sM[(NU + l) * M_LD + NU + tid] * sv[SV_DX + l];
                        sv[SV_TMP + tid] = s;
                    }
                    __syncthreads();
                    if (tid < NX) {
                        double s = sv[SV_PN + tid];
                        for (int l = 0; l <= tid; l++) s += sM[(NU + tid) * M_LD + NU + l] * sv[SV_TMP + l];
                        vb[(size_t)(k - 1) * V_STAGE + V_DPI + tid] = s;
                    }
                }
                if (!term) {
                    // rhs = lu + Lxu^T dx ; back substitution with Luu^T (warp 0)
                    if (tid < 32) {
                        double rhs = 0.0;
                        if (tid < NU) {
                            rhs = sM[NZ * M_LD + tid];
                            for (int l = 0; l < NX; l++) rhs += sM[(NU + l) * M_LD + tid] * sv[SV_DX + l];
                        }
                        double du = 0.0;
                        for (int i = NU - 1; i >= 0; i--) {
                            const double d = sM[i * M_LD + i];
                            const double inv = d != 0.0 ? 1.0 / d : 0.0;
                            const double dui = -__shfl_sync(0xffffffffu, rhs, i) * inv;
                            if (tid == i) du = dui;
                            if (tid < i) rhs += sM[i * M_LD + tid] * dui;
                        }
                        if (tid < NU) sv[SV_TMP + NX + tid] = du;   // temp slot, copied below
                    }
                    __syncthreads();
                    if (tid < NU) sv[SV_Z + tid] = sv[SV_TMP + NX + tid];
                } else {
                    if (tid < NU) sv[SV_Z + tid] = 0.0;
                }
                if (tid < NX) sv[SV_Z + NU + tid] = sv[SV_DX + tid];
                __syncthreads();
                for (int i = tid; i < NZ; i += QP_THREADS) vk[V_DZ + i] = sv[SV_Z + i];
                if (!term) {
                    const double s = rowdot4(sU, G_LD, sv + SV_Z, NX, NZ, tid);
                    if ((tid & 3) == 0 && (tid >> 2) < NX) sv[SV_DXN + (tid >> 2)] = s + sv[SV_RB + (tid >> 2)];
                }
                __syncthreads();
                // constraint rows: load C into the union buffer
                if (tid == 0) {
                    mbar_arrive_expect_tx(&bar, C_SZ * 8);
                    bulk_g2s(sU, Cb + (size_t)k * C_SZ, C_SZ * 8, &bar);
                }
                mbar_wait(&bar, phase); phase ^= 1;
                {
                    const double s = rowdot4(sU, C_LD, sv + SV_Z, NG, NZ, tid);
                    if ((tid & 3) == 0 && (tid >> 2) < NG) sv[SV_CZ + (tid >> 2)] = s;
                }
                __syncthreads();
                if (tid < 2 * NG) {
                    const int e = tid / NG, j = tid % NG;
                    const double msk = Db[(size_t)k * D_SZ + 2 * NG + tid];
                    double dtv = 0.0, dl = 0.0;
                    if (msk > 0.5) {
                        const double lam = qlam[(size_t)k * 2 * NG + tid], t = qt[(size_t)k * 2 * NG + tid];
                        const double cdz = sv[SV_CZ + j];
                        double rm = lam * t;
                        if (corr == 1) rm += vk[V_DT + tid] * vk[V_DLAM + tid] - sigma_mu;
                        dtv = (e ? -cdz : cdz) + vk[V_RD + tid];
                        dl = -(rm + lam * dtv) / t;
                        if (dtv < 0.0) a_p = fmin(a_p, -t / dtv);
                        if (dl < 0.0) a_d = fmin(a_d, -lam / dl);
                    }
                    vk[V_DT + tid] = dtv; vk[V_DLAM + tid] = dl;
                }
                if (tid < NX) sv[SV_DX + tid] = sv[SV_DXN + tid];
                __syncthreads();
            }
            a_p = block_reduce(a_p, 1, red, tid);
            a_d = block_reduce(a_d, 1, red, tid);
            if (corr == 0) {
                const double alpha_aff = fmin(a_p, a_d);
                double s = 0.0;
                for (int i = tid; i < (N + 1) * 2 * NG; i += QP_THREADS) {
                    const int k = i / (2 * NG), id = i % (2 * NG);
                    if (Db[(size_t)k * D_SZ + 2 * NG + id] > 0.5) {
                        const double* vk = vb + (size_t)k * V_STAGE;
                        s += (qlam[i] + alpha_aff * vk[V_DLAM + id]) * (qt[i] + alpha_aff * vk[V_DT + id]);
                    }
                }
                const double mu_aff = nc > 0.0 ? block_reduce(s, 2, red, tid) / nc : 0.0;
                double sg = mu > 0.0 ? mu_aff / mu : 0.0;
                sg = sg * sg * sg;
                sigma_mu = sg * mu;
            } else {
                alpha_p = a_p; alpha_d = a_d;
            }
        }
        alpha_p *= (1.0 - alpha_p) * 0.99 + alpha_p * 0.9999999;
        alpha_d *= (1.0 - alpha_d) * 0.99 + alpha_d * 0.9999999;
        if (alpha_p < IPM_ALPHA_MIN && alpha_d < IPM_ALPHA_MIN) { status = 2; break; }
        // =============================== PASS E: update ===============================
        for (int i = tid; i < (N + 1) * NZ; i += QP_THREADS) {
            const int k = i / NZ, c = i % NZ;
            qz[i] += alpha_p * vb[(size_t)k * V_STAGE + V_DZ + c];
        }
        for (int i = tid; i < N * NX; i += QP_THREADS) {
            const int k = i / NX, c = i % NX;
            qpi[i] += alpha_d * vb[(size_t)k * V_STAGE + V_DPI + c];
        }
        for (int i = tid; i < (N + 1) * 2 * NG; i += QP_THREADS) {
            const int k = i / (2 * NG), id = i % (2 * NG);
            if (Db[(size_t)k * D_SZ + 2 * NG + id] > 0.5) {
                const double* vk = vb + (size_t)k * V_STAGE;
                qt[i] = fmax(qt[i] + alpha_p * vk[V_DT + id], IPM_T_MIN);
                qlam[i] = fmax(qlam[i] + alpha_d * vk[V_DLAM + id], IPM_LAM_MIN);
            }
        }
        __syncthreads();
    }

    // ---- epilogue: stats, SQP full step ----
    if (tid == 0) {
        a.qp_iter[b] = it; a.qp_status[b] = status;
        a.qp_res[(size_t)b * 4 + 0] = res_g; a.qp_res[(size_t)b * 4 + 1] = res_b;
        a.qp_res[(size_t)b * 4 + 2] = res_d; a.qp_res[(size_t)b * 4 + 3] = res_m;
        if (a.apply && a.sqp_iter) a.sqp_iter[b] += 1;
        bulk_store_wait_all();
    }
    if (a.apply && (status == 0 || status == 1)) {
        double* x = a.x + (size_t)b * (N + 1) * NX;
        double* u = a.u + (size_t)b * N * NU;
        double* pi = a.pi + (size_t)b * N * NX;
        double* lam = a.lam + (size_t)b * (N + 1) * 2 * NG;
        for (int i = tid; i < (N + 1) * NX; i += QP_THREADS) x[i] += qz[(size_t)(i / NX) * NZ + NU + i % NX];
        for (int i = tid; i < N * NU; i += QP_THREADS) u[i] += qz[(size_t)(i / NU) * NZ + i % NU];
        for (int i = tid; i < N * NX; i += QP_THREADS) pi[i] = qpi[i];
        for (int i = tid; i < (N + 1) * 2 * NG; i += QP_THREADS) {
            const int k = i / (2 * NG), id = i % (2 * NG);
            lam[i] = Db[(size_t)k * D_SZ + 2 * NG + id] > 0.5 ? qlam[i] : 0.0;
        }
    }
}
