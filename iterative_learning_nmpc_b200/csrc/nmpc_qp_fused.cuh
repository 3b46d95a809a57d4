// nmpc_qp_fused.cuh -- the product's OCP-QP interior point: ONE persistent CTA per OCP runs the
// whole Mehrotra predictor-corrector IPM (all iterations, all stages) with a Riccati recursion
// specialised to the structure of this OCP family.
//
// HPIPM `speed` semantics (oracle:
// oracle/nmpc_oracle.c::qp_ipm) -- the Newton directions are identical in exact arithmetic -- but
// the arithmetic is organised around what the stage matrices of THIS problem look like:
//
//   dynamics   q+ = q + dt v + c a,  v+ = v + dt a,  h+ = h + Hq q + Hf f        (c = dt^2/2)
//              -> G = [B A] is never formed: G'PG is O(n^2) combinations of blocks of P plus
//                 rank-6 corrections through Hq (6x18), Hf (6x12)
//   cost       diagonal in (a, f, v, h), dense only in the 18x18 q block
//   constraints friction rows touch one foot's 3 force columns; contact / patch / centroidal rows
//              touch (q, v) (+ an identity column in h)
//              -> H + C' Gamma C = diag(18) + four 3x3 blocks in u, dense 42x42 in x, zero u-x block
//
// so a stage is described by a 1692-double "stage record" (SR) instead of 11.6k doubles of dense
// blocks, and the Riccati step is the CLASSICAL recursion on P (42x42): Cholesky of the 30x30
// input block only (30 sequential pivots instead of 72), Lxu = Mxu Luu^-T, P = Mxx - Lxu Lxu'.
//
// Fusion: the stage-parallel work of an IPM iteration (residuals, Gamma, barrier Hessian) is done
// by the same CTA right before it factorises the stage, all reductions over stages (norms, mu,
// step lengths, sigma) are CTA-local, and nothing but the per-stage vectors and the factors
// (needed again by the forward / corrector sweeps) goes back to HBM.  One kernel launch per QP.
#pragma once
#include "nmpc_qp_layout.cuh"

// ---------------------------------------------------------------------------------------------
// stage record (doubles)
// ---------------------------------------------------------------------------------------------
#define SR_HQ 0                      // [6][18]  d h+ / d q
#define SR_HF (SR_HQ + 6 * NQ)       // [6][12]  d h+ / d f
#define SR_B (SR_HF + 6 * 12)        // [42]     dynamics defect f(x,u) - x+
#define SR_RD (SR_B + NX)            // [30]     Hessian diagonal, u block
#define SR_QV (SR_RD + NU)           // [18]     Hessian diagonal, v block
#define SR_QH (SR_QV + NQ)           // [6]      Hessian diagonal, h block
#define SR_QQ (SR_QH + 6)            // [18][18] Hessian, q block (full symmetric)
#define SR_GR (SR_QQ + NQ * NQ)      // [72]     gradient
#define SR_CF (SR_GR + NZ)           // [20][3]  friction rows over their foot's force
#define SR_CXLD 37                   // row stride of CX: 36 + 1, so that thread-per-row and thread-per-column reads are both free of bank conflicts
#define SR_CX (SR_CF + 20 * 3)       // [22][37] contact / patch / centroidal rows over (q, v)
#define SR_D (SR_CX + 22 * SR_CXLD)       // lg[42] ug[42] mask_l[42] mask_u[42]
// time-opt extension of the stage record (enable_time_opt: the node duration dt_k is a 31st input; zero otherwise)
#define SR_X (SR_D + 4 * NG)         // 1714
#define SRX_B 0                      // [42] column of the dynamics over dt: d x+ / d dt = [v + dt a; a; hdot]
#define SRX_H 42                     // Hessian entry of dt (weight + regularisation)
#define SRX_G 43                     // gradient entry: weight (dt - dt_nodes)
#define SRX_LB 44                    // dt_min - dt
#define SRX_UB 45                    // dt_max - dt
#define SRX_DT 46                    // dt_k of this linearisation
#define SRX_C2 47                    // dt_k^2 / 2
#define SRX_SZ 48
#define SR_SZ (SR_X + SRX_SZ)        // 1762
#define NGX 22                       // constraint rows on the state
#define NGF 20                       // friction rows

// factor record (doubles): u panels of the stage factor (Luu, Lxu, row 72 = lu) | cost-to-go Hessian P_k
#define P_LD 43
#define FR_MU 0
#define FR_P (FR_MU + LP_UPANELS)            // 1830
#define FR_SZ 3680                           // 1830 + 42 * 43 = 3636, padded

static_assert(SR_SZ % 2 == 0 && SR_X % 2 == 0 && SR_CF % 2 == 0 && FR_P % 2 == 0 && FR_SZ % 2 == 0, "16-byte alignment of bulk copies");

// ---------------------------------------------------------------------------------------------
// dense stage blocks -> stage record (used for externally supplied QPs and, for now, after the
// dense linearisation).  One CTA per (stage, problem).  With `check` set, every dense entry is
// compared with the value the record implies; a mismatch raises *flag (the caller then falls back
// to the generic dense path, nmpc_qp_ipm.cuh).
// ---------------------------------------------------------------------------------------------
struct SrArgs {
    const double* G; const double* M; const double* C; const double* D;
    double* SR;
    int32_t* flag;
    const int32_t* done;
    int N, b0, check;
    double dt;
};

__device__ __forceinline__ bool sr_differs(double a, double b) { return !(a == b) && !(a != a && b != b); }

__global__ void __launch_bounds__(128) nmpc_dense2sr_kernel(SrArgs a)
{
    const int k = blockIdx.x, bl = blockIdx.y, tid = threadIdx.x, N = a.N;
    if (a.done && a.done[a.b0 + bl]) return;
    const bool term = (k == N);
    const double* Gp = term ? nullptr : a.G + ((size_t)bl * N + k) * G_SZ;
    const double* Mp = a.M + ((size_t)bl * (N + 1) + k) * M_SZ;
    const double* Cp = a.C + ((size_t)bl * (N + 1) + k) * C_SZ;
    const double* Dp = a.D + ((size_t)bl * (N + 1) + k) * D_SZ;
    double* S = a.SR + ((size_t)bl * (N + 1) + k) * SR_SZ;
    for (int e = tid; e < 6 * NQ; e += 128) S[SR_HQ + e] = term ? 0.0 : Gp[(2 * NQ + e / NQ) * G_LD + IQ + e % NQ];
    for (int e = tid; e < 6 * 12; e += 128) S[SR_HF + e] = term ? 0.0 : Gp[(2 * NQ + e / 12) * G_LD + IFRC + e % 12];
    for (int e = tid; e < NX; e += 128) S[SR_B + e] = term ? 0.0 : Gp[e * G_LD + NZ];
    for (int e = tid; e < NU; e += 128) S[SR_RD + e] = Mp[e * M_LD + e];
    for (int e = tid; e < NQ; e += 128) S[SR_QV + e] = Mp[(IV + e) * M_LD + IV + e];
    for (int e = tid; e < 6; e += 128) S[SR_QH + e] = Mp[(IH + e) * M_LD + IH + e];
    for (int e = tid; e < NQ * NQ; e += 128) S[SR_QQ + e] = Mp[(IQ + e / NQ) * M_LD + IQ + e % NQ];
    for (int e = tid; e < NZ; e += 128) S[SR_GR + e] = Mp[NZ * M_LD + e];
    for (int e = tid; e < NGF * 3; e += 128) { const int j = e / 3; S[SR_CF + e] = Cp[j * C_LD + IFRC + 3 * (j / 5) + e % 3]; }
    for (int e = tid; e < NGX * 36; e += 128) S[SR_CX + SR_CXLD * (e / 36) + e % 36] = Cp[(NGF + e / 36) * C_LD + IQ + e % 36];
    for (int e = tid; e < 4 * NG; e += 128) S[SR_D + e] = Dp[e];
    if (!a.check) return;
    // ---- structure check: the dense blocks must be exactly what the record expands to ----
    bool bad = false;
    const double dt = a.dt, c2 = 0.5 * dt * dt;
    if (!term) {
        for (int e = tid; e < NX * NZ; e += 128) {
            const int l = e / NZ, c = e % NZ;
            double ex = 0.0;
            if (l < NQ) { if (c == IA + l) ex = c2; else if (c == IQ + l) ex = 1.0; else if (c == IV + l) ex = dt; }
            else if (l < 2 * NQ) { const int i = l - NQ; if (c == IA + i) ex = dt; else if (c == IV + i) ex = 1.0; }
            else {
                const int i = l - 2 * NQ;
                if (c >= IFRC && c < IQ) ex = Gp[l * G_LD + c];
                else if (c >= IQ && c < IV) ex = Gp[l * G_LD + c];
                else if (c == IH + i) ex = 1.0;
            }
            bad |= sr_differs(Gp[l * G_LD + c], ex);
        }
    }
    for (int e = tid; e < NZ * NZ; e += 128) {
        const int i = e / NZ, j = e % NZ;
        const bool in_q = (i >= IQ && i < IV && j >= IQ && j < IV);
        if (i != j && !in_q) bad |= sr_differs(Mp[i * M_LD + j], 0.0);
        if (in_q) bad |= sr_differs(Mp[i * M_LD + j], Mp[j * M_LD + i]);
    }
    for (int e = tid; e < NG * NZ; e += 128) {
        const int j = e / NZ, c = e % NZ;
        bool allowed;
        double fixed = 0.0;
        bool has_fixed = false;
        if (j < NGF) allowed = (c >= IFRC + 3 * (j / 5) && c < IFRC + 3 * (j / 5) + 3);
        else {
            allowed = (c >= IQ && c < IH);
            if (j >= GR_CONS && c >= IH) { has_fixed = true; fixed = (c == IH + (j - GR_CONS)) ? 1.0 : 0.0; }
        }
        const double v = Cp[j * C_LD + c];
        // a row with both sides masked out contributes nothing (Gamma = gamma = 0): anything goes
        if (!(Dp[2 * NG + j] > 0.5 || Dp[3 * NG + j] > 0.5)) continue;
        if (has_fixed) bad |= sr_differs(v, fixed);
        else if (!allowed) bad |= sr_differs(v, 0.0);
    }
    if (bad) atomicOr(a.flag, 1);
}

// ---------------------------------------------------------------------------------------------
// fused IPM kernel
// ---------------------------------------------------------------------------------------------
// C z for the current stage: sv[FV_CZ + j] = sum_c C(j, c) zz[c], zz over [u; x]
__device__ __forceinline__ void fq_cz(const double* S, const double* zz, double* out, int tid)
{
    if (tid < NG) {
        const int j = tid;
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        if (j < NGF) {
            const double* cf = S + SR_CF + 3 * j;
            const double* f = zz + IFRC + 3 * (j / 5);
            s0 = cf[0] * f[0] + cf[1] * f[1] + cf[2] * f[2];
        } else {
            const double* cx = S + SR_CX + SR_CXLD * (j - NGF);
            const double* xx = zz + IQ;
#pragma unroll
            for (int c = 0; c < 36; c += 4) {
                s0 += cx[c] * xx[c]; s1 += cx[c + 1] * xx[c + 1]; s2 += cx[c + 2] * xx[c + 2]; s3 += cx[c + 3] * xx[c + 3];
            }
            if (j >= GR_CONS) s0 += zz[IH + j - GR_CONS];
        }
        out[j] = (s0 + s1) + (s2 + s3);
    }
}

// (C' g)[c] for a multiplier-like vector g over the 42 constraint rows
__device__ __forceinline__ double fq_ct(const double* S, const double* g, int c)
{
    double s = 0.0;
    if (c < IFRC) return 0.0;
    if (c < IQ) {
        const int ft = (c - IFRC) / 3, i = (c - IFRC) % 3;
#pragma unroll
        for (int r = 0; r < 5; r++) s += S[SR_CF + 3 * (5 * ft + r) + i] * g[5 * ft + r];
        return s;
    }
    if (c < IH) {
        const double* cx = S + SR_CX + (c - IQ);
        double s1 = 0.0;
#pragma unroll
        for (int jj = 0; jj < NGX; jj += 2) { s += cx[SR_CXLD * jj] * g[NGF + jj]; s1 += cx[SR_CXLD * (jj + 1)] * g[NGF + jj + 1]; }
        return s + s1;
    }
    return g[GR_CONS + c - IH];
}

// (C' g1)[c] and (C' g2)[c] with one pass over the matrix entries
__device__ __forceinline__ void fq_ct2(const double* S, const double* g1, const double* g2, int c, double& o1, double& o2)
{
    double s = 0.0, q = 0.0;
    if (c < IFRC) { o1 = 0.0; o2 = 0.0; return; }
    if (c < IQ) {
        const int ft = (c - IFRC) / 3, i = (c - IFRC) % 3;
#pragma unroll
        for (int r = 0; r < 5; r++) { const double m = S[SR_CF + 3 * (5 * ft + r) + i]; s = fma(m, g1[5 * ft + r], s); q = fma(m, g2[5 * ft + r], q); }
        o1 = s; o2 = q; return;
    }
    if (c < IH) {
        const double* cx = S + SR_CX + (c - IQ);
        double s1 = 0.0, q1 = 0.0;
#pragma unroll
        for (int jj = 0; jj < NGX; jj += 2) {
            const double m0 = cx[SR_CXLD * jj], m1 = cx[SR_CXLD * (jj + 1)];
            s = fma(m0, g1[NGF + jj], s); s1 = fma(m1, g1[NGF + jj + 1], s1);
            q = fma(m0, g2[NGF + jj], q); q1 = fma(m1, g2[NGF + jj + 1], q1);
        }
        o1 = s + s1; o2 = q + q1; return;
    }
    o1 = g1[GR_CONS + c - IH]; o2 = g2[GR_CONS + c - IH];
}

// (G' w)[c] for w over the next state (structured dynamics)
__device__ __forceinline__ double fq_gt(const double* S, const double* w, int c, double dt, double c2)
{
    if (c < IFRC) return c2 * w[c] + dt * w[NQ + c];
    if (c < IQ) {
        const int j = c - IFRC;
        double s = 0.0;
#pragma unroll
        for (int h = 0; h < 6; h++) s += S[SR_HF + 12 * h + j] * w[2 * NQ + h];
        return s;
    }
    if (c < IV) {
        const int i = c - IQ;
        double s = w[i];
#pragma unroll
        for (int h = 0; h < 6; h++) s += S[SR_HQ + NQ * h + i] * w[2 * NQ + h];
        return s;
    }
    if (c < IH) { const int i = c - IV; return dt * w[i] + w[NQ + i]; }
    return w[2 * NQ + c - IH];
}

// (G zz)[l] for zz over [u; x] (no affine term)
__device__ __forceinline__ double fq_g(const double* S, const double* zz, int l, double dt, double c2)
{
    if (l < NQ) return zz[IQ + l] + dt * zz[IV + l] + c2 * zz[IA + l];
    if (l < 2 * NQ) { const int i = l - NQ; return zz[IV + i] + dt * zz[IA + i]; }
    const int i = l - 2 * NQ;
    double s = zz[IH + i], s1 = 0.0;
    const double* hq = S + SR_HQ + NQ * i;
    const double* hf = S + SR_HF + 12 * i;
#pragma unroll
    for (int j = 0; j < NQ; j += 2) { s += hq[j] * zz[IQ + j]; s1 += hq[j + 1] * zz[IQ + j + 1]; }
#pragma unroll
    for (int j = 0; j < 12; j += 2) { s += hf[j] * zz[IFRC + j]; s1 += hf[j + 1] * zz[IFRC + j + 1]; }
    return s + s1;
}

// (H zz)[c]: Gauss-Newton Hessian times a vector over [u; x]
__device__ __forceinline__ double fq_h(const double* S, const double* zz, int c)
{
    if (c < IQ) return S[SR_RD + c] * zz[c];
    if (c < IV) {
        const double* row = S + SR_QQ + NQ * (c - IQ);
        double s = 0.0, s1 = 0.0;
#pragma unroll
        for (int j = 0; j < NQ; j += 2) { s += row[j] * zz[IQ + j]; s1 += row[j + 1] * zz[IQ + j + 1]; }
        return s + s1;
    }
    if (c < IH) return S[SR_QV + c - IV] * zz[c];
    return S[SR_QH + c - IH] * zz[c];
}

__device__ __noinline__ double fq_block_max(double v, double* red, int tid)
{
    v = warp_max(v);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    return nanmax(nanmax(red[0], red[1]), nanmax(red[2], red[3]));
}
__device__ __noinline__ double fq_block_min(double v, double* red, int tid)
{
    v = warp_min(v);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    return fmin(fmin(red[0], red[1]), fmin(red[2], red[3]));
}
__device__ __noinline__ double fq_block_sum(double v, double* red, int tid)
{
    v = warp_sum(v);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    return ((red[0] + red[1]) + red[2]) + red[3];
}

// ---------------------------------------------------------------------------------------------
// register tiles of the stage matrix, one entry per thread: {row type, column type, row block, column block}.  Tiles of
// one block type are kept in one warp wherever they fit (a warp executes every assembly formula it holds, one after the
// other): warp 0 meets 2 u and 1 x formula, warps 1 and 3 two of each, warp 2 three and two.  Row type 7 (u) / 3 (x)
// marks a thread without a tile; for those c_xtile[tid][2] numbers the 23 threads without a state tile.
//   u tiles (3 x 6): rows of z = [a f q v h] (types 0..4) x columns a / f (types 0 / 1)
//   x tiles (3 x 3): rows q / v / h (types 0..2) x columns q / v / h of the 42 x 42 state block
// ---------------------------------------------------------------------------------------------
__constant__ unsigned char c_utile[128][4] = {{2,0,10,0}, {2,0,11,0}, {2,0,12,0}, {2,0,13,0}, {2,0,14,0}, {2,0,15,0}, {2,0,10,1}, {2,0,11,1}, {2,0,12,1}, {2,0,13,1}, {2,0,14,1}, {2,0,15,1}, {2,0,10,2}, {2,0,11,2}, {2,0,12,2}, {2,0,13,2}, {2,0,14,2}, {2,0,15,2}, {1,0,6,0}, {1,0,7,0}, {1,0,8,0}, {1,0,9,0}, {1,0,6,1}, {1,0,7,1}, {1,0,8,1}, {1,0,9,1}, {1,0,6,2}, {1,0,7,2}, {1,0,8,2}, {1,0,9,2}, {7,0,0,0}, {7,0,0,0}, {3,0,16,0}, {3,0,17,0}, {3,0,18,0}, {3,0,19,0}, {3,0,20,0}, {3,0,21,0}, {3,0,16,1}, {3,0,17,1}, {3,0,18,1}, {3,0,19,1}, {3,0,20,1}, {3,0,21,1}, {3,0,16,2}, {3,0,17,2}, {3,0,18,2}, {3,0,19,2}, {3,0,20,2}, {3,0,21,2}, {0,0,0,0}, {0,0,1,0}, {0,0,2,0}, {0,0,3,0}, {0,0,4,0}, {0,0,5,0}, {0,0,2,1}, {0,0,3,1}, {0,0,4,1}, {0,0,5,1}, {0,0,4,2}, {0,0,5,2}, {7,0,0,0}, {7,0,0,0}, {2,1,10,3}, {2,1,11,3}, {2,1,12,3}, {2,1,13,3}, {2,1,14,3}, {2,1,15,3}, {2,1,10,4}, {2,1,11,4}, {2,1,12,4}, {2,1,13,4}, {2,1,14,4}, {2,1,15,4}, {3,1,16,3}, {3,1,17,3}, {3,1,18,3}, {3,1,19,3}, {3,1,20,3}, {3,1,21,3}, {3,1,16,4}, {3,1,17,4}, {3,1,18,4}, {3,1,19,4}, {3,1,20,4}, {3,1,21,4}, {4,0,22,0}, {4,0,23,0}, {4,0,22,1}, {4,0,23,1}, {4,0,22,2}, {4,0,23,2}, {7,0,0,0}, {7,0,0,0}, {1,1,6,3}, {1,1,7,3}, {1,1,8,3}, {1,1,9,3}, {1,1,8,4}, {1,1,9,4}, {4,1,22,3}, {4,1,23,3}, {4,1,22,4}, {4,1,23,4}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}, {7,0,0,0}};
__constant__ unsigned char c_xtile[128][4] = {{1,0,6,0}, {1,0,6,1}, {1,0,6,2}, {1,0,6,3}, {1,0,6,4}, {1,0,6,5}, {1,0,7,0}, {1,0,7,1}, {1,0,7,2}, {1,0,7,3}, {1,0,7,4}, {1,0,7,5}, {1,0,8,0}, {1,0,8,1}, {1,0,8,2}, {1,0,8,3}, {1,0,8,4}, {1,0,8,5}, {1,0,9,0}, {1,0,9,1}, {1,0,9,2}, {1,0,9,3}, {1,0,9,4}, {1,0,9,5}, {1,0,10,0}, {1,0,10,1}, {1,0,10,2}, {1,0,10,3}, {1,0,10,4}, {1,0,10,5}, {1,0,11,0}, {1,0,11,1}, {1,0,11,2}, {1,0,11,3}, {1,0,11,4}, {1,0,11,5}, {0,0,0,0}, {0,0,1,0}, {0,0,1,1}, {0,0,2,0}, {0,0,2,1}, {0,0,2,2}, {0,0,3,0}, {0,0,3,1}, {0,0,3,2}, {0,0,3,3}, {0,0,4,0}, {0,0,4,1}, {0,0,4,2}, {0,0,4,3}, {0,0,4,4}, {0,0,5,0}, {0,0,5,1}, {0,0,5,2}, {0,0,5,3}, {0,0,5,4}, {0,0,5,5}, {3,0,0,0}, {3,0,1,0}, {3,0,2,0}, {3,0,3,0}, {3,0,4,0}, {3,0,5,0}, {3,0,6,0}, {1,1,6,6}, {1,1,7,6}, {1,1,7,7}, {1,1,8,6}, {1,1,8,7}, {1,1,8,8}, {1,1,9,6}, {1,1,9,7}, {1,1,9,8}, {1,1,9,9}, {1,1,10,6}, {1,1,10,7}, {1,1,10,8}, {1,1,10,9}, {1,1,10,10}, {1,1,11,6}, {1,1,11,7}, {1,1,11,8}, {1,1,11,9}, {1,1,11,10}, {1,1,11,11}, {2,2,12,12}, {2,2,13,12}, {2,2,13,13}, {3,0,7,0}, {3,0,8,0}, {3,0,9,0}, {3,0,10,0}, {3,0,11,0}, {3,0,12,0}, {3,0,13,0}, {3,0,14,0}, {2,0,12,0}, {2,0,12,1}, {2,0,12,2}, {2,0,12,3}, {2,0,12,4}, {2,0,12,5}, {2,0,13,0}, {2,0,13,1}, {2,0,13,2}, {2,0,13,3}, {2,0,13,4}, {2,0,13,5}, {2,1,12,6}, {2,1,12,7}, {2,1,12,8}, {2,1,12,9}, {2,1,12,10}, {2,1,12,11}, {2,1,13,6}, {2,1,13,7}, {2,1,13,8}, {2,1,13,9}, {2,1,13,10}, {2,1,13,11}, {3,0,15,0}, {3,0,16,0}, {3,0,17,0}, {3,0,18,0}, {3,0,19,0}, {3,0,20,0}, {3,0,21,0}, {3,0,22,0}};

// blocks of P_{k+1} (sP, ld 43; state order q 0..17, v 18..35, h 36..41) and of U = P[:,h] [Hf | Hq] (sU, ld 30)
#define PQQ(i, j) sP[(i) * P_LD + (j)]
#define PQV(i, j) sP[(i) * P_LD + NQ + (j)]
#define PVQ(i, j) sP[(NQ + (i)) * P_LD + (j)]
#define PVV(i, j) sP[(NQ + (i)) * P_LD + NQ + (j)]
#define PHQ(h, j) sP[(2 * NQ + (h)) * P_LD + (j)]
#define PHV(h, j) sP[(2 * NQ + (h)) * P_LD + NQ + (j)]
#define PHH(h, g) sP[(2 * NQ + (h)) * P_LD + 2 * NQ + (g)]
#define UF(r, j) sU[(r) * NU + (j)]
#define UQ(r, j) sU[(r) * NU + 12 + (j)]

// (G' P G) on a 3 x 6 tile: rows i0.. of block type RT (a f q v h), columns j0.. of block type CT (a f)
template <int RT, int CT>
__device__ __forceinline__ void fq_asm_u(const double* sP, const double* sU, const double* S, int i0, int j0,
                                         double dt, double c2, double (&t)[3][6])
{
#pragma unroll
    for (int r = 0; r < 3; r++) {
        const int i = i0 + r;
#pragma unroll
        for (int c = 0; c < 6; c++) {
            const int j = j0 + c;
            double v;
            if (CT == 0) {
                if (RT == 0) v = c2 * c2 * PQQ(i, j) + c2 * dt * (PQV(i, j) + PVQ(i, j)) + dt * dt * PVV(i, j);
                else if (RT == 1) v = c2 * UF(j, i) + dt * UF(NQ + j, i);
                else if (RT == 2) v = c2 * (PQQ(i, j) + UQ(j, i)) + dt * (PQV(i, j) + UQ(NQ + j, i));
                else if (RT == 3) v = c2 * dt * PQQ(i, j) + c2 * PVQ(i, j) + dt * dt * PQV(i, j) + dt * PVV(i, j);
                else v = c2 * PHQ(i, j) + dt * PHV(i, j);
            } else {
                if (RT == 1) {
                    v = 0.0;
#pragma unroll
                    for (int h = 0; h < 6; h++) v += S[SR_HF + 12 * h + i] * UF(2 * NQ + h, j);
                } else if (RT == 2) {
                    v = UF(i, j);
#pragma unroll
                    for (int h = 0; h < 6; h++) v += S[SR_HQ + NQ * h + i] * UF(2 * NQ + h, j);
                } else if (RT == 3) v = dt * UF(i, j) + UF(NQ + i, j);
                else v = UF(2 * NQ + i, j);
            }
            t[r][c] = v;
        }
    }
}

// (G' P G) on a 3 x 3 tile of the state block: row type RT, column type CT in (q v h)
template <int RT, int CT>
__device__ __forceinline__ void fq_asm_x(const double* sP, const double* sU, const double* S, int i0, int j0,
                                         double dt, double (&t)[3][3])
{
#pragma unroll
    for (int r = 0; r < 3; r++) {
        const int i = i0 + r;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const int j = j0 + c;
            double v;
            if (RT == 0) {
                v = PQQ(i, j) + UQ(i, j) + UQ(j, i);
#pragma unroll
                for (int h = 0; h < 6; h++) v += S[SR_HQ + NQ * h + i] * UQ(2 * NQ + h, j);
            } else if (RT == 1 && CT == 0) v = dt * PQQ(i, j) + PVQ(i, j) + dt * UQ(i, j) + UQ(NQ + i, j);
            else if (RT == 1) v = dt * dt * PQQ(i, j) + dt * (PQV(i, j) + PVQ(i, j)) + PVV(i, j);
            else if (CT == 0) v = PHQ(i, j) + UQ(2 * NQ + i, j);
            else if (CT == 1) v = dt * PHQ(i, j) + PHV(i, j);
            else v = PHH(i, j);
            t[r][c] = v;
        }
    }
}


// ---------------------------------------------------------------------------------------------
// fused IPM kernel: one persistent CTA per OCP-QP.  An IPM iteration is a sequence of PASSES over the stages:
//
//   R   k = N..0  bulk   apply the previous step; residuals rd rm rg rb, Gamma, q~, norms, mu       (pipelined)
//       exit tests
//   F   k = N..0  CHAIN  Riccati factorisation on P (the only pass with a long per-stage chain); leaves per stage
//                        the feedback gain K = Luu^-T Lxu', W = Luu^-1, P_k and P_{k+1} rb_k
//   A1  k = 0..N  chain  affine direction: du = -(kff + K dx), dx+ = G dz + rb   (two short phases per stage)
//   A2  all k     bulk   C dz, dt, dlam, step lengths, sums for the centring parameter               (pipelined)
//   C1  all k     bulk   corrector complementarity residual, gamma_c, m0 = rg + C' gamma_c + G' P+ rb (pipelined)
//   C2  k = N..0  chain  p_k = m_x - K' m_u with m = m0 + G' p+;  off the chain: kff = W' W m_u
//   D1  k = 0..N  chain  corrector direction (as A1)
//   D2  all k     bulk   dpi_{k-1} = P_k dx_k + p_k, C dz, dt, dlam, step lengths                    (pipelined)
//
// Only F carries the long dependent chain; in every other pass the serial dependence between stages is a handful
// of short dot products (no triangular solve: the gains are explicit), and everything that does not depend on the
// neighbouring stage is done in a bulk pass whose inputs arrive through a ring of bulk (TMA) copies several stages
// ahead.  Every pass is a separate non-inlined function: the register peak of the factorisation does not leak.
// ---------------------------------------------------------------------------------------------
#define FQ_THREADS 128
#define FQ_WARPS (FQ_THREADS / 32)
#define N_UTILES 100                 // 3x6 tiles of rows 0..71 x columns 0..29 at or below the diagonal
#define N_XTILES 105                 // 3x3 tiles of the lower triangle of the 42x42 state block

struct FqArgs {
    const double* SR;               // [Bc][N+1][SR_SZ]
    double* FR;                     // [Bc][N+1][FR_SZ]
    double* vec;                    // [Bc][(N+1)*V_STAGE]
    double* qz; double* qpi; double* qlam; double* qt;       // QP iterate, global problem index
    double* qx;                     // [B][N][QX_SZ] time-opt: QP iterate of dt and of its two bounds
    int32_t* qp_iter; int32_t* qp_status; int32_t* qp_ncond; double* qp_res;
    const double* dx0_ext; const double* x0; const double* x;
    const int32_t* sqp_done;
    int N, b0, max_iter, warm, cond_pred_corr;   // warm: HPIPM warm_start level 0 / 1 / 2 (see nmpc_b200.h)
    int bl0;                        // first problem of this launch inside the chunk (a chunk is launched wave by wave)
    // SQP full step at the end of the QP (apply != 0): (x, u) += dz, multipliers replaced by the QP's, dt += tau
    int apply; double* sx; double* su; double* spi; double* slam; int32_t* sqp_iter; double* dtn; double* lamt;
    double tol, dt, c2;              // c2 = dt^2 / 2
};

// per-CTA context in shared memory: base pointers of this OCP (no registers held across the passes)
struct FqCtx {
    const double* SRb; double* FRb; double* vb; double* qz; double* qpi; double* qlam; double* qt; double* qx;
    double dt, c2;
    int N;
};

extern __shared__ __align__(16) double fq_smem[];
__shared__ FqCtx fq_ctx;
__shared__ __align__(8) uint64_t fq_bars[4];

#define FQ_SMEM_DOUBLES 7096
#define FQ_SMEM_BYTES (FQ_SMEM_DOUBLES * 8)
// Dynamic shared memory that caps the number of co-resident CTAs per SM at 3 / 2 / 1 (233472 bytes per SM, 1 KB of it
// reserved per CTA, ~150 bytes static): the last, partial wave of a launch plan is spread over all SMs with it.
#define FQ_SMEM_CAP3 (60 * 1024)
#define FQ_SMEM_CAP2 (80 * 1024)
#define FQ_SMEM_CAP1 (120 * 1024)
#define FQ_ISSUER 96                         // the thread that issues the bulk copies of the rings: first lane of warp 3, whose
                                             // lanes have the least work in every pass (thread 0 is on every critical path)
#define FQ_RED (FQ_SMEM_DOUBLES - 16)        // 16 reduction slots, the same place in every pass

// ---- pass R: ring of 2 slots (a stage of this pass takes far longer than the latency of its copies) ----
#define R_SR 0
#define R_Z (R_SR + SR_SZ)
#define R_DZ (R_Z + NZ)
#define R_PI (R_DZ + NZ)                     // pi_{k-1}
#define R_DPI (R_PI + NX)
#define R_LAM (R_DPI + NX)
#define R_T (R_LAM + 2 * NG)
#define R_DT (R_T + 2 * NG)                  // dt | dlam
#define R_XQ (R_DT + 4 * NG)                 // time-opt: dt iterate record (QX_SZ)
#define R_XV (R_XQ + QX_SZ)                  // time-opt: V_X block of the stage
#define R_SLOT (R_XV + VX_SZ)
#define R_DEPTH 2
#define RV_BASE (R_DEPTH * R_SLOT)           // 6768
#define RV_ZC 0                              // [2][72] updated z_k, ping-pong by stage parity (x part = "next state" of stage k-1)
#define RV_PIN (RV_ZC + 2 * NZ)              // [2][42] updated pi_{k-1}, ping-pong
#define RV_CZ (RV_PIN + 2 * NX)              // 42: C z, then lam_l - lam_u
#define RV_GAMV (RV_CZ + NG)                 // 42
#define RV_X (RV_GAMV + NG)                  // time-opt scalars: tau, lam_l - lam_u, gamma
#define RV_END (RV_X + 4)
static_assert(RV_BASE + RV_END <= FQ_RED, "pass R shared memory");
static_assert(R_SLOT % 2 == 0 && R_PI % 2 == 0 && R_DPI % 2 == 0 && R_LAM % 2 == 0, "alignment");

// ---- pass F ----
#define F_P 0                                // P_{k+1}, full symmetric, ld 43
#define F_M (F_P + NX * P_LD)                // u panels of the stage matrix; before the factorisation U = P[:,h][Hf|Hq]; after it K | W
#define F_S (F_M + LP_UPANELS)               // stage record | QT GAM KFF RB of the stage (186)
#define F_SV (F_S + SR_SZ)
#define F_XV (F_SV + 186)                    // time-opt: V_X block of the stage
#define F_V (F_XV + VX_SZ)
#define FV_QT 0                              // 72  working linear term m (row 72 of the stage matrix)
#define FV_W (FV_QT + NZ)                    // 42
#define FV_PN (FV_W + NX)                    // 42  p_{k+1}
#define FV_PB (FV_PN + NX)                   // 42  time-opt: P+ b
#define FV_XS (FV_PB + NX)                   // 4   time-opt scalars: 1 / d, s
#define FV_END (FV_XS + 4)
static_assert(F_V + FV_END <= FQ_RED, "pass F shared memory");

// ---- passes A1 / D1: ring of 3 slots ----
#define A_MU 0                               // u panels of the stage factor: Luu, Lxu
#define A_H (A_MU + LP_UPANELS)              // Hq | Hf (180)
#define A_LU (A_H + 180)                     // lu of this direction
#define A_RB (A_LU + NU)
#define A_XS (A_RB + NX)                     // time-opt: stage record extension
#define A_XV (A_XS + SRX_SZ)                 // time-opt: V_X block
#define A_SLOT (A_XV + VX_SZ)
#define A_DEPTH 3
#define AV_BASE (A_DEPTH * A_SLOT)
#define AV_DX 0                              // [2][42]
#define AV_DZ (AV_DX + 2 * NX)               // 72
#define AV_X (AV_DZ + NZ)                    // time-opt scalar: d tau
#define AV_RS (AV_X + 2)                     // 32  right-hand side of the blocked back substitution
#define AV_XQ (AV_RS + 32)                   // 32  its solution
static_assert(AV_BASE + AV_XQ + 32 <= FQ_RED, "pass A1 shared memory");

// ---- pass A2: ring of 4 slots ----
#define B_C 0                                // constraint part of the stage record: SR_CF .. SR_X (1042)
#define SR_CSZ (SR_X - SR_CF)
#define B_DZ (B_C + SR_CSZ)                  // DZ RD RM (240)
#define B_RD (B_DZ + NZ)
#define B_RM (B_RD + 2 * NG)
#define B_LAM (B_RM + 2 * NG)
#define B_T (B_LAM + 2 * NG)
#define B_XV (B_T + 2 * NG)
#define B_XQ (B_XV + VX_SZ)
#define B_SLOT (B_XQ + QX_SZ)
#define B_DEPTH 4
#define BV_BASE (B_DEPTH * B_SLOT)
static_assert(BV_BASE + NG <= FQ_RED, "pass A2 shared memory");

// ---- pass C1: ring of 3 slots ----
#define C_H 0
#define C_C (C_H + 180)
#define C_RG (C_C + SR_CSZ + SRX_SZ)         // RG PRB (114); the constraint part is followed by the time-opt extension
#define C_PRB (C_RG + NZ)
#define C_DT (C_PRB + NX)                    // DT DLAM (168)
#define C_DLAM (C_DT + 2 * NG)
#define C_RD (C_DLAM + 2 * NG)
#define C_LAM (C_RD + 2 * NG)
#define C_T (C_LAM + 2 * NG)
#define C_XV (C_T + 2 * NG)
#define C_XQ (C_XV + VX_SZ)
#define C_SLOT (C_XQ + QX_SZ)
#define C_DEPTH 3
#define CV_BASE (C_DEPTH * C_SLOT)           // 6936
static_assert(CV_BASE + NG <= FQ_RED, "pass C1 shared memory");

// ---- pass C2: ring of 3 slots ----
#define E_MU 0
#define E_H (E_MU + LP_UPANELS)
#define E_M0 (E_H + 180)
#define E_XS (E_M0 + NZ)
#define E_XV (E_XS + SRX_SZ)
#define E_SLOT (E_XV + VX_SZ)
#define E_DEPTH 3
#define EV_BASE (E_DEPTH * E_SLOT)           // 6246
#define EV_M 0                               // 72
#define EV_PN (EV_M + NZ)                    // [2][42]
#define EV_LU (EV_PN + 2 * NX)               // 32
#define EV_PT (EV_LU + 32)                   // 42  time-opt: p+ after the elimination of dt
#define EV_X (EV_PT + NX)                    // time-opt scalar s
#define EV_RS (EV_X + 2)                     // 32  right-hand side of the blocked forward substitution
static_assert(EV_BASE + EV_RS + 32 <= FQ_RED, "pass C2 shared memory");

// ---- pass D2: ring of 2 slots ----
#define D_P 0
#define D_C (D_P + NX * P_LD)
#define D_DZ (D_C + SR_CSZ)
#define D_RD (D_DZ + NZ)
#define D_RM (D_RD + 2 * NG)
#define D_LAM (D_RM + 2 * NG)
#define D_T (D_LAM + 2 * NG)
#define D_PV (D_T + 2 * NG)
#define D_XV (D_PV + NX)
#define D_XQ (D_XV + VX_SZ)
#define D_SLOT (D_XQ + QX_SZ)
#define D_DEPTH 2
#define DV_BASE (D_DEPTH * D_SLOT)
static_assert(DV_BASE + NG <= FQ_RED, "pass D2 shared memory");

__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

__device__ __noinline__ void fq_mbar_wait(uint32_t slot, uint32_t parity) { mbar_wait(&fq_bars[slot], parity); }

__device__ __forceinline__ void fq_ring_wait(int slot, uint32_t& ph)
{
    fq_mbar_wait(slot, (ph >> slot) & 1u);
    ph ^= (1u << slot);
}

// end of a pass: this thread's global and shared writes become visible to the bulk copies of the next pass
__device__ __forceinline__ void fq_pass_end()
{
    fence_proxy_async_all();
    __syncthreads();
}

#ifdef NMPC_TIMING
__device__ unsigned long long g_fq_cycles[32];
__shared__ unsigned long long fq_tacc[32];
__shared__ long long fq_tprev;
#define FMARK(ph_) do { if (threadIdx.x == 0) { const long long t_ = clock64(); fq_tacc[ph_] += (unsigned long long)(t_ - fq_tprev); fq_tprev = t_; } } while (0)
#else
#define FMARK(ph_) do { } while (0)
#endif

// =============================================================================================
// pass R: update with the previous step, residuals of every stage, norms
// out: [0] |rg| [1] |rb| [2] |rd| [3] |rm| [4] sum lam t [5] number of constraint sides
// =============================================================================================
template <bool TOPT>
__device__ __noinline__ void fq_issue_R(int k, int slot, bool upd)
{
    const FqCtx& c = fq_ctx;
    double* sl = fq_smem + slot * R_SLOT;
    uint64_t* bar = &fq_bars[slot];
    uint32_t bytes = (SR_SZ + NZ + 4 * NG) * 8;
    if (upd) bytes += (NZ + 4 * NG) * 8;
    if (k > 0) bytes += (upd ? 2 : 1) * NX * 8;
    if (TOPT && k < c.N) bytes += (QX_SZ + VX_SZ) * 8;
    mbar_arrive_expect_tx(bar, bytes);
    if (TOPT && k < c.N) {
        bulk_g2s(sl + R_XQ, c.qx + (size_t)k * QX_SZ, QX_SZ * 8, bar);
        bulk_g2s(sl + R_XV, c.vb + (size_t)k * V_STAGE + V_X, VX_SZ * 8, bar);
    }
    bulk_g2s(sl + R_SR, c.SRb + (size_t)k * SR_SZ, SR_SZ * 8, bar);
    bulk_g2s(sl + R_Z, c.qz + (size_t)k * NZ, NZ * 8, bar);
    bulk_g2s(sl + R_LAM, c.qlam + (size_t)k * 2 * NG, 2 * NG * 8, bar);
    bulk_g2s(sl + R_T, c.qt + (size_t)k * 2 * NG, 2 * NG * 8, bar);
    if (k > 0) bulk_g2s(sl + R_PI, c.qpi + (size_t)(k - 1) * NX, NX * 8, bar);
    if (upd) {
        const double* vk = c.vb + (size_t)k * V_STAGE;
        bulk_g2s(sl + R_DZ, vk + V_DZ, NZ * 8, bar);
        bulk_g2s(sl + R_DT, vk + V_DT, 4 * NG * 8, bar);
        if (k > 0) bulk_g2s(sl + R_DPI, vk - V_STAGE + V_DPI, NX * 8, bar);
    }
}

template <bool TOPT>
__device__ __noinline__ uint32_t fq_pass_R(uint32_t ph, int it, int warm, double alpha_p, double alpha_d, double* out)
{
    const int tid = threadIdx.x;
    const int N = fq_ctx.N;
    double* rv = fq_smem + RV_BASE;
    double* red = fq_smem + FQ_RED;
    const bool upd = it > 0;
    double n_g = 0.0, n_b = 0.0, n_d = 0.0, n_m = 0.0, s_m = 0.0;
    int cnt = 0;
    if (tid == FQ_ISSUER)
        for (int i = 0; i < R_DEPTH - 1 && i <= N; i++) fq_issue_R<TOPT>(N - i, i % R_DEPTH, upd);
    for (int i = 0; i <= N; i++) {
        const int k = N - i, slot = i % R_DEPTH;
        const bool term = (k == N);
        if (tid == FQ_ISSUER && i + R_DEPTH - 1 <= N) fq_issue_R<TOPT>(k - (R_DEPTH - 1), (i + R_DEPTH - 1) % R_DEPTH, upd);
        FMARK(1);
        fq_ring_wait(slot, ph);
        FMARK(18);
        const double* sl = fq_smem + slot * R_SLOT;
        const double* S = sl + R_SR;
        const double dt = TOPT ? S[SR_X + SRX_DT] : fq_ctx.dt, c2 = TOPT ? S[SR_X + SRX_C2] : fq_ctx.c2;
        double* vk = fq_ctx.vb + (size_t)k * V_STAGE;
        double* zc = rv + RV_ZC + (k & 1) * NZ;
        // time-opt: thread 127 carries the node duration tau = dt step and its two bounds through the steps
        const bool xthr = TOPT && !term && tid == FQ_THREADS - 1;
        double x_tau = 0.0;
        const double* zxn = rv + RV_ZC + ((k + 1) & 1) * NZ + NU;
        double* pik = rv + RV_PIN + (k & 1) * NX;
        const double* pin = rv + RV_PIN + ((k + 1) & 1) * NX;
        // ---- step 1: the updated primal iterate and dynamics multiplier of the stage ----
        if (tid < NZ) {
            double z = sl[R_Z + tid];
            if (upd) { z = fma(alpha_p, sl[R_DZ + tid], z); fq_ctx.qz[(size_t)k * NZ + tid] = z; }
            zc[tid] = z;
        } else if (tid >= 80 && tid < 80 + NX && k > 0) {
            const int l = tid - 80;
            double p = sl[R_PI + l];
            if (upd) { p = fma(alpha_d, sl[R_DPI + l], p); fq_ctx.qpi[(size_t)(k - 1) * NX + l] = p; }
            pik[l] = p;
        }
        if (xthr) {
            x_tau = sl[R_XQ];
            if (upd) { x_tau = fma(alpha_p, sl[R_XV + VX_DTAU], x_tau); fq_ctx.qx[(size_t)k * QX_SZ] = x_tau; }
            rv[RV_X] = x_tau;
        }
        __syncthreads();
        FMARK(19);
        // ---- step 2: C z; dynamics residual ----
        fq_cz(S, zc, rv + RV_CZ, tid);
        if (!term && tid >= 80 && tid < 80 + NX) {
            const int l = tid - 80;
            double rb = fq_g(S, zc, l, dt, c2) + S[SR_B + l] - zxn[l];
            if (TOPT) rb = fma(S[SR_X + SRX_B + l], rv[RV_X], rb);
            vk[V_RB + l] = rb;
            n_b = nanmax(n_b, fabs(rb));
        }
        __syncthreads();
        FMARK(20);
        // ---- step 3: both sides of constraint row j: slacks, multipliers, rd, rm, Gamma, gamma ----
        if (tid < NG) {
            const int j = tid;
            const double cz = rv[RV_CZ + j];
            double Gam = 0.0, gam = 0.0, dl = 0.0;
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int row = e * NG + j;
                const bool msk = S[SR_D + 2 * NG + row] > 0.5;
                double t = sl[R_T + row], lam = sl[R_LAM + row];
                if (upd && msk) {
                    t = fmax(fma(alpha_p, sl[R_DT + row], t), IPM_T_MIN);
                    lam = fmax(fma(alpha_d, sl[R_DT + 2 * NG + row], lam), IPM_LAM_MIN);
                    fq_ctx.qt[(size_t)k * 2 * NG + row] = t; fq_ctx.qlam[(size_t)k * 2 * NG + row] = lam;
                }
                const double slack = e ? (S[SR_D + NG + j] - cz) : (cz - S[SR_D + j]);
                if (it == 0) {               // HPIPM INIT_VAR: levels 0 / 1 from the slacks, level 2 keeps lam, t clipped at thr0
                    if (msk) {
                        if (warm < 2 || !(t > 0.0) || !(lam > 0.0)) { t = fmax(slack, IPM_THR0); lam = IPM_MU0 * __drcp_rn(t); }
                        else { t = fmax(t, IPM_THR0); lam = fmax(lam, IPM_THR0); }
                    } else { t = 1.0; lam = 0.0; }
                    fq_ctx.qt[(size_t)k * 2 * NG + row] = t; fq_ctx.qlam[(size_t)k * 2 * NG + row] = lam;
                }
                const double rd = msk ? (slack - t) : 0.0;
                const double rm = msk ? lam * t : 0.0;
                vk[V_RD + row] = rd; vk[V_RM + row] = rm;
                s_m += rm; cnt += msk ? 1 : 0;
                n_d = nanmax(n_d, fabs(rd));
                n_m = nanmax(n_m, fabs(rm));
                if (msk) {
                    const double ti = __drcp_rn(t);          // one reciprocal per side instead of two divisions
                    const double g = (rm + lam * rd) * ti;
                    Gam += lam * ti;
                    if (e) { gam -= g; dl -= lam; } else { gam += g; dl += lam; }
                }
            }
            vk[V_GAM + j] = Gam;
            rv[RV_GAMV + j] = gam; rv[RV_CZ + j] = dl;
        }
        if (xthr) {          // the two bounds of dt: lb <= tau <= ub
            double Gam = 0.0, gam = 0.0, dl = 0.0;
#pragma unroll
            for (int e = 0; e < 2; e++) {
                double t = sl[R_XQ + 4 + e], lam = sl[R_XQ + 2 + e];
                if (upd) {
                    t = fmax(fma(alpha_p, sl[R_XV + VX_DT + e], t), IPM_T_MIN);
                    lam = fmax(fma(alpha_d, sl[R_XV + VX_DL + e], lam), IPM_LAM_MIN);
                }
                const double slack = e ? (S[SR_X + SRX_UB] - x_tau) : (x_tau - S[SR_X + SRX_LB]);
                if (it == 0) {
                    if (warm < 2 || !(t > 0.0) || !(lam > 0.0)) { t = fmax(slack, IPM_THR0); lam = IPM_MU0 * __drcp_rn(t); }
                    else { t = fmax(t, IPM_THR0); lam = fmax(lam, IPM_THR0); }
                }
                if (upd || it == 0) { fq_ctx.qx[(size_t)k * QX_SZ + 4 + e] = t; fq_ctx.qx[(size_t)k * QX_SZ + 2 + e] = lam; }
                const double rd = slack - t, rm = lam * t;
                vk[V_X + VX_RD + e] = rd; vk[V_X + VX_RM + e] = rm;
                s_m += rm; cnt += 1;
                n_d = nanmax(n_d, fabs(rd));
                n_m = nanmax(n_m, fabs(rm));
                const double ti = __drcp_rn(t);
                const double g = (rm + lam * rd) * ti;
                Gam += lam * ti;
                if (e) { gam -= g; dl -= lam; } else { gam += g; dl += lam; }
            }
            vk[V_X + VX_GAM] = Gam;
            rv[RV_X + 1] = dl; rv[RV_X + 2] = gam;
        }
        __syncthreads();
        FMARK(21);
        // ---- step 4: stationarity residual and the linear term q~ = rg + C' gamma ----
        if (tid < NZ) {
            const int c = tid;
            double s = S[SR_GR + c] + fq_h(S, zc, c);
            if (c >= NU && k > 0) s -= pik[c - NU];
            double cdl, qp;
            fq_ct2(S, rv + RV_CZ, rv + RV_GAMV, c, cdl, qp);
            s -= cdl;
            if (!term) s += fq_gt(S, pin, c, dt, c2);
            if ((k == 0 && c >= NU) || (term && c < NU)) s = 0.0;
            vk[V_RG + c] = s; vk[V_QT + c] = qp + s;
            n_g = nanmax(n_g, fabs(s));
        }
        if (xthr) {          // stationarity in dt: g + h tau + b' pi_k - (lam_l - lam_u)
            double s0 = fma(S[SR_X + SRX_H], x_tau, S[SR_X + SRX_G]) - rv[RV_X + 1], s1 = 0.0;
#pragma unroll
            for (int l = 0; l < NX; l += 2) { s0 = fma(S[SR_X + SRX_B + l], pin[l], s0); s1 = fma(S[SR_X + SRX_B + l + 1], pin[l + 1], s1); }
            const double sg = s0 + s1;
            vk[V_X + VX_RG] = sg; vk[V_X + VX_QT] = sg + rv[RV_X + 2];
            n_g = nanmax(n_g, fabs(sg));
        }
        __syncthreads();
        FMARK(22);
    }
    out[0] = fq_block_max(n_g, red, tid);
    out[1] = fq_block_max(n_b, red, tid);
    out[2] = fq_block_max(n_d, red, tid);
    out[3] = fq_block_max(n_m, red, tid);
    out[4] = fq_block_sum(s_m, red, tid);
    out[5] = (it == 0) ? fq_block_sum((double)cnt, red, tid) : 0.0;
    fq_pass_end();
    return ph;
}

// =============================================================================================
// pass F: Riccati factorisation sweep (classical recursion on P), k = N..0
// =============================================================================================
template <bool TOPT>
__device__ __noinline__ void fq_issue_F(int k)
{
    const FqCtx& c = fq_ctx;
    uint64_t* bar = &fq_bars[0];
    mbar_arrive_expect_tx(bar, (SR_SZ + 186 + (TOPT ? VX_SZ : 0)) * 8);
    bulk_g2s(fq_smem + F_S, c.SRb + (size_t)k * SR_SZ, SR_SZ * 8, bar);
    bulk_g2s(fq_smem + F_SV, c.vb + (size_t)k * V_STAGE + V_QT, 186 * 8, bar);
    if (TOPT) bulk_g2s(fq_smem + F_XV, c.vb + (size_t)k * V_STAGE + V_X, VX_SZ * 8, bar);
}

template <bool TOPT>
__device__ __noinline__ uint32_t fq_pass_F(uint32_t ph)
{
    const int tid = threadIdx.x;
    const int N = fq_ctx.N;
    double* sP = fq_smem + F_P;
    double* sM = fq_smem + F_M;
    double* sU = sM;
    const double* S = fq_smem + F_S;
    const double* Sv = fq_smem + F_SV;           // [V_QT 72 | V_GAM 42 | V_KFF 30 | V_RB 42] of the stage
    double* fv = fq_smem + F_V;

    // tile ownership, packed into one register (unpacked at each use: nothing but `tcode` stays live across a stage)
    unsigned tcode = 0;
    if (c_utile[tid][0] != 7)
        tcode |= c_utile[tid][0] | (c_utile[tid][1] << 3) | (c_utile[tid][2] << 4) | (c_utile[tid][3] << 9) | (1u << 24);
    if (c_xtile[tid][0] != 3)
        tcode |= (c_xtile[tid][0] << 12) | (c_xtile[tid][1] << 14) | (c_xtile[tid][2] << 16) | (c_xtile[tid][3] << 20) | (1u << 25);
    const int free_ord = (c_xtile[tid][0] == 3) ? c_xtile[tid][2] : -1;      // ordinal among the threads without a state tile
#define FQ_UNPACK_U                                                                                              \
    unsigned tc_ = tcode; asm volatile("" : "+r"(tc_));                                                          \
    const bool has_ut = (tc_ >> 24) & 1u;                                                                        \
    const int urt = tc_ & 7u, uct = (tc_ >> 3) & 1u, urb = (tc_ >> 4) & 31u, ucb = (tc_ >> 9) & 7u;              \
    const int ur0 = 3 * urb, uc0 = 6 * ucb;                                                                      \
    (void)urt; (void)uct; (void)ur0; (void)uc0; (void)has_ut
#define FQ_UNPACK_X                                                                                              \
    unsigned tx_ = tcode; asm volatile("" : "+r"(tx_));                                                          \
    const bool has_xt = (tx_ >> 25) & 1u;                                                                        \
    const int xrt = (tx_ >> 12) & 3u, xct = (tx_ >> 14) & 3u, xr = (tx_ >> 16) & 15u, xc = (tx_ >> 20) & 15u;    \
    (void)xrt; (void)xct; (void)xr; (void)xc; (void)has_xt

    if (tid == 0) fq_issue_F<TOPT>(N);
    for (int k = N; k >= 0; k--) {
        const bool term = (k == N);
        double* vk = fq_ctx.vb + (size_t)k * V_STAGE;
        fq_ring_wait(0, ph);                                 // SR_k and the stage vectors have landed
        FMARK(2);
        double xa[3][3], ua[3][6];
        const double dt = TOPT ? S[SR_X + SRX_DT] : fq_ctx.dt, c2 = TOPT ? S[SR_X + SRX_C2] : fq_ctx.c2;
        if (tid == 0) bulk_store_wait_read();                // the previous factor record has left sM / sP
        if (!term) {
            // P+ rb (kept for the corrector) and w = P+ rb + p+.
            // Time-opt: the node duration is eliminated from the stage before anything else.  With b = d x+ / d dt,
            // d = h_tt + Gamma_t + b' P+ b and s = q~_t + b' p+, the rest of the stage sees the cost-to-go
            // P~ = P+ - (P+ b)(P+ b)' / d,  p~ = p+ - (P+ b) s / d, and dt follows in the forward sweeps from
            // tau = -(s + (P+ b)' (A dx + B du + rb)) / d.  One sweep over the rows of P+ gives P+ b and P+ rb;
            // P~ rb = P+ rb - (P+ b) ((P+ b)' rb) / d needs no second one.
            const double* bc = S + SR_X + SRX_B;
            double prb = 0.0, pb = 0.0;
            if (tid < NX) {
                const double* row = sP + tid * P_LD;
                double s0 = 0.0, s1 = 0.0, b0 = 0.0, b1 = 0.0;
#pragma unroll
                for (int j = 0; j < NX; j += 2) {
                    s0 += row[j] * Sv[V_RB + j]; s1 += row[j + 1] * Sv[V_RB + j + 1];
                    if (TOPT) { b0 = fma(row[j], bc[j], b0); b1 = fma(row[j + 1], bc[j + 1], b1); }
                }
                prb = s0 + s1;
                if (TOPT) { pb = b0 + b1; fv[FV_PB + tid] = pb; vk[V_X + VX_PB + tid] = pb; }
            }
            if (TOPT) {
                __syncthreads();                             // also orders wait_read before the writes to sP
                if (tid < 32) {
                    double d0 = 0.0, s0 = 0.0, t0 = 0.0;
                    for (int j = tid; j < NX; j += 32) {
                        d0 = fma(bc[j], fv[FV_PB + j], d0); s0 = fma(bc[j], fv[FV_PN + j], s0);
                        t0 = fma(Sv[V_RB + j], fv[FV_PB + j], t0);
                    }
                    d0 = warp_sum(d0); s0 = warp_sum(s0); t0 = warp_sum(t0);
                    if (tid == 0) {
                        const double d = S[SR_X + SRX_H] + fq_smem[F_XV + VX_GAM] + IPM_REG_PRIM + d0;
                        const double dinv = d > 0.0 ? 1.0 / d : 0.0;
                        const double sx = fq_smem[F_XV + VX_QT] + s0;
                        fv[FV_XS] = dinv; fv[FV_XS + 1] = sx; fv[FV_XS + 2] = t0;
                        vk[V_X + VX_DINV] = dinv; vk[V_X + VX_S] = sx;
                    }
                }
                __syncthreads();
                const double dinv = fv[FV_XS], sx = fv[FV_XS + 1], t0 = fv[FV_XS + 2];
                for (int e = tid; e < NX * NX; e += FQ_THREADS) {
                    const int ii = e / NX, jj = e % NX;
                    sP[ii * P_LD + jj] = fma(-fv[FV_PB + ii] * dinv, fv[FV_PB + jj], sP[ii * P_LD + jj]);
                }
                if (tid < NX) {
                    fv[FV_PN + tid] = fma(-pb, sx * dinv, fv[FV_PN + tid]);
                    prb = fma(-pb, t0 * dinv, prb);
                }
            }
            if (tid < NX) {
                vk[V_PRB + tid] = prb;
                fv[FV_W + tid] = prb + fv[FV_PN + tid];
            }
            __syncthreads();                                 // also orders wait_read before the writes to sU
            if (tid < 4 * NU) {                              // U = P+[:, h] [Hf | Hq]
                const int j = tid % NU, g = tid / NU;
                const double* hc = (j < 12) ? (S + SR_HF + j) : (S + SR_HQ + (j - 12));
                const int ld = (j < 12) ? 12 : NQ;
                double hcv[6];
#pragma unroll
                for (int h = 0; h < 6; h++) hcv[h] = hc[ld * h];
                for (int r = g; r < NX; r += 4) {
                    const double* ph_ = sP + r * P_LD + 2 * NQ;
                    double s = 0.0;
#pragma unroll
                    for (int h = 0; h < 6; h++) s += ph_[h] * hcv[h];
                    sU[r * NU + j] = s;
                }
            }
            if (tid < NZ) fv[FV_QT + tid] = Sv[V_QT + tid] + fq_gt(S, fv + FV_W, tid, dt, c2);     // m = q~ + G' w
            __syncthreads();
            FMARK(4);
            // ---- u tile: rows ur0..+2 (z index), columns uc0..+5 (u index) ----
            FQ_UNPACK_U;
            const int uti = ur0 - (urt == 0 ? IA : urt == 1 ? IFRC : urt == 2 ? IQ : urt == 3 ? IV : IH);
            const int utj = uc0 - (uct == 0 ? IA : IFRC);
            if (has_ut) {
                switch (uct * 5 + urt) {
                case 0: fq_asm_u<0, 0>(sP, sU, S, uti, utj, dt, c2, ua); break;
                case 1: fq_asm_u<1, 0>(sP, sU, S, uti, utj, dt, c2, ua); break;
                case 2: fq_asm_u<2, 0>(sP, sU, S, uti, utj, dt, c2, ua); break;
                case 3: fq_asm_u<3, 0>(sP, sU, S, uti, utj, dt, c2, ua); break;
                case 4: fq_asm_u<4, 0>(sP, sU, S, uti, utj, dt, c2, ua); break;
                case 6: fq_asm_u<1, 1>(sP, sU, S, uti, utj, dt, c2, ua); break;
                case 7: fq_asm_u<2, 1>(sP, sU, S, uti, utj, dt, c2, ua); break;
                case 8: fq_asm_u<3, 1>(sP, sU, S, uti, utj, dt, c2, ua); break;
                default: fq_asm_u<4, 1>(sP, sU, S, uti, utj, dt, c2, ua); break;
                }
                if (urb / 2 == ucb) {                        // diagonal tile: Hessian diagonal + regularisation
#pragma unroll
                    for (int r = 0; r < 3; r++)
#pragma unroll
                        for (int c = 0; c < 6; c++)
                            if (ur0 + r == uc0 + c) ua[r][c] += S[SR_RD + ur0 + r] + IPM_REG_PRIM;
                }
                if (ur0 >= IFRC && ur0 < IQ && uc0 >= IFRC) {    // friction rows: 3x3 block of the foot
                    const int ft = (ur0 - IFRC) / 3;
                    const int cofs = IFRC + 3 * ft - uc0;    // first column of the foot inside this tile
                    if (cofs >= 0 && cofs < 6) {
#pragma unroll 1
                        for (int r5 = 0; r5 < 5; r5++) {                   // rolled: a short body that stays in the instruction cache
                            const double g = Sv[V_GAM + 5 * ft + r5];
                            const double* cf = S + SR_CF + 3 * (5 * ft + r5);
#pragma unroll
                            for (int r = 0; r < 3; r++)
#pragma unroll
                                for (int c = 0; c < 3; c++) {
                                    const double add = g * cf[r] * cf[c];
                                    if (cofs == 0) ua[r][c] += add; else ua[r][3 + c] += add;
                                }
                        }
                    }
                }
            }
        } else {
            if (tid < NZ) fv[FV_QT + tid] = Sv[V_QT + tid];
            __syncthreads();                                 // orders wait_read before the writes to sP below
        }
        // ---- x tile: rows 3xr..+2, columns 3xc..+2 of the state block ----
        {
            FQ_UNPACK_X;
            const int xti = 3 * xr - (xrt == 0 ? 0 : xrt == 1 ? NQ : 2 * NQ);
            const int xtj = 3 * xc - (xct == 0 ? 0 : xct == 1 ? NQ : 2 * NQ);
            if (has_xt) {
                const int i0 = 3 * xr, j0x = 3 * xc;
                if (term) {
#pragma unroll
                    for (int r = 0; r < 3; r++)
#pragma unroll
                        for (int c = 0; c < 3; c++) xa[r][c] = 0.0;
                } else {
                    switch (xrt * 3 + xct) {
                    case 0: fq_asm_x<0, 0>(sP, sU, S, xti, xtj, dt, xa); break;
                    case 3: fq_asm_x<1, 0>(sP, sU, S, xti, xtj, dt, xa); break;
                    case 4: fq_asm_x<1, 1>(sP, sU, S, xti, xtj, dt, xa); break;
                    case 6: fq_asm_x<2, 0>(sP, sU, S, xti, xtj, dt, xa); break;
                    case 7: fq_asm_x<2, 1>(sP, sU, S, xti, xtj, dt, xa); break;
                    default: fq_asm_x<2, 2>(sP, sU, S, xti, xtj, dt, xa); break;
                    }
                }
#pragma unroll
                for (int r = 0; r < 3; r++)
#pragma unroll
                    for (int c = 0; c < 3; c++) {
                        const int ii = i0 + r, jj = j0x + c;
                        double v = 0.0;
                        if (xrt == 0) v = S[SR_QQ + NQ * ii + jj];
                        else if (ii == jj) v = (xrt == 1) ? S[SR_QV + ii - NQ] : S[SR_QH + ii - 2 * NQ];
                        if (!term && ii == jj) v += IPM_REG_PRIM;
                        xa[r][c] += v;
                    }
                if (i0 < 2 * NQ) {                           // (q,v) x (q,v): all 22 state rows
                    for (int g = 0; g < NGX; g++) {
                        const double gm = Sv[V_GAM + NGF + g];
                        if (gm == 0.0) continue;
                        const double* row = S + SR_CX + SR_CXLD * g;
                        const double a0 = row[i0] * gm, a1 = row[i0 + 1] * gm, a2 = row[i0 + 2] * gm;
                        const double b0 = row[j0x], b1 = row[j0x + 1], b2 = row[j0x + 2];
                        xa[0][0] += a0 * b0; xa[0][1] += a0 * b1; xa[0][2] += a0 * b2;
                        xa[1][0] += a1 * b0; xa[1][1] += a1 * b1; xa[1][2] += a1 * b2;
                        xa[2][0] += a2 * b0; xa[2][1] += a2 * b1; xa[2][2] += a2 * b2;
                    }
                } else {                                     // h rows: only the centroidal rows reach them
#pragma unroll
                    for (int r = 0; r < 3; r++) {
                        const int hi = i0 + r - 2 * NQ;
                        const double gm = Sv[V_GAM + GR_CONS + hi];
#pragma unroll
                        for (int c = 0; c < 3; c++) {
                            const int jj = j0x + c;
                            if (jj < 2 * NQ) xa[r][c] += gm * S[SR_CX + SR_CXLD * (16 + hi) + jj];
                            else if (jj == i0 + r) xa[r][c] += gm;
                        }
                    }
                }
            }
        }
        __syncthreads();                 // every read of SR_k, the stage vectors, P+, U is done
        FMARK(5);
        if (k > 0 && tid == 0) fq_issue_F<TOPT>(k - 1);            // prefetch the next stage under the factorisation
        if (!term) {
            // ---- right-looking Cholesky of the 30 input columns; x tiles take the same rank-6 updates ----
            for (int jb = 0; jb < 5; jb++) {
                const int j0 = PB * jb;
                double* pan = sM + P_OFF(jb);                // rows j0..72 of columns j0..j0+5
                {
                    FQ_UNPACK_U;
                    if (has_ut && ucb == jb) {
#pragma unroll
                        for (int r = 0; r < 3; r++)
#pragma unroll
                            for (int c = 0; c < 6; c++) pan[(ur0 + r - j0) * PB + c] = ua[r][c];
                    }
                }
                if (tid < PB) pan[(NZ - j0) * PB + tid] = fv[FV_QT + j0 + tid];
                __syncthreads();
                // Panel factorisation, right-looking in registers: every row thread factors the 6x6 diagonal block
                // redundantly and carries its own row along as a seventh row, so the dependent chain per pivot is one
                // rsqrt, one multiply and one fused multiply-add.  Non-positive pivots zero their column (BLASFEO
                // potrf rule).
                const int nrow = NZ + 1 - j0;
                double drow[PB] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
                if (tid < nrow) {
                    double A[PB][PB], rv[PB];
#pragma unroll
                    for (int aa = 0; aa < PB; aa++)
#pragma unroll
                        for (int c = 0; c <= aa; c++) A[aa][c] = pan[aa * PB + c];
                    load6(pan + tid * PB, rv);
#pragma unroll
                    for (int j = 0; j < PB; j++) {
                        const double v = A[j][j];
                        const double inv = (v > 0.0) ? rsqrt(v) : 0.0;
                        A[j][j] = inv;                       // the sweeps only ever divide by the diagonal: keep 1 / L_jj
#pragma unroll
                        for (int i = j + 1; i < PB; i++) A[i][j] *= inv;
                        rv[j] *= inv;
#pragma unroll
                        for (int c = j + 1; c < PB; c++) {
#pragma unroll
                            for (int i = c; i < PB; i++) A[i][c] -= A[i][j] * A[c][j];
                            rv[c] -= rv[j] * A[c][j];
                        }
                    }
                    if (tid >= PB) {
#pragma unroll
                        for (int jj = 0; jj < PB; jj++) pan[tid * PB + jj] = rv[jj];
                    } else {
#pragma unroll
                        for (int aa = 0; aa < PB; aa++)
#pragma unroll
                            for (int c = 0; c <= aa; c++) if (tid == aa) drow[c] = A[aa][c];
                    }
                }
                __syncthreads();
                if (tid < PB) {
#pragma unroll
                    for (int c = 0; c < PB; c++) pan[tid * PB + c] = drow[c];
                }
                FQ_UNPACK_U;
                FQ_UNPACK_X;
                if (has_ut && ucb > jb) {
                    const double* pa = pan + (ur0 - j0) * PB;
                    const double* pb = pan + (uc0 - j0) * PB;
#pragma unroll
                    for (int kk = 0; kk < PB; kk++) {
                        const double a0 = pa[kk], a1 = pa[PB + kk], a2 = pa[2 * PB + kk];
#pragma unroll
                        for (int c = 0; c < 6; c++) {
                            const double bv = pb[c * PB + kk];
                            ua[0][c] -= a0 * bv; ua[1][c] -= a1 * bv; ua[2][c] -= a2 * bv;
                        }
                    }
                }
                if (has_xt) {
                    const double* pa = pan + (NU + 3 * xr - j0) * PB;
                    const double* pb = pan + (NU + 3 * xc - j0) * PB;
                    double av[3][PB], bv[3][PB];
#pragma unroll
                    for (int r = 0; r < 3; r++) { load6(pa + r * PB, av[r]); load6(pb + r * PB, bv[r]); }
#pragma unroll
                    for (int r = 0; r < 3; r++)
#pragma unroll
                        for (int c = 0; c < 3; c++)
#pragma unroll
                            for (int kk = 0; kk < PB; kk++) xa[r][c] -= av[r][kk] * bv[c][kk];
                }
                if (tid < NZ - j0 - PB) {                    // row 72: m_c -= lu . L(c, j0..j0+5)
                    const double* pa = pan + (NZ - j0) * PB;
                    const double* pb = pan + (PB + tid) * PB;
                    double av[6], bv[6];
                    load6(pa, av);
                    load6(pb, bv);
                    double sm_ = fv[FV_QT + j0 + PB + tid];
#pragma unroll
                    for (int kk = 0; kk < PB; kk++) sm_ -= av[kk] * bv[kk];
                    fv[FV_QT + j0 + PB + tid] = sm_;
                }
                // the next panel's owners write a different panel; the m row is ordered by the next barrier
            }
        }
        __syncthreads();
        FMARK(6);
        // ---- P_k (both triangles) -> sP, p_k -> fv ----
        {
            FQ_UNPACK_X;
            if (has_xt) {
#pragma unroll
                for (int r = 0; r < 3; r++)
#pragma unroll
                    for (int c = 0; c < 3; c++) {
                        const int ii = 3 * xr + r, jj = 3 * xc + c;
                        if (ii >= jj) { sP[ii * P_LD + jj] = xa[r][c]; sP[jj * P_LD + ii] = xa[r][c]; }
                    }
            }
        }
        if (tid < NX) fv[FV_PN + tid] = fv[FV_QT + NU + tid];
        if (!term && free_ord >= 0) {
            // Inverses of the five 6 x 6 diagonal blocks of Luu, for the blocked triangular solves of passes A1 / D1 /
            // C2 (5 block steps instead of 30 scalar steps on the dependent chain).  The diagonal already holds
            // 1 / L_jj; the strictly lower part of the inverse goes, transposed, into the unused upper triangle of the
            // block, so nothing that is still being read is overwritten.  Done by the threads that own no state tile.
            for (int col = free_ord; col < NU; col += FQ_THREADS - N_XTILES) {
                double* blk = sM + P_OFF(col / PB);
                const int cc = col % PB;
                double x[PB];
#pragma unroll
                for (int r = 0; r < PB; r++) {
                    double sacc = (r == cc) ? 1.0 : 0.0;
#pragma unroll
                    for (int m = 0; m < r; m++) sacc = fma(-blk[r * PB + m], x[m], sacc);
                    x[r] = sacc * blk[r * PB + r];
                }
#pragma unroll
                for (int r = 1; r < PB; r++)
                    if (r > cc) blk[cc * PB + r] = x[r];
            }
        }
        if (!term && tid < NU) vk[V_KFF + tid] = sM[lp_idx(NZ, tid)];     // lu of the affine direction, for pass A1
        fence_proxy_async();
        __syncthreads();
        if (tid == 0) {
            double* fr = fq_ctx.FRb + (size_t)k * FR_SZ;
            if (!term) bulk_s2g(fr + FR_MU, sM, LP_UPANELS * 8);
            bulk_s2g(fr + FR_P, sP, NX * P_LD * 8);
        }
        FMARK(7);
    }
    if (tid == 0) bulk_store_wait_all();                     // the factor records are in global memory
    fq_pass_end();
    return ph;
#undef FQ_UNPACK_U
#undef FQ_UNPACK_X
}

// =============================================================================================
// passes A1 / D1: forward chain  du = -Luu^-T (lu + Lxu' dx),  dx+ = G [du; dx] + rb
// =============================================================================================
template <bool TOPT>
__device__ __noinline__ void fq_issue_A(int k, int slot)
{
    const FqCtx& c = fq_ctx;
    double* sl = fq_smem + slot * A_SLOT;
    uint64_t* bar = &fq_bars[slot];
    mbar_arrive_expect_tx(bar, (LP_UPANELS + 180 + NU + NX + (TOPT ? SRX_SZ + VX_SZ : 0)) * 8);
    if (TOPT) {
        bulk_g2s(sl + A_XS, c.SRb + (size_t)k * SR_SZ + SR_X, SRX_SZ * 8, bar);
        bulk_g2s(sl + A_XV, c.vb + (size_t)k * V_STAGE + V_X, VX_SZ * 8, bar);
    }
    bulk_g2s(sl + A_MU, c.FRb + (size_t)k * FR_SZ + FR_MU, LP_UPANELS * 8, bar);
    bulk_g2s(sl + A_H, c.SRb + (size_t)k * SR_SZ + SR_HQ, 180 * 8, bar);
    bulk_g2s(sl + A_LU, c.vb + (size_t)k * V_STAGE + V_KFF, (NU + NX) * 8, bar);
}

template <bool TOPT>
__device__ __noinline__ uint32_t fq_pass_A1(uint32_t ph)
{
    const int tid = threadIdx.x, lane = tid & 31;
    const int N = fq_ctx.N;
    double* av = fq_smem + AV_BASE;
    if (tid < NX) av[AV_DX + tid] = 0.0;                     // dx_0 = 0 (the initial state is fixed)
    if (tid == FQ_ISSUER)
        for (int i = 0; i < A_DEPTH - 1 && i < N; i++) fq_issue_A<TOPT>(i, i % A_DEPTH);
    __syncthreads();
    for (int k = 0; k <= N; k++) {
        const bool term = (k == N);
        const int slot = k % A_DEPTH;
        double* vk = fq_ctx.vb + (size_t)k * V_STAGE;
        const double* dx = av + AV_DX + (k & 1) * NX;
        double* dxn = av + AV_DX + ((k + 1) & 1) * NX;
        double* dz = av + AV_DZ;
        if (!term) {
            if (tid == FQ_ISSUER && k + A_DEPTH - 1 < N) fq_issue_A<TOPT>(k + A_DEPTH - 1, (k + A_DEPTH - 1) % A_DEPTH);
            FMARK(8);
            fq_ring_wait(slot, ph);
            FMARK(23);
            const double* sl = fq_smem + slot * A_SLOT;
            const double* sM = sl + A_MU;
            const double dt = TOPT ? sl[A_XS + SRX_DT] : fq_ctx.dt, c2 = TOPT ? sl[A_XS + SRX_C2] : fq_ctx.c2;
            if (tid < 32) {             // du = -Luu^-T (lu + Lxu' dx)
                double rhs = 0.0;
                if (lane < NU) {
                    double r1 = 0.0;
                    rhs = sl[A_LU + lane];
#pragma unroll
                    for (int l = 0; l < NX; l += 2) {
                        rhs = fma(sM[lp_idx(NU + l, lane)], dx[l], rhs);
                        r1 = fma(sM[lp_idx(NU + l + 1, lane)], dx[l + 1], r1);
                    }
                    rhs += r1;
                }
                // Back substitution with Luu', blocked by the 6 x 6 diagonal blocks whose inverses the factorisation left
                // in the panels: per block, 6 broadcasts + 6 multiply-adds solve the block, 6 + 6 more update the lanes
                // before it.  Row li of a stored diagonal block reads [L(li, < li) | 1 / L_ii | inverse(> li, li)], so
                // the coefficients of lane ll's block solve are one row with the leading entries masked.
                const int ll = lane < NU ? lane : NU - 1;
                const int lb = ll / PB, li = ll % PB;
                double cf[PB];
                load6(sM + P_OFF(lb) + li * PB, cf);
#pragma unroll
                for (int m = 0; m < PB; m++) cf[m] = (m >= li) ? cf[m] : 0.0;
                // L(i, ll) for i >= ll sits at colb + 6 i (panel-packed, 6 doubles per row)
                const double* colb = sM + P_OFF(lb) - PB * PB * lb + li;
                // The lanes exchange through shared memory (two 32-double arrays of this warp) rather than with shuffles:
                // a block step is 12 loads + 12 multiply-adds instead of 24 shuffles and their lane arithmetic, in a
                // rolled body that stays in the instruction cache.  Same operations in the same order as before.
                double* rs = av + AV_RS;
                double* xq = av + AV_XQ;
                rs[lane] = rhs;
                __syncwarp();
                FMARK(25);
#pragma unroll 1
                for (int bk = NU / PB - 1; bk >= 0; bk--) {
                    if (lb == bk) {                          // the lanes of block bk solve it
                        const double* r6 = rs + PB * bk;
                        double a0 = 0.0, a1 = 0.0;
#pragma unroll
                        for (int m = 0; m < PB; m += 2) { a0 = fma(cf[m], r6[m], a0); a1 = fma(cf[m + 1], r6[m + 1], a1); }
                        xq[lane] = a0 + a1;
                    }
                    __syncwarp();
                    if (lane < PB * bk) {                    // the lanes before it take the block out of their right-hand side
                        const double* x6 = xq + PB * bk;
                        const double* lc = colb + PB * PB * bk;
                        double u0 = 0.0, u1 = 0.0;
#pragma unroll
                        for (int m = 0; m < PB; m += 2) { u0 = fma(lc[PB * m], x6[m], u0); u1 = fma(lc[PB * (m + 1)], x6[m + 1], u1); }
                        rs[lane] -= u0 + u1;
                    }
                    __syncwarp();
                }
                const double xs = xq[lane < NU ? lane : NU - 1];
                FMARK(26);
                const double du = -xs;
                if (lane < NU) dz[lane] = du;
            } else if (tid < 32 + NX) dz[NU + tid - 32] = dx[tid - 32];
            __syncthreads();
            FMARK(24);
            if (tid < NZ) vk[V_DZ + tid] = dz[tid];
            if (tid < 2 * NQ) dxn[tid] = fq_g(sl + A_H, dz, tid, dt, c2) + sl[A_RB + tid];
            else if (tid >= 64 && tid < 96) {                // momentum rows: four threads per row (warp 2, all lanes shuffle)
                const int hi = min((tid - 64) >> 2, 5), part = tid & 3;
                const double* hq = sl + A_H + SR_HQ + NQ * hi;
                const double* hf = sl + A_H + SR_HF + 12 * hi;
                double s0 = 0.0, s1 = 0.0;
#pragma unroll
                for (int j = part; j < NQ; j += 8) {
                    s0 = fma(hq[j], dz[IQ + j], s0);
                    if (j + 4 < NQ) s1 = fma(hq[j + 4], dz[IQ + j + 4], s1);
                }
#pragma unroll
                for (int j = part; j < 12; j += 8) {
                    s0 = fma(hf[j], dz[IFRC + j], s0);
                    if (j + 4 < 12) s1 = fma(hf[j + 4], dz[IFRC + j + 4], s1);
                }
                double s = s0 + s1;
                s += __shfl_xor_sync(0xffffffffu, s, 1);
                s += __shfl_xor_sync(0xffffffffu, s, 2);
                if (part == 0 && tid < 64 + 24) dxn[2 * NQ + hi] = s + dz[IH + hi] + sl[A_RB + 2 * NQ + hi];
            }
            __syncthreads();
            if (TOPT) {
                // time-opt: dxn holds y0 = A dx + B du + rb; tau = -(s + (P+ b)' y0) / d, then dx+ = y0 + b tau
                if (tid >= 96) {
                    double a0 = 0.0;
                    for (int j = tid - 96; j < NX; j += 32) a0 = fma(sl[A_XV + VX_PB + j], dxn[j], a0);
                    a0 = warp_sum(a0);
                    if (tid == 96) {
                        const double tau = -(sl[A_XV + VX_S] + a0) * sl[A_XV + VX_DINV];
                        av[AV_X] = tau; vk[V_X + VX_DTAU] = tau;
                    }
                }
                __syncthreads();
                if (tid < NX) dxn[tid] = fma(sl[A_XS + SRX_B + tid], av[AV_X], dxn[tid]);
                __syncthreads();
            }
        } else {
            if (tid < NU) vk[V_DZ + tid] = 0.0;
            else if (tid < NZ) vk[V_DZ + tid] = dx[tid - NU];
        }
    }
    fq_pass_end();
    return ph;
}

// =============================================================================================
// passes A2 / D2: C dz, slack and multiplier directions, step lengths (and, D2, dpi_{k-1} = P_k dx_k + p_k)
// out: [0] alpha_p [1] alpha_d [2..4] sum (lam + a dlam)(t + a dt) = c0 + a c1 + a^2 c2
// =============================================================================================
template <bool CORR, bool TOPT>
__device__ __noinline__ void fq_issue_B(int k, int slot)
{
    const FqCtx& c = fq_ctx;
    constexpr int SLOT = CORR ? D_SLOT : B_SLOT;
    constexpr int OC = CORR ? D_C : B_C, ODZ = CORR ? D_DZ : B_DZ, OLAM = CORR ? D_LAM : B_LAM, OT = CORR ? D_T : B_T;
    double* sl = fq_smem + slot * SLOT;
    uint64_t* bar = &fq_bars[slot];
    uint32_t bytes = (SR_CSZ + NZ + 8 * NG) * 8;
    if (CORR && k > 0) bytes += (NX * P_LD + NX) * 8;
    if (TOPT && k < c.N) bytes += (VX_SZ + QX_SZ) * 8;
    mbar_arrive_expect_tx(bar, bytes);
    if (TOPT && k < c.N) {
        bulk_g2s(sl + (CORR ? D_XV : B_XV), c.vb + (size_t)k * V_STAGE + V_X, VX_SZ * 8, bar);
        bulk_g2s(sl + (CORR ? D_XQ : B_XQ), c.qx + (size_t)k * QX_SZ, QX_SZ * 8, bar);
    }
    bulk_g2s(sl + OC, c.SRb + (size_t)k * SR_SZ + SR_CF, SR_CSZ * 8, bar);
    bulk_g2s(sl + ODZ, c.vb + (size_t)k * V_STAGE + V_DZ, (NZ + 4 * NG) * 8, bar);
    bulk_g2s(sl + OLAM, c.qlam + (size_t)k * 2 * NG, 2 * NG * 8, bar);
    bulk_g2s(sl + OT, c.qt + (size_t)k * 2 * NG, 2 * NG * 8, bar);
    if (CORR && k > 0) {
        bulk_g2s(sl + D_P, c.FRb + (size_t)k * FR_SZ + FR_P, NX * P_LD * 8, bar);
        bulk_g2s(sl + D_PV, c.vb + (size_t)k * V_STAGE + V_P, NX * 8, bar);
    }
}

template <bool CORR, bool TOPT>
__device__ __noinline__ uint32_t fq_pass_B(uint32_t ph, double* out)
{
    constexpr int SLOT = CORR ? D_SLOT : B_SLOT, DEPTH = CORR ? D_DEPTH : B_DEPTH;
    constexpr int OC = CORR ? D_C : B_C, ODZ = CORR ? D_DZ : B_DZ, ORD = CORR ? D_RD : B_RD, ORM = CORR ? D_RM : B_RM;
    constexpr int OLAM = CORR ? D_LAM : B_LAM, OT = CORR ? D_T : B_T;
    const int tid = threadIdx.x;
    const int N = fq_ctx.N;
    double* cz = fq_smem + (CORR ? DV_BASE : BV_BASE);
    double* red = fq_smem + FQ_RED;
    double r_p = 1.0, r_d = 1.0, cs0 = 0.0, cs1 = 0.0, cs2 = 0.0;
    if (tid == FQ_ISSUER)
        for (int i = 0; i < DEPTH - 1 && i <= N; i++) fq_issue_B<CORR, TOPT>(i, i % DEPTH);
    for (int k = 0; k <= N; k++) {
        const int slot = k % DEPTH;
        if (!CORR) FMARK(9);
        if (tid == FQ_ISSUER && k + DEPTH - 1 <= N) fq_issue_B<CORR, TOPT>(k + DEPTH - 1, (k + DEPTH - 1) % DEPTH);
        if (!CORR) FMARK(14);
        fq_ring_wait(slot, ph);
        if (!CORR) FMARK(15);
        const double* sl = fq_smem + slot * SLOT;
        const double* Sc = sl + OC - SR_CF;                  // the helpers index the constraint part by its stage-record offsets
        double* vk = fq_ctx.vb + (size_t)k * V_STAGE;
        fq_cz(Sc, sl + ODZ, cz, tid);                        // threads 0..41
        if (CORR && k > 0 && tid >= 64 && tid < 64 + NX) {   // dpi_{k-1} = P_k dx_k + p_k
            const int i = tid - 64;
            const double* row = sl + D_P + i * P_LD;
            const double* dx = sl + ODZ + NU;
            double s0 = sl[D_PV + i], s1 = 0.0;
#pragma unroll
            for (int j = 0; j < NX; j += 2) { s0 = fma(row[j], dx[j], s0); s1 = fma(row[j + 1], dx[j + 1], s1); }
            vk[V_DPI - V_STAGE + i] = s0 + s1;
        }
        __syncthreads();
        if (!CORR) FMARK(16);
        if (tid < 2 * NG) {
            const int e = tid / NG, j = tid % NG;
            double dtv = 0.0, dl = 0.0;
            if (Sc[SR_D + 2 * NG + tid] > 0.5) {
                const double lam = sl[OLAM + tid], t = sl[OT + tid];
                const double cdz = cz[j];
                dtv = (e ? -cdz : cdz) + sl[ORD + tid];
                dl = -(sl[ORM + tid] + lam * dtv) * __drcp_rn(t);
                // step lengths: the largest ratios -dt / t, -dlam / lam are tracked (one division at the end of the pass
                // instead of two per row); fmax drops a NaN ratio the way the comparison dt < 0 did
                r_p = fmax(r_p, -dtv * __drcp_rn(t));
                r_d = fmax(r_d, -dl * __drcp_rn(lam));
                cs0 = fma(lam, t, cs0); cs1 += fma(lam, dtv, t * dl); cs2 = fma(dl, dtv, cs2);
            }
            vk[V_DT + tid] = dtv; vk[V_DLAM + tid] = dl;
        } else if (TOPT && k < N && tid == FQ_THREADS - 1) {     // the two bounds of dt
            const double* xv = sl + (CORR ? D_XV : B_XV);
            const double* xq = sl + (CORR ? D_XQ : B_XQ);
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const double lam = xq[2 + e], t = xq[4 + e];
                const double dtv = (e ? -xv[VX_DTAU] : xv[VX_DTAU]) + xv[VX_RD + e];
                const double dl = -(xv[VX_RM + e] + lam * dtv) * __drcp_rn(t);
                r_p = fmax(r_p, -dtv * __drcp_rn(t));
                r_d = fmax(r_d, -dl * __drcp_rn(lam));
                cs0 = fma(lam, t, cs0); cs1 += fma(lam, dtv, t * dl); cs2 = fma(dl, dtv, cs2);
                vk[V_X + VX_DT + e] = dtv; vk[V_X + VX_DL + e] = dl;
            }
        }
        __syncthreads();
        if (!CORR) FMARK(17);
    }
    out[0] = 1.0 / fq_block_max(r_p, red, tid);              // min(1, min_i -t_i / dt_i)
    out[1] = 1.0 / fq_block_max(r_d, red, tid);
    out[2] = fq_block_sum(cs0, red, tid);
    out[3] = fq_block_sum(cs1, red, tid);
    out[4] = fq_block_sum(cs2, red, tid);
    fq_pass_end();
    return ph;
}

// =============================================================================================
// pass C1: corrector complementarity residual, gamma_c, m0 = rg + C' gamma_c + G' (P+ rb)
// =============================================================================================
template <bool TOPT>
__device__ __noinline__ void fq_issue_C1(int k, int slot)
{
    const FqCtx& c = fq_ctx;
    double* sl = fq_smem + slot * C_SLOT;
    uint64_t* bar = &fq_bars[slot];
    const double* vk = c.vb + (size_t)k * V_STAGE;
    mbar_arrive_expect_tx(bar, (180 + SR_CSZ + (TOPT ? SRX_SZ : 0) + NZ + NX + 10 * NG + ((TOPT && k < c.N) ? VX_SZ + QX_SZ : 0)) * 8);
    if (TOPT && k < c.N) {
        bulk_g2s(sl + C_XV, vk + V_X, VX_SZ * 8, bar);
        bulk_g2s(sl + C_XQ, c.qx + (size_t)k * QX_SZ, QX_SZ * 8, bar);
    }
    bulk_g2s(sl + C_H, c.SRb + (size_t)k * SR_SZ + SR_HQ, 180 * 8, bar);
    bulk_g2s(sl + C_C, c.SRb + (size_t)k * SR_SZ + SR_CF, (SR_CSZ + (TOPT ? SRX_SZ : 0)) * 8, bar);
    bulk_g2s(sl + C_RG, vk + V_RG, (NZ + NX) * 8, bar);
    bulk_g2s(sl + C_DT, vk + V_DT, 4 * NG * 8, bar);
    bulk_g2s(sl + C_RD, vk + V_RD, 2 * NG * 8, bar);
    bulk_g2s(sl + C_LAM, c.qlam + (size_t)k * 2 * NG, 2 * NG * 8, bar);
    bulk_g2s(sl + C_T, c.qt + (size_t)k * 2 * NG, 2 * NG * 8, bar);
}

template <bool TOPT>
__device__ __noinline__ uint32_t fq_pass_C1(uint32_t ph, int cpass, double sigma_mu)
{
    const int tid = threadIdx.x;
    const int N = fq_ctx.N;
    double* gamc = fq_smem + CV_BASE;
    if (tid == FQ_ISSUER)
        for (int i = 0; i < C_DEPTH - 1 && i <= N; i++) fq_issue_C1<TOPT>(i, i % C_DEPTH);
    for (int k = 0; k <= N; k++) {
        const int slot = k % C_DEPTH;
        const bool term = (k == N);
        if (tid == FQ_ISSUER && k + C_DEPTH - 1 <= N) fq_issue_C1<TOPT>(k + C_DEPTH - 1, (k + C_DEPTH - 1) % C_DEPTH);
        fq_ring_wait(slot, ph);
        const double* sl = fq_smem + slot * C_SLOT;
        const double* Sc = sl + C_C - SR_CF;
        double* vk = fq_ctx.vb + (size_t)k * V_STAGE;
        if (tid < NG) {
            double g = 0.0;
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int row = e * NG + tid;
                if (Sc[SR_D + 2 * NG + row] > 0.5) {
                    const double lam = sl[C_LAM + row], t = sl[C_T + row];
                    const double rmc = fma(lam, t, cpass == 0 ? sl[C_DT + row] * sl[C_DLAM + row] : 0.0) - sigma_mu;
                    vk[V_RM + row] = rmc;
                    const double v = (rmc + lam * sl[C_RD + row]) * __drcp_rn(t);
                    g += e ? -v : v;
                }
            }
            gamc[tid] = g;
        } else if (TOPT && !term && tid == FQ_THREADS - 1) {     // the two bounds of dt
            double g = 0.0;
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const double lam = sl[C_XQ + 2 + e], t = sl[C_XQ + 4 + e];
                const double rmc = fma(lam, t, cpass == 0 ? sl[C_XV + VX_DT + e] * sl[C_XV + VX_DL + e] : 0.0) - sigma_mu;
                vk[V_X + VX_RM + e] = rmc;
                const double v = (rmc + lam * sl[C_XV + VX_RD + e]) * __drcp_rn(t);
                g += e ? -v : v;
            }
            vk[V_X + VX_M0] = sl[C_XV + VX_RG] + g;
        }
        __syncthreads();
        if (tid < NZ) {
            double s = sl[C_RG + tid] + fq_ct(Sc, gamc, tid);
            // time-opt: the extension of the stage record follows the constraint part and is fetched with it
            if (!term) s += fq_gt(sl + C_H, sl + C_PRB, tid, TOPT ? sl[C_C + SR_CSZ + SRX_DT] : fq_ctx.dt, TOPT ? sl[C_C + SR_CSZ + SRX_C2] : fq_ctx.c2);
            vk[V_M0 + tid] = s;
        }
        __syncthreads();
    }
    fq_pass_end();
    return ph;
}

// =============================================================================================
// pass C2: corrector backward chain  m = m0 + G' p+,  lu = Luu^-1 m_u,  p = m_x - Lxu lu
// =============================================================================================
template <bool TOPT>
__device__ __noinline__ void fq_issue_C2(int k, int slot)
{
    const FqCtx& c = fq_ctx;
    double* sl = fq_smem + slot * E_SLOT;
    uint64_t* bar = &fq_bars[slot];
    mbar_arrive_expect_tx(bar, (LP_UPANELS + 180 + NZ + (TOPT ? SRX_SZ + VX_SZ : 0)) * 8);
    if (TOPT) {
        bulk_g2s(sl + E_XS, c.SRb + (size_t)k * SR_SZ + SR_X, SRX_SZ * 8, bar);
        bulk_g2s(sl + E_XV, c.vb + (size_t)k * V_STAGE + V_X, VX_SZ * 8, bar);
    }
    bulk_g2s(sl + E_MU, c.FRb + (size_t)k * FR_SZ + FR_MU, LP_UPANELS * 8, bar);
    bulk_g2s(sl + E_H, c.SRb + (size_t)k * SR_SZ + SR_HQ, 180 * 8, bar);
    bulk_g2s(sl + E_M0, c.vb + (size_t)k * V_STAGE + V_M0, NZ * 8, bar);
}

template <bool TOPT>
__device__ __noinline__ uint32_t fq_pass_C2(uint32_t ph)
{
    const int tid = threadIdx.x;
    const int N = fq_ctx.N;
    double* ev = fq_smem + EV_BASE;
    // terminal stage: p_N = m_x (read straight from the workspace)
    {
        double* vk = fq_ctx.vb + (size_t)N * V_STAGE;
        if (tid < NX) { const double pv = vk[V_M0 + NU + tid]; ev[EV_PN + (N & 1) * NX + tid] = pv; vk[V_P + tid] = pv; }
    }
    if (tid == FQ_ISSUER)
        for (int i = 0; i < E_DEPTH - 1 && i < N; i++) fq_issue_C2<TOPT>(N - 1 - i, i % E_DEPTH);
    __syncthreads();
    for (int i = 0; i < N; i++) {
        const int k = N - 1 - i, slot = i % E_DEPTH;
        if (tid == FQ_ISSUER && i + E_DEPTH - 1 < N) fq_issue_C2<TOPT>(k - (E_DEPTH - 1), (i + E_DEPTH - 1) % E_DEPTH);
        fq_ring_wait(slot, ph);
        const double* sl = fq_smem + slot * E_SLOT;
        const double* sM = sl + E_MU;
        double* vk = fq_ctx.vb + (size_t)k * V_STAGE;
        double* ms = ev + EV_M;
        double* lus = ev + EV_LU;
        const double* pnn = ev + EV_PN + ((k + 1) & 1) * NX;      // p_{k+1} (time-opt: replaced by p~ below)
        double* pnk = ev + EV_PN + (k & 1) * NX;
        const double dt = TOPT ? sl[E_XS + SRX_DT] : fq_ctx.dt, c2 = TOPT ? sl[E_XS + SRX_C2] : fq_ctx.c2;
        if (TOPT) {
            // time-opt: s = m0_t + b' p+, then p~ = p+ - (P+ b) s / d stands in for p+ (see pass F)
            if (tid >= 96) {
                double a0 = 0.0;
                for (int j = tid - 96; j < NX; j += 32) a0 = fma(sl[E_XS + SRX_B + j], pnn[j], a0);
                a0 = warp_sum(a0);
                if (tid == 96) { const double sx = sl[E_XV + VX_M0] + a0; ev[EV_X] = sx; vk[V_X + VX_S] = sx; }
            }
            __syncthreads();
            if (tid < NX) ev[EV_PT + tid] = fma(-sl[E_XV + VX_PB + tid], ev[EV_X] * sl[E_XV + VX_DINV], pnn[tid]);
            __syncthreads();
            pnn = ev + EV_PT;
        }
        if (tid < NZ) ms[tid] = sl[E_M0 + tid] + fq_gt(sl + E_H, pnn, tid, dt, c2);
        __syncthreads();
        if (tid < 32) {                                      // lu = Luu^-1 m_u (column-oriented, one warp)
            // blocked forward substitution (see pass A1): row li of the inverse of the diagonal block is the column li of
            // the stored block above the diagonal, 1 / L_ii on it
            double r0 = tid < NU ? ms[tid] : 0.0;
            const int ll = tid < NU ? tid : NU - 1;
            const int lb = ll / PB, li = ll % PB;
            double cf[PB];
            {
                const double* blk = sM + P_OFF(lb);
#pragma unroll
                for (int m = 0; m < PB; m++) cf[m] = (m <= li) ? blk[(m < li ? m : li) * PB + li] : 0.0;
            }
            // exchange through shared memory, as in pass A1: the solution goes straight into lus
            double* rs = ev + EV_RS;
            rs[tid] = r0;
            __syncwarp();
#pragma unroll 1
            for (int bk = 0; bk < NU / PB; bk++) {
                if (lb == bk) {
                    const double* r6 = rs + PB * bk;
                    double a0 = 0.0, a1 = 0.0;
#pragma unroll
                    for (int m = 0; m < PB; m += 2) { a0 = fma(cf[m], r6[m], a0); a1 = fma(cf[m + 1], r6[m + 1], a1); }
                    lus[tid] = a0 + a1;                      // lanes 30, 31 duplicate lane 29 (same value, their own slot)
                }
                __syncwarp();
                if (lb > bk) {
                    const double* y6 = lus + PB * bk;
                    double lv[PB];
                    load6(sM + P_OFF(bk) + (ll - PB * bk) * PB, lv);         // L(ll, 6 bk ..)
                    double u0 = 0.0, u1 = 0.0;
#pragma unroll
                    for (int m = 0; m < PB; m += 2) { u0 = fma(lv[m], y6[m], u0); u1 = fma(lv[m + 1], y6[m + 1], u1); }
                    rs[tid] -= u0 + u1;
                }
                __syncwarp();
            }
            const double m0 = lus[ll];
            if (tid < NU) { lus[tid] = m0; vk[V_KFF + tid] = m0; }
        }
        __syncthreads();
        if (tid < NX) {                                      // p = m_x - Lxu lu
            double s0 = ms[NU + tid], s1 = 0.0;
#pragma unroll
            for (int jb = 0; jb < 5; jb++) {
                double lv[6];
                load6(sM + P_OFF(jb) + (NU + tid - PB * jb) * PB, lv);
#pragma unroll
                for (int c = 0; c < 6; c += 2) { s0 -= lv[c] * lus[PB * jb + c]; s1 -= lv[c + 1] * lus[PB * jb + c + 1]; }
            }
            const double pv = s0 + s1;
            pnk[tid] = pv; vk[V_P + tid] = pv;
        }
        __syncthreads();
    }
    fq_pass_end();
    return ph;
}

// =============================================================================================
// the kernel
// =============================================================================================
template <bool TOPT>
__global__ void __launch_bounds__(FQ_THREADS, 4) nmpc_qp_fused_kernel(FqArgs a)
{
    const int tid = threadIdx.x;
    const int bl = a.bl0 + blockIdx.x, b = a.b0 + bl, N = a.N;
    if (a.sqp_done && a.sqp_done[b]) return;             // finished problems keep their statistics
    if (tid == 0) {
        fq_ctx.SRb = a.SR + (size_t)bl * (N + 1) * SR_SZ;
        fq_ctx.FRb = a.FR + (size_t)bl * (N + 1) * FR_SZ;
        fq_ctx.vb = a.vec + (size_t)bl * (N + 1) * V_STAGE;
        fq_ctx.qz = a.qz + (size_t)b * (N + 1) * NZ;
        fq_ctx.qpi = a.qpi + (size_t)b * N * NX;
        fq_ctx.qlam = a.qlam + (size_t)b * (N + 1) * 2 * NG;
        fq_ctx.qt = a.qt + (size_t)b * (N + 1) * 2 * NG;
        fq_ctx.qx = a.qx + (size_t)b * N * QX_SZ;
        fq_ctx.dt = a.dt; fq_ctx.c2 = a.c2; fq_ctx.N = N;
#pragma unroll
        for (int i = 0; i < 4; i++) mbar_init(&fq_bars[i], 1);
        fence_barrier_init();
#ifdef NMPC_TIMING
        for (int i = 0; i < 32; i++) fq_tacc[i] = 0ULL;
        fq_tprev = clock64();
#endif
    }
    __syncthreads();

    // ---- start of the QP ----
    {
        double* qz = fq_ctx.qz;
        double* qpi = fq_ctx.qpi;
        if (a.warm == 0)
            for (int i = tid; i < (N + 1) * NZ; i += FQ_THREADS) qz[i] = 0.0;
        if (a.warm < 2)
            for (int i = tid; i < N * NX; i += FQ_THREADS) qpi[i] = 0.0;
        __syncthreads();
        if (tid < NX)
            qz[NU + tid] = a.dx0_ext ? a.dx0_ext[(size_t)b * NX + tid]
                                     : a.x0[(size_t)b * NX + tid] - a.x[(size_t)b * (N + 1) * NX + tid];
        if (tid < NU) qz[(size_t)N * NZ + tid] = 0.0;
        if (TOPT && a.warm == 0)
            for (int i = tid; i < N; i += FQ_THREADS) fq_ctx.qx[(size_t)i * QX_SZ] = 0.0;
        fq_pass_end();
    }

    uint32_t ph = 0;
    int status = 1, iters = 0, nc_total = 0, ncond = 0;
    double alpha_p = 1.0, alpha_d = 1.0;

    for (int it = 0; it <= a.max_iter; it++) {
        const bool last = (it >= a.max_iter);
        if (it > 0) {
            alpha_p *= (1.0 - alpha_p) * 0.99 + alpha_p * 0.9999999;
            alpha_d *= (1.0 - alpha_d) * 0.99 + alpha_d * 0.9999999;
            if (alpha_p < IPM_ALPHA_MIN && alpha_d < IPM_ALPHA_MIN) { status = 2; break; }
        }
        iters = it;
        FMARK(0);
        // ================= update + residuals of every stage, norms, mu, exit tests =================
        double o[6];
        ph = fq_pass_R<TOPT>(ph, it, a.warm, alpha_p, alpha_d, o);
        FMARK(1);
        if (it == 0) nc_total = (int)(o[5] + 0.5);
        const double mu = nc_total > 0 ? o[4] / nc_total : 0.0;
        {
            bool bad = false, conv = true;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                if (isinf(o[i]) || !(o[i] == o[i])) bad = true;
                if (!(o[i] < a.tol)) conv = false;
                if (tid == 0) a.qp_res[(size_t)b * 4 + i] = o[i];         // residuals of the last iterate that was evaluated
            }
            if (bad) { status = 3; break; }
            if (conv) { status = 0; break; }
            if (last) { status = 1; break; }
        }
        // ================= factorisation sweep =================
        ph = fq_pass_F<TOPT>(ph);
        FMARK(3);
        // ================= affine direction, step lengths, centring parameter =================
        ph = fq_pass_A1<TOPT>(ph);
        FMARK(8);
        ph = fq_pass_B<false, TOPT>(ph, o);
        FMARK(9);
        double sigma_mu, mu_aff;
        {
            const double alpha_aff = fmin(o[0], o[1]);
            const double s = fma(alpha_aff, fma(alpha_aff, o[4], o[3]), o[2]);
            mu_aff = nc_total > 0 ? s / nc_total : 0.0;
            double sg = mu > 0.0 ? mu_aff / mu : 0.0;
            sg = sg * sg * sg;
            sigma_mu = fmax(sg * mu, IPM_TAU_MIN);
        }
        // Corrector on the stored factorisation.  Pass 0 is Mehrotra's centring-corrector direction (HPIPM
        // COMPUTE_CENTERING_CORRECTION_QP); pass 1 runs only when the conditional predictor-corrector rejects it
        // (HPIPM arg->cond_pred_corr): if the complementarity predicted along it exceeds twice the affine
        // prediction, the centring direction alone (COMPUTE_CENTERING_QP: rm = lam t - sigma mu) is solved for.
        for (int cpass = 0; cpass < 2; cpass++) {
            ph = fq_pass_C1<TOPT>(ph, cpass, sigma_mu);
            FMARK(10);
            ph = fq_pass_C2<TOPT>(ph);
            FMARK(11);
            ph = fq_pass_A1<TOPT>(ph);
            FMARK(12);
            ph = fq_pass_B<true, TOPT>(ph, o);
            FMARK(13);
            alpha_p = o[0]; alpha_d = o[1];
            if (cpass == 1 || !a.cond_pred_corr) break;
            const double al = fmin(alpha_p, alpha_d);
            const double mu_c = nc_total > 0 ? fma(al, fma(al, o[4], o[3]), o[2]) / nc_total : 0.0;
            if (!(mu_c > 2.0 * mu_aff)) break;
            ncond++;
        }
    }
#ifdef NMPC_TIMING
    if (bl == 0 && tid == 0)
        for (int i = 0; i < 32; i++) g_fq_cycles[i] += fq_tacc[i];
#endif
    // ---- statistics of this QP ----
    if (tid == 0) {
        a.qp_iter[b] = iters;
        a.qp_status[b] = status;
        a.qp_ncond[b] = ncond;
    }
    // ---- end of a QP inside an SQP iteration: full step (x, u) += dz, multipliers replaced by the QP's; the masks of
    // the inequality rows come from the stage record.  The last pass R left the iterate in global memory. ----
    if (a.apply) {
        __syncthreads();
        if (tid == 0 && a.sqp_iter) a.sqp_iter[b] += 1;
        if (status == 0 || status == 1) {
            const double* qz = fq_ctx.qz;
            const double* qpi = fq_ctx.qpi;
            const double* qlam = fq_ctx.qlam;
            const double* Sb = fq_ctx.SRb;
            double* x = a.sx + (size_t)b * (N + 1) * NX;
            double* u = a.su + (size_t)b * N * NU;
            double* pi = a.spi + (size_t)b * N * NX;
            double* lam = a.slam + (size_t)b * (N + 1) * 2 * NG;
            for (int i = tid; i < (N + 1) * NX; i += FQ_THREADS) x[i] += qz[(size_t)(i / NX) * NZ + NU + i % NX];
            for (int i = tid; i < N * NU; i += FQ_THREADS) u[i] += qz[(size_t)(i / NU) * NZ + i % NU];
            for (int i = tid; i < N * NX; i += FQ_THREADS) pi[i] = qpi[i];
            for (int i = tid; i < (N + 1) * 2 * NG; i += FQ_THREADS) {
                const int k = i / (2 * NG), id = i % (2 * NG);
                lam[i] = Sb[(size_t)k * SR_SZ + SR_D + 2 * NG + id] > 0.5 ? qlam[i] : 0.0;
            }
            if (TOPT)
                for (int k = tid; k < N; k += FQ_THREADS) {
                    const double* q = fq_ctx.qx + (size_t)k * QX_SZ;
                    a.dtn[(size_t)b * N + k] += q[0];
                    a.lamt[((size_t)b * N + k) * 2] = q[2]; a.lamt[((size_t)b * N + k) * 2 + 1] = q[3];
                }
        }
    }
}
