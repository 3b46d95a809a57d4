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
#define SR_CX (SR_CF + 20 * 3)       // [22][36] contact / patch / centroidal rows over (q, v)
#define SR_D (SR_CX + 22 * 36)       // lg[42] ug[42] mask_l[42] mask_u[42]
#define SR_SZ (SR_D + 4 * NG)        // 1692
#define NGX 22                       // constraint rows on the state
#define NGF 20                       // friction rows

// factor record (doubles): u panels of the stage factor (Luu, Lxu, row 72 = lu) | P | p
#define FR_MU 0
#define P_LD 43
#define FR_P (FR_MU + LP_UPANELS)            // 1830
#define FR_PV (FR_P + NX * P_LD)             // 3636
#define FR_SZ (FR_PV + NX + 2)               // 3680

static_assert(SR_SZ % 2 == 0 && FR_P % 2 == 0 && FR_PV % 2 == 0 && FR_SZ % 2 == 0, "16-byte alignment of bulk copies");

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
    for (int e = tid; e < NGX * 36; e += 128) S[SR_CX + e] = Cp[(NGF + e / 36) * C_LD + IQ + e % 36];
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
#define FQ_THREADS 128
#define FQ_WARPS (FQ_THREADS / 32)
#define N_UTILES 100                 // 3x6 tiles of rows 0..71 x columns 0..29 at or below the diagonal
#define N_XTILES 105                 // 3x3 tiles of the lower triangle of the 42x42 state block

struct FqArgs {
    const double* SR;               // [Bc][N+1][SR_SZ]
    double* FR;                     // [Bc][N+1][FR_SZ]
    double* vec;                    // [Bc][(N+1)*V_STAGE]
    double* qz; double* qpi; double* qlam; double* qt;       // QP iterate, global problem index
    int32_t* qp_iter; int32_t* qp_status; int32_t* qp_ncond; double* qp_res;
    const double* dx0_ext; const double* x0; const double* x;
    const int32_t* sqp_done;
    int N, b0, max_iter, warm, cond_pred_corr;   // warm: HPIPM warm_start level 0 / 1 / 2 (see nmpc_b200.h)
    double tol, dt, c2;              // c2 = dt^2 / 2 (host-computed: both stay constant-bank operands)
};

// shared memory carve-up (doubles)
#define FS_P 0                               // P_{k+1}, full symmetric, ld 43
#define FS_M (FS_P + NX * P_LD)              // u panels of the stage matrix (LP_UPANELS); before the
                                             // factorisation the same space holds U and PGh
#define FS_SR (FS_M + LP_UPANELS)            // stage record
#define FS_V (FS_SR + SR_SZ)                 // vectors
#define FS_U FS_M                            // U[42][30] = P[:,h] [Hf | Hq]   (ld 30)
#define FS_PGH (FS_U + NX * NU)              // PGh[6][72] = (P G)[h rows]     (1260 + 432 <= 1830)
// vectors
#define FV_Z 0
#define FV_ZXN (FV_Z + NZ)
#define FV_PIK (FV_ZXN + NX)
#define FV_PIN (FV_PIK + NX)
#define FV_LAM (FV_PIN + NX)
#define FV_T (FV_LAM + 2 * NG)
#define FV_CZ (FV_T + 2 * NG)
#define FV_RD (FV_CZ + NG)
#define FV_RM (FV_RD + 2 * NG)
#define FV_GAM (FV_RM + 2 * NG)
#define FV_GAMV (FV_GAM + NG)
#define FV_DL (FV_GAMV + NG)
#define FV_RG (FV_DL + NG)
#define FV_QT (FV_RG + NZ)                   // m: linear term, row 72 of the stage matrix (NZ)
#define FV_RB (FV_QT + NZ)
#define FV_W (FV_RB + NX)
#define FV_PN (FV_W + NX)                    // p_{k+1}
#define FV_DX (FV_PN + NX)
#define FV_DXN (FV_DX + NX)
#define FV_DZ (FV_DXN + NX)
#define FV_RED (FV_DZ + NZ)                  // 16 reduction slots
#define FV_END (FV_RED + 16)
// per-thread running norms / complementarity sum of the backward pass (5 x 128): in shared memory rather than in
// registers that would be spilled across the panel factorisation of every stage
#define FS_ACC (FS_V + FV_END)
#define FQ_SMEM_BYTES ((FS_ACC + 5 * FQ_THREADS) * 8)

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
            const double* cx = S + SR_CX + 36 * (j - NGF);
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
        for (int jj = 0; jj < NGX; jj += 2) { s += cx[36 * jj] * g[NGF + jj]; s1 += cx[36 * (jj + 1)] * g[NGF + jj + 1]; }
        return s + s1;
    }
    return g[GR_CONS + c - IH];
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

__device__ __forceinline__ double fq_block_max(double v, double* red, int tid)
{
    v = warp_max(v);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    return nanmax(nanmax(red[0], red[1]), nanmax(red[2], red[3]));
}
__device__ __forceinline__ double fq_block_min(double v, double* red, int tid)
{
    v = warp_min(v);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    return fmin(fmin(red[0], red[1]), fmin(red[2], red[3]));
}
__device__ __forceinline__ double fq_block_sum(double v, double* red, int tid)
{
    v = warp_sum(v);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    return ((red[0] + red[1]) + red[2]) + red[3];
}

// ---------------------------------------------------------------------------------------------
// forward sweep, one stage.  In: sM = u panels of stage k (Luu, Lxu, row 72 = lu), S = SR_k,
// sv[FV_DX] = dx_k, sv[FV_RB] = rb_k; CORR: sP = P_k, sv[FV_PN] = p_k.
// Out: dz_k, dt_k, dlam_k (and CORR: dpi_{k-1}) to the stage vectors, step-length candidates
// folded into a_p / a_d, sv[FV_DX] = dx_{k+1}.
// ---------------------------------------------------------------------------------------------
template <bool CORR>
__device__ __forceinline__ void fq_forward_stage(const double* sM, const double* sP, const double* S, double* sv,
                                                 double* vk, double* dpi_prev, const double* qlam, const double* qt,
                                                 bool term, int tid, double dt, double c2, double& a_p, double& a_d,
                                                 double (&cs)[3])
{
    const int lane = tid & 31;
    // this thread's multiplier data, requested before the serial chain below so that the loads are not on it
    double g_lam = 0.0, g_t = 1.0, g_rd = 0.0, g_rm = 0.0;
    if (tid < 2 * NG) { g_lam = qlam[tid]; g_t = qt[tid]; g_rd = vk[V_RD + tid]; g_rm = vk[V_RM + tid]; }
    if (CORR && dpi_prev && tid >= 64 && tid < 64 + NX) {           // dpi_{k-1} = P_k dx_k + p_k
        const int i = tid - 64;
        const double* row = sP + i * P_LD;
        double s0 = sv[FV_PN + i], s1 = 0.0;
#pragma unroll
        for (int j = 0; j < NX; j += 2) { s0 += row[j] * sv[FV_DX + j]; s1 += row[j + 1] * sv[FV_DX + j + 1]; }
        dpi_prev[i] = s0 + s1;
    }
    if (!term) {
        if (tid < 32) {             // du = -Luu^-T (lu + Lxu' dx)
            double rhs = 0.0;
            if (lane < NU) {
                double r1 = 0.0;
                rhs = sM[lp_idx(NZ, lane)];
                for (int l = 0; l < NX; l += 2) {
                    rhs += sM[lp_idx(NU + l, lane)] * sv[FV_DX + l];
                    r1 += sM[lp_idx(NU + l + 1, lane)] * sv[FV_DX + l + 1];
                }
                rhs += r1;
            }
            // back substitution with Luu': branch-free, one shuffle + multiply + fused multiply-add per step;
            // lane l keeps rhs_l * (1 / L_ll) ready so the broadcast value is already the solution component
            double du = 0.0, myinv = 0.0;
            const int ll = lane < NU ? lane : NU - 1;
            { const double d = sM[lp_idx(ll, ll)]; myinv = d != 0.0 ? __drcp_rn(d) : 0.0; }
            // L(i, ll) for i >= ll sits at colb + 6 i (panel-packed, 6 doubles per row): one register instead of a
            // table of 30 per-lane addresses, which the compiler hoisted out of the stage loop and kept in local memory
            const double* colb = sM + P_OFF(ll / PB) - PB * PB * (ll / PB) + ll % PB;
#pragma unroll
            for (int i = NU - 1; i >= 0; i--) {
                const double dui = -__shfl_sync(0xffffffffu, rhs * myinv, i);
                const double lv = (lane < i) ? colb[PB * i] : 0.0;
                du = (lane == i) ? dui : du;
                rhs = (lane < i) ? fma(lv, dui, rhs) : rhs;
            }
            if (lane < NU) sv[FV_DZ + lane] = du;
        }
    } else if (tid < NU) sv[FV_DZ + tid] = 0.0;
    if (tid >= 32 && tid < 32 + NX) sv[FV_DZ + NU + tid - 32] = sv[FV_DX + tid - 32];
    __syncthreads();
    if (tid < NZ) vk[V_DZ + tid] = sv[FV_DZ + tid];
    fq_cz(S, sv + FV_DZ, sv + FV_CZ, tid);                           // threads 0..41
    if (!term && tid >= 64 && tid < 64 + NX) {                      // dx+ = G dz + rb
        const int l = tid - 64;
        sv[FV_DXN + l] = fq_g(S, sv + FV_DZ, l, dt, c2) + sv[FV_RB + l];
    }
    __syncthreads();
    if (tid < 2 * NG) {
        const int e = tid / NG, j = tid % NG;
        double dtv = 0.0, dl = 0.0;
        if (S[SR_D + 2 * NG + tid] > 0.5) {
            const double lam = g_lam, t = g_t;
            const double cdz = sv[FV_CZ + j];
            dtv = (e ? -cdz : cdz) + g_rd;
            dl = -(g_rm + lam * dtv) / t;
            if (dtv < 0.0) a_p = fmin(a_p, -t / dtv);
            if (dl < 0.0) a_d = fmin(a_d, -lam / dl);
            // sum (lam + a dlam)(t + a dt) = cs0 + a cs1 + a^2 cs2: neither the centring parameter (affine sweep) nor the
            // conditional predictor-corrector test (corrector sweep) needs an extra pass over the stages
            cs[0] = fma(lam, t, cs[0]); cs[1] += fma(lam, dtv, t * dl); cs[2] = fma(dl, dtv, cs[2]);
        }
        vk[V_DT + tid] = dtv; vk[V_DLAM + tid] = dl;
    }
    if (!term && tid < NX) sv[FV_DX + tid] = sv[FV_DXN + tid];
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// register tiles of the stage matrix, sorted by block type so that a warp meets at most three or
// four different assembly formulas: {row type, column type, row block, column block}
//   u tiles (3 x 6): rows of z = [a f q v h] (types 0..4) x columns a / f (types 0 / 1)
//   x tiles (3 x 3): rows q / v / h (types 0..2) x columns q / v / h of the 42 x 42 state block
// ---------------------------------------------------------------------------------------------
__constant__ unsigned char c_utile[100][4] = {{0,0,0,0}, {0,0,1,0}, {0,0,2,0}, {0,0,3,0}, {0,0,4,0}, {0,0,5,0}, {0,0,2,1}, {0,0,3,1}, {0,0,4,1}, {0,0,5,1}, {0,0,4,2}, {0,0,5,2}, {1,0,6,0}, {1,0,7,0}, {1,0,8,0}, {1,0,9,0}, {1,0,6,1}, {1,0,7,1}, {1,0,8,1}, {1,0,9,1}, {1,0,6,2}, {1,0,7,2}, {1,0,8,2}, {1,0,9,2}, {2,0,10,0}, {2,0,11,0}, {2,0,12,0}, {2,0,13,0}, {2,0,14,0}, {2,0,15,0}, {2,0,10,1}, {2,0,11,1}, {2,0,12,1}, {2,0,13,1}, {2,0,14,1}, {2,0,15,1}, {2,0,10,2}, {2,0,11,2}, {2,0,12,2}, {2,0,13,2}, {2,0,14,2}, {2,0,15,2}, {3,0,16,0}, {3,0,17,0}, {3,0,18,0}, {3,0,19,0}, {3,0,20,0}, {3,0,21,0}, {3,0,16,1}, {3,0,17,1}, {3,0,18,1}, {3,0,19,1}, {3,0,20,1}, {3,0,21,1}, {3,0,16,2}, {3,0,17,2}, {3,0,18,2}, {3,0,19,2}, {3,0,20,2}, {3,0,21,2}, {4,0,22,0}, {4,0,23,0}, {4,0,22,1}, {4,0,23,1}, {4,0,22,2}, {4,0,23,2}, {1,1,6,3}, {1,1,7,3}, {1,1,8,3}, {1,1,9,3}, {1,1,8,4}, {1,1,9,4}, {2,1,10,3}, {2,1,11,3}, {2,1,12,3}, {2,1,13,3}, {2,1,14,3}, {2,1,15,3}, {2,1,10,4}, {2,1,11,4}, {2,1,12,4}, {2,1,13,4}, {2,1,14,4}, {2,1,15,4}, {3,1,16,3}, {3,1,17,3}, {3,1,18,3}, {3,1,19,3}, {3,1,20,3}, {3,1,21,3}, {3,1,16,4}, {3,1,17,4}, {3,1,18,4}, {3,1,19,4}, {3,1,20,4}, {3,1,21,4}, {4,1,22,3}, {4,1,23,3}, {4,1,22,4}, {4,1,23,4}};
__constant__ unsigned char c_xtile[105][4] = {{0,0,0,0}, {0,0,1,0}, {0,0,1,1}, {0,0,2,0}, {0,0,2,1}, {0,0,2,2}, {0,0,3,0}, {0,0,3,1}, {0,0,3,2}, {0,0,3,3}, {0,0,4,0}, {0,0,4,1}, {0,0,4,2}, {0,0,4,3}, {0,0,4,4}, {0,0,5,0}, {0,0,5,1}, {0,0,5,2}, {0,0,5,3}, {0,0,5,4}, {0,0,5,5}, {1,0,6,0}, {1,0,6,1}, {1,0,6,2}, {1,0,6,3}, {1,0,6,4}, {1,0,6,5}, {1,0,7,0}, {1,0,7,1}, {1,0,7,2}, {1,0,7,3}, {1,0,7,4}, {1,0,7,5}, {1,0,8,0}, {1,0,8,1}, {1,0,8,2}, {1,0,8,3}, {1,0,8,4}, {1,0,8,5}, {1,0,9,0}, {1,0,9,1}, {1,0,9,2}, {1,0,9,3}, {1,0,9,4}, {1,0,9,5}, {1,0,10,0}, {1,0,10,1}, {1,0,10,2}, {1,0,10,3}, {1,0,10,4}, {1,0,10,5}, {1,0,11,0}, {1,0,11,1}, {1,0,11,2}, {1,0,11,3}, {1,0,11,4}, {1,0,11,5}, {1,1,6,6}, {1,1,7,6}, {1,1,7,7}, {1,1,8,6}, {1,1,8,7}, {1,1,8,8}, {1,1,9,6}, {1,1,9,7}, {1,1,9,8}, {1,1,9,9}, {1,1,10,6}, {1,1,10,7}, {1,1,10,8}, {1,1,10,9}, {1,1,10,10}, {1,1,11,6}, {1,1,11,7}, {1,1,11,8}, {1,1,11,9}, {1,1,11,10}, {1,1,11,11}, {2,0,12,0}, {2,0,12,1}, {2,0,12,2}, {2,0,12,3}, {2,0,12,4}, {2,0,12,5}, {2,0,13,0}, {2,0,13,1}, {2,0,13,2}, {2,0,13,3}, {2,0,13,4}, {2,0,13,5}, {2,1,12,6}, {2,1,12,7}, {2,1,12,8}, {2,1,12,9}, {2,1,12,10}, {2,1,12,11}, {2,1,13,6}, {2,1,13,7}, {2,1,13,8}, {2,1,13,9}, {2,1,13,10}, {2,1,13,11}, {2,2,12,12}, {2,2,13,12}, {2,2,13,13}};

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

#ifdef NMPC_TIMING
__device__ unsigned long long g_fq_cycles[24];
#define FMARK(ph) do { const long long t_ = clock64(); f_acc[ph] += (unsigned long long)(t_ - f_prev); f_prev = t_; } while (0)
#else
#define FMARK(ph) do { } while (0)
#endif

__device__ __forceinline__ int fq_opaque(int v) { asm volatile("" : "+r"(v)); return v; }

__global__ void __launch_bounds__(FQ_THREADS, 4) nmpc_qp_fused_kernel(FqArgs a)
{
#ifdef NMPC_TIMING
    unsigned long long f_acc[24];
#pragma unroll
    for (int i = 0; i < 24; i++) f_acc[i] = 0ULL;
    long long f_prev = clock64();
#endif
    extern __shared__ __align__(16) double smem[];
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x;
    const int bl = blockIdx.x, b = a.b0 + bl, N = a.N;
    if (a.sqp_done && a.sqp_done[b]) return;             // finished problems keep their statistics
    double* sP = smem + FS_P;
    double* sM = smem + FS_M;
    double* S = smem + FS_SR;
    double* sv = smem + FS_V;
    double* sU = smem + FS_U;
    double* red = sv + FV_RED;
    const double dt = a.dt, c2 = a.c2;
    const int tid_k = tid;

    // Base pointers of this OCP are NOT kept in registers for the lifetime of the kernel (seven 64-bit values that the
    // register peak of the panel factorisation would push into local memory): every use recomputes them from the
    // kernel parameters (constant bank) and the block index, behind an opaque move so that they are not hoisted.
#define SRb (a.SR + (size_t)fq_opaque(bl) * (N + 1) * SR_SZ)
#define FRb (a.FR + (size_t)fq_opaque(bl) * (N + 1) * FR_SZ)
#define vb (a.vec + (size_t)fq_opaque(bl) * (N + 1) * V_STAGE)
#define qz (a.qz + (size_t)fq_opaque(b) * (N + 1) * NZ)
#define qpi (a.qpi + (size_t)fq_opaque(b) * N * NX)
#define qlamb (a.qlam + (size_t)fq_opaque(b) * (N + 1) * 2 * NG)
#define qtb (a.qt + (size_t)fq_opaque(b) * (N + 1) * 2 * NG)

    uint32_t phase = 0;
    if (tid == 0) { mbar_init(&bar, 1); fence_barrier_init(); }

    // ---- start of the QP (nmpc_qp_init_kernel) ----
    if (a.warm == 0)
        for (int i = tid; i < (N + 1) * NZ; i += FQ_THREADS) qz[i] = 0.0;
    if (a.warm < 2)
        for (int i = tid; i < N * NX; i += FQ_THREADS) qpi[i] = 0.0;
    __syncthreads();
    if (tid < NX)
        qz[NU + tid] = a.dx0_ext ? a.dx0_ext[(size_t)b * NX + tid]
                                 : a.x0[(size_t)b * NX + tid] - a.x[(size_t)b * (N + 1) * NX + tid];
    if (tid < NU) qz[(size_t)N * NZ + tid] = 0.0;
    __syncthreads();

    // tile ownership, packed into one register.  The unpacked fields (and everything derived from them) used to be
    // kernel-lifetime values that the register peak of the panel factorisation pushed into local memory; with four
    // CTAs per SM those reloads miss the L1 left over next to 211 KB of shared memory.  Each use site below unpacks
    // what it needs behind an opaque move, so nothing but `tcode` stays live across the stage.
    unsigned tcode = 0;
    if (tid < N_UTILES)
        tcode |= c_utile[tid][0] | (c_utile[tid][1] << 3) | (c_utile[tid][2] << 4) | (c_utile[tid][3] << 9) | (1u << 24);
    if (tid < N_XTILES)
        tcode |= (c_xtile[tid][0] << 12) | (c_xtile[tid][1] << 14) | (c_xtile[tid][2] << 16) | (c_xtile[tid][3] << 20) | (1u << 25);
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

    int status = 1, iters = 0, nc_total = 0, ncond = 0;
    double alpha_p = 1.0, alpha_d = 1.0, mu = 0.0;

    for (int it = 0; it <= a.max_iter; it++) {
        const bool last = (it >= a.max_iter);
        // ================= update with the step of the previous iteration (nmpc_qp_update_kernel) =====
        if (it > 0) {
            alpha_p *= (1.0 - alpha_p) * 0.99 + alpha_p * 0.9999999;
            alpha_d *= (1.0 - alpha_d) * 0.99 + alpha_d * 0.9999999;
            if (alpha_p < IPM_ALPHA_MIN && alpha_d < IPM_ALPHA_MIN) { status = 2; break; }
            for (int k = 0; k <= N; k++) {
                const int tid = fq_opaque(tid_k);        // per-stage copy: thread-invariant addresses are not hoisted (and spilled)
                const double* vk = vb + (size_t)k * V_STAGE;
                const double* Dm = SRb + (size_t)k * SR_SZ + SR_D + 2 * NG;
                if (tid < NZ) qz[(size_t)k * NZ + tid] += alpha_p * vk[V_DZ + tid];
                if (k < N && tid >= 80 && tid < 80 + NX) qpi[(size_t)k * NX + tid - 80] += alpha_d * vk[V_DPI + tid - 80];
                if (tid < 2 * NG && Dm[tid] > 0.5) {
                    double* qt = qtb + (size_t)k * 2 * NG;
                    double* ql = qlamb + (size_t)k * 2 * NG;
                    qt[tid] = fmax(qt[tid] + alpha_p * vk[V_DT + tid], IPM_T_MIN);
                    ql[tid] = fmax(ql[tid] + alpha_d * vk[V_DLAM + tid], IPM_LAM_MIN);
                }
            }
            __syncthreads();
        }
        iters = it;
        FMARK(0);

        // ================= backward pass: residuals of every stage (+ factorisation unless last) =====
        double* acc = smem + FS_ACC + tid;       // [0] n_g  [1] n_b  [2] n_d  [3] n_m  [4] sum lam t   (stride FQ_THREADS)
#pragma unroll
        for (int i = 0; i < 5; i++) acc[i * FQ_THREADS] = 0.0;
        int cnt = 0;
        if (tid == 0) {
            bulk_store_wait_read();
            mbar_arrive_expect_tx(&bar, SR_SZ * 8);
            bulk_g2s(S, SRb + (size_t)N * SR_SZ, SR_SZ * 8, &bar);
        }
        for (int k = N; k >= 0; k--) {
            const int tid = fq_opaque(tid_k);        // per-stage copy: thread-invariant addresses are not hoisted (and spilled)
            const bool term = (k == N);
            double* vk = vb + (size_t)k * V_STAGE;
            double* qlam = qlamb + (size_t)k * 2 * NG;
            double* qt = qtb + (size_t)k * 2 * NG;
            if (tid < NZ) sv[FV_Z + tid] = qz[(size_t)k * NZ + tid];
            if (tid >= 80 && tid < 80 + NX) {
                const int i = tid - 80;
                sv[FV_ZXN + i] = term ? 0.0 : qz[(size_t)(k + 1) * NZ + NU + i];
                sv[FV_PIK + i] = (k > 0) ? qpi[(size_t)(k - 1) * NX + i] : 0.0;
                sv[FV_PIN + i] = term ? 0.0 : qpi[(size_t)k * NX + i];
            }
            if (tid < 2 * NG) { sv[FV_LAM + tid] = qlam[tid]; sv[FV_T + tid] = qt[tid]; }
            __syncthreads();
            FMARK(1);
            mbar_wait(&bar, phase); phase ^= 1;                      // SR_k has landed
            FMARK(2);

            fq_cz(S, sv + FV_Z, sv + FV_CZ, tid);
            __syncthreads();
            if (tid < 2 * NG) {
                const int e = tid / NG, j = tid % NG;
                const double msk = S[SR_D + 2 * NG + tid];
                const double cz = sv[FV_CZ + j];
                const double slack = e ? (S[SR_D + NG + j] - cz) : (cz - S[SR_D + j]);
                double t = sv[FV_T + tid], lam = sv[FV_LAM + tid];
                if (it == 0) {               // HPIPM INIT_VAR: levels 0 / 1 from the slacks, level 2 keeps lam, t clipped at thr0
                    if (msk > 0.5) {
                        if (a.warm < 2 || !(t > 0.0) || !(lam > 0.0)) { t = fmax(slack, IPM_THR0); lam = IPM_MU0 / t; }
                        else { t = fmax(t, IPM_THR0); lam = fmax(lam, IPM_THR0); }
                    } else { t = 1.0; lam = 0.0; }
                    sv[FV_T + tid] = t; sv[FV_LAM + tid] = lam;
                    qt[tid] = t; qlam[tid] = lam;
                }
                const double rd = msk > 0.5 ? (slack - t) : 0.0;
                const double rm = msk > 0.5 ? lam * t : 0.0;
                sv[FV_RD + tid] = rd; sv[FV_RM + tid] = rm;
                vk[V_RD + tid] = rd; vk[V_RM + tid] = rm;
                acc[4 * FQ_THREADS] += rm; cnt += (msk > 0.5) ? 1 : 0;
                acc[2 * FQ_THREADS] = nanmax(acc[2 * FQ_THREADS], fabs(rd));
                acc[3 * FQ_THREADS] = nanmax(acc[3 * FQ_THREADS], fabs(rm));
            }
            __syncthreads();
            if (tid < NG) {
                const int j = tid;
                const double ml = S[SR_D + 2 * NG + j], mu_ = S[SR_D + 3 * NG + j];
                const double ll = sv[FV_LAM + j], lu = sv[FV_LAM + NG + j];
                const double tl = sv[FV_T + j], tu = sv[FV_T + NG + j];
                double Gam = 0.0, gam = 0.0, dl = 0.0;
                if (ml > 0.5) { Gam += ll / tl; gam += (sv[FV_RM + j] + ll * sv[FV_RD + j]) / tl; dl += ll; }
                if (mu_ > 0.5) { Gam += lu / tu; gam -= (sv[FV_RM + NG + j] + lu * sv[FV_RD + NG + j]) / tu; dl -= lu; }
                sv[FV_GAM + j] = Gam; sv[FV_GAMV + j] = gam; sv[FV_DL + j] = dl;
            }
            __syncthreads();
            if (tid < NZ) {
                const int c = tid;
                double s = S[SR_GR + c] + fq_h(S, sv + FV_Z, c);
                if (c >= NU && k > 0) s -= sv[FV_PIK + c - NU];
                s -= fq_ct(S, sv + FV_DL, c);
                const double qp = fq_ct(S, sv + FV_GAMV, c);
                if (!term) s += fq_gt(S, sv + FV_PIN, c, dt, c2);
                if ((k == 0 && c >= NU) || (term && c < NU)) s = 0.0;
                vk[V_RG + c] = s; sv[FV_QT + c] = qp + s;
                acc[0] = nanmax(acc[0], fabs(s));
            } else if (!term && tid >= 80 && tid < 80 + NX) {
                const int l = tid - 80;
                const double rb = fq_g(S, sv + FV_Z, l, dt, c2) + S[SR_B + l] - sv[FV_ZXN + l];
                sv[FV_RB + l] = rb; vk[V_RB + l] = rb;
                acc[FQ_THREADS] = nanmax(acc[FQ_THREADS], fabs(rb));
            }
            __syncthreads();
            if (last) {
                if (k > 0 && tid == 0) {
                    mbar_arrive_expect_tx(&bar, SR_SZ * 8);
                    bulk_g2s(S, SRb + (size_t)(k - 1) * SR_SZ, SR_SZ * 8, &bar);
                }
                continue;
            }

            FMARK(3);
            // ---------------- Riccati step of stage k (classical recursion on P) ----------------
            double xa[3][3], ua[3][6];
            if (tid == 0) bulk_store_wait_read();                    // the previous factor record has left sM / sP
            if (!term) {
                if (tid < NX) {                                      // w = P+ rb + p+
                    const double* row = sP + tid * P_LD;
                    double s0 = sv[FV_PN + tid], s1 = 0.0;
#pragma unroll
                    for (int j = 0; j < NX; j += 2) { s0 += row[j] * sv[FV_RB + j]; s1 += row[j + 1] * sv[FV_RB + j + 1]; }
                    sv[FV_W + tid] = s0 + s1;
                }
                __syncthreads();                                     // also orders wait_read before the writes to sU
                if (tid < 4 * NU) {                                  // U = P+[:, h] [Hf | Hq]
                    // thread = (column j, row group g): the column of [Hf | Hq] stays in registers, the six P entries
                    // of a row are one broadcast per warp (12 loads per entry before, 6 per row now)
                    const int j = tid % NU, g = tid / NU;
                    const double* hc = (j < 12) ? (S + SR_HF + j) : (S + SR_HQ + (j - 12));
                    const int ld = (j < 12) ? 12 : NQ;
                    double hcv[6];
#pragma unroll
                    for (int h = 0; h < 6; h++) hcv[h] = hc[ld * h];
                    for (int r = g; r < NX; r += 4) {
                        const double* ph = sP + r * P_LD + 2 * NQ;
                        double s = 0.0;
#pragma unroll
                        for (int h = 0; h < 6; h++) s += ph[h] * hcv[h];
                        sU[r * NU + j] = s;
                    }
                }
                if (tid < NZ) sv[FV_QT + tid] += fq_gt(S, sv + FV_W, tid, dt, c2);       // m = q tilde + G' w
                __syncthreads();
                FMARK(4);
                // ---- u tile: rows ur0..+2 (z index), columns uc0..+5 (u index) ----
                FQ_UNPACK_U;
                // first row / column of the tile inside its block types
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
                    if (urb / 2 == ucb) {                            // diagonal tile: Hessian diagonal + regularisation
#pragma unroll
                        for (int r = 0; r < 3; r++)
#pragma unroll
                            for (int c = 0; c < 6; c++)
                                if (ur0 + r == uc0 + c) ua[r][c] += S[SR_RD + ur0 + r] + IPM_REG_PRIM;
                    }
                    if (ur0 >= IFRC && ur0 < IQ && uc0 >= IFRC) {    // friction rows: 3x3 block of the foot
                        const int ft = (ur0 - IFRC) / 3;
                        const int cofs = IFRC + 3 * ft - uc0;        // first column of the foot inside this tile
                        if (cofs >= 0 && cofs < 6) {
                            for (int r5 = 0; r5 < 5; r5++) {
                                const double g = sv[FV_GAM + 5 * ft + r5];
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
                if (i0 < 2 * NQ) {                                   // (q,v) x (q,v): all 22 state rows
                    for (int g = 0; g < NGX; g++) {
                        const double gm = sv[FV_GAM + NGF + g];
                        if (gm == 0.0) continue;
                        const double* row = S + SR_CX + 36 * g;
                        const double a0 = row[i0] * gm, a1 = row[i0 + 1] * gm, a2 = row[i0 + 2] * gm;
                        const double b0 = row[j0x], b1 = row[j0x + 1], b2 = row[j0x + 2];
                        xa[0][0] += a0 * b0; xa[0][1] += a0 * b1; xa[0][2] += a0 * b2;
                        xa[1][0] += a1 * b0; xa[1][1] += a1 * b1; xa[1][2] += a1 * b2;
                        xa[2][0] += a2 * b0; xa[2][1] += a2 * b1; xa[2][2] += a2 * b2;
                    }
                } else {                                             // h rows: only the centroidal rows reach them
#pragma unroll
                    for (int r = 0; r < 3; r++) {
                        const int hi = i0 + r - 2 * NQ;
                        const double gm = sv[FV_GAM + GR_CONS + hi];
#pragma unroll
                        for (int c = 0; c < 3; c++) {
                            const int jj = j0x + c;
                            if (jj < 2 * NQ) xa[r][c] += gm * S[SR_CX + 36 * (16 + hi) + jj];
                            else if (jj == i0 + r) xa[r][c] += gm;
                        }
                    }
                }
            }
            }
            __syncthreads();                 // every read of SR_k, P+, U, PGh is done
            FMARK(5);
            if (k > 0 && tid == 0) {         // prefetch the next stage record under the factorisation
                mbar_arrive_expect_tx(&bar, SR_SZ * 8);
                bulk_g2s(S, SRb + (size_t)(k - 1) * SR_SZ, SR_SZ * 8, &bar);
            }
            if (!term) {
                // ---- right-looking Cholesky of the 30 input columns; x tiles take the same rank-6 updates ----
                for (int jb = 0; jb < 5; jb++) {
                    const int j0 = PB * jb;
                    double* pan = sM + P_OFF(jb);                    // rows j0..72 of columns j0..j0+5
                    {
                    FQ_UNPACK_U;
                    if (has_ut && ucb == jb) {
#pragma unroll
                        for (int r = 0; r < 3; r++)
#pragma unroll
                            for (int c = 0; c < 6; c++) pan[(ur0 + r - j0) * PB + c] = ua[r][c];
                    }
                    }
                    if (tid < PB) pan[(NZ - j0) * PB + tid] = sv[FV_QT + j0 + tid];
                    __syncthreads();
                    // Panel factorisation, right-looking in registers: every row thread factors the 6x6 diagonal
                    // block redundantly and carries its own row along as a seventh row, so the dependent chain per
                    // pivot is one rsqrt, one multiply and one fused multiply-add.  Non-positive pivots zero their
                    // column (BLASFEO potrf rule).
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
                            double d = 0.0, inv = 0.0;
                            if (v > 0.0) { inv = rsqrt(v); d = v * inv; }
                            A[j][j] = d;
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
                    if (tid < NZ - j0 - PB) {                        // row 72: m_c -= lu . L(c, j0..j0+5)
                        const double* pa = pan + (NZ - j0) * PB;
                        const double* pb = pan + (PB + tid) * PB;
                        double av[6], bv[6];
                        load6(pa, av);
                        load6(pb, bv);
                        double sm_ = sv[FV_QT + j0 + PB + tid];
#pragma unroll
                        for (int kk = 0; kk < PB; kk++) sm_ -= av[kk] * bv[kk];
                        sv[FV_QT + j0 + PB + tid] = sm_;
                    }
                    // the next panel's owners write a different panel; the m row is ordered by the next barrier
                }
            }
            __syncthreads();
            FMARK(6);
            // ---- P_k (both triangles) -> sP, p_k -> sv; factor record -> HBM ----
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
            if (tid < NX) { const double pv = sv[FV_QT + NU + tid]; sv[FV_PN + tid] = pv; vk[V_P + tid] = pv; }
            fence_proxy_async();
            __syncthreads();
            if (tid == 0) {
                double* fr = FRb + (size_t)k * FR_SZ;
                if (!term) bulk_s2g(fr + FR_MU, sM, LP_UPANELS * 8);
                bulk_s2g(fr + FR_P, sP, NX * P_LD * 8);
            }
        }

        FMARK(7);
        // ================= norms, mu, exit tests (nmpc_qp_decide_kernel) =================
        double res4[4];
        res4[0] = fq_block_max(acc[0], red, tid);
        res4[1] = fq_block_max(acc[FQ_THREADS], red, tid);
        res4[2] = fq_block_max(acc[2 * FQ_THREADS], red, tid);
        res4[3] = fq_block_max(acc[3 * FQ_THREADS], red, tid);
        const double mu_sum = fq_block_sum(acc[4 * FQ_THREADS], red, tid);
        if (it == 0) nc_total = (int)(fq_block_sum((double)cnt, red, tid) + 0.5);
        mu = nc_total > 0 ? mu_sum / nc_total : 0.0;
        {
            bool bad = false, conv = true;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                if (isinf(res4[i]) || !(res4[i] == res4[i])) bad = true;
                if (!(res4[i] < a.tol)) conv = false;
                if (tid == 0) a.qp_res[(size_t)b * 4 + i] = res4[i];      // residuals of the last iterate that was evaluated
            }
            if (bad) { status = 3; break; }
            if (conv) { status = 0; break; }
            if (last) { status = 1; break; }
        }

        FMARK(8);
        // ================= forward sweep: affine direction, step lengths =================
        double a_p = 1.0, a_d = 1.0;
        double cs[3] = {0.0, 0.0, 0.0};
        if (tid < NX) sv[FV_DX + tid] = 0.0;
        if (tid == 0) bulk_store_wait_all();                         // the factor records are in global memory
        __syncthreads();
        for (int k = 0; k <= N; k++) {
            const int tid = fq_opaque(tid_k);        // per-stage copy: thread-invariant addresses are not hoisted (and spilled)
            const bool term = (k == N);
            double* vk = vb + (size_t)k * V_STAGE;
            if (tid == 0) {
                const uint32_t bytes = (SR_SZ + (term ? 0 : LP_UPANELS)) * 8;
                mbar_arrive_expect_tx(&bar, bytes);
                bulk_g2s(S, SRb + (size_t)k * SR_SZ, SR_SZ * 8, &bar);
                if (!term) bulk_g2s(sM, FRb + (size_t)k * FR_SZ + FR_MU, LP_UPANELS * 8, &bar);
            }
            if (!term && tid < NX) sv[FV_RB + tid] = vk[V_RB + tid];
            __syncthreads();
            mbar_wait(&bar, phase); phase ^= 1;
            fq_forward_stage<false>(sM, sP, S, sv, vk, nullptr, qlamb + (size_t)k * 2 * NG, qtb + (size_t)k * 2 * NG,
                                    term, tid, dt, c2, a_p, a_d, cs);
        }
        a_p = fq_block_min(a_p, red, tid);
        a_d = fq_block_min(a_d, red, tid);
        const double alpha_aff = fmin(a_p, a_d);

        FMARK(9);
        // ================= centring parameter (nmpc_qp_sigma_kernel) =================
        double sigma_mu, mu_aff;
        {
            const double s0 = fq_block_sum(cs[0], red, tid);
            const double s1 = fq_block_sum(cs[1], red, tid);
            const double s2 = fq_block_sum(cs[2], red, tid);
            const double s = fma(alpha_aff, fma(alpha_aff, s2, s1), s0);
            mu_aff = nc_total > 0 ? s / nc_total : 0.0;
            double sg = mu > 0.0 ? mu_aff / mu : 0.0;
            sg = sg * sg * sg;
            sigma_mu = fmax(sg * mu, IPM_TAU_MIN);
        }

        FMARK(10);
        // Corrector on the stored factorisation.  Pass 0 is Mehrotra's centring-corrector direction (HPIPM
        // COMPUTE_CENTERING_CORRECTION_QP); pass 1 runs only when the conditional predictor-corrector rejects it
        // (HPIPM arg->cond_pred_corr): if the complementarity predicted along it exceeds twice the affine
        // prediction, the centring direction alone (COMPUTE_CENTERING_QP: rm = lam t - sigma mu) is solved for.
        for (int cpass = 0; cpass < 2; cpass++) {
        // ================= corrector: backward vector sweep on the stored factorisation =================
        for (int k = N; k >= 0; k--) {
            const int tid = fq_opaque(tid_k);        // per-stage copy: thread-invariant addresses are not hoisted (and spilled)
            const bool term = (k == N);
            double* vk = vb + (size_t)k * V_STAGE;
            if (tid == 0) {
                // SR_k, the u panels of stage k, and (once, for the terminal stage) nothing else: P_{k+1} is
                // already in sP from the previous step of this sweep
                const uint32_t bytes = (SR_SZ + (term ? 0 : LP_UPANELS)) * 8;
                mbar_arrive_expect_tx(&bar, bytes);
                bulk_g2s(S, SRb + (size_t)k * SR_SZ, SR_SZ * 8, &bar);
                if (!term) bulk_g2s(sM, FRb + (size_t)k * FR_SZ + FR_MU, LP_UPANELS * 8, &bar);
            }
            double g_lam = 0.0, g_t = 1.0, g_dt = 0.0, g_dl = 0.0, g_rd = 0.0;
            if (tid < 2 * NG) {                                      // requested before the waits below
                g_lam = qlamb[(size_t)k * 2 * NG + tid]; g_t = qtb[(size_t)k * 2 * NG + tid];
                g_dt = vk[V_DT + tid]; g_dl = vk[V_DLAM + tid]; g_rd = vk[V_RD + tid];
            }
            if (!term && tid < NX) sv[FV_RB + tid] = vk[V_RB + tid];
            if (tid < NZ) sv[FV_RG + tid] = vk[V_RG + tid];
            __syncthreads();
            FMARK(13);
            mbar_wait(&bar, phase); phase ^= 1;
            FMARK(14);
            // corrector complementarity residual, gamma
            if (tid < 2 * NG) {
                double v = 0.0;
                if (S[SR_D + 2 * NG + tid] > 0.5) {
                    const double lam = g_lam, t = g_t;
                    const double rmc = fma(lam, t, cpass == 0 ? g_dt * g_dl : 0.0) - sigma_mu;
                    vk[V_RM + tid] = rmc;
                    v = (rmc + lam * g_rd) / t;
                }
                sv[FV_RD + tid] = v;                                 // scratch: per-side gamma
            }
            __syncthreads();
            if (tid < NG) sv[FV_GAMV + tid] = sv[FV_RD + tid] - sv[FV_RD + NG + tid];
            if (!term && tid >= 64 && tid < 64 + NX) {               // w = P+ rb + p+
                const int i = tid - 64;
                const double* row = sP + i * P_LD;
                double s0 = sv[FV_PN + i], s1 = 0.0;
#pragma unroll
                for (int j = 0; j < NX; j += 2) { s0 += row[j] * sv[FV_RB + j]; s1 += row[j + 1] * sv[FV_RB + j + 1]; }
                sv[FV_W + i] = s0 + s1;
            }
            __syncthreads();
            FMARK(15);
            if (tid < NZ) {                                          // m = rg + C' gamma + G' w
                double s = sv[FV_RG + tid] + fq_ct(S, sv + FV_GAMV, tid);
                if (!term) s += fq_gt(S, sv + FV_W, tid, dt, c2);
                sv[FV_QT + tid] = s;
            }
            __syncthreads();
            FMARK(16);
            if (tid == 0) {                                          // sP is free: fetch P_k for the next step
                mbar_arrive_expect_tx(&bar, NX * P_LD * 8);
                bulk_g2s(sP, FRb + (size_t)k * FR_SZ + FR_P, NX * P_LD * 8, &bar);
            }
            __syncwarp();         // reconverge warp 0 before its shuffles: a split warp takes the slow shuffle path
            if (!term) {
                if (tid < 32) {                                      // lu = Luu^-1 m_u (column-oriented, one warp)
                    double m0 = tid < NU ? sv[FV_QT + tid] : 0.0, myinv = 0.0;
                    const int ll = tid < NU ? tid : NU - 1;
                    { const double d = sM[lp_idx(ll, ll)]; myinv = d != 0.0 ? __drcp_rn(d) : 0.0; }
#pragma unroll
                    for (int j = 0; j < NU; j++) {
                        const double xj = __shfl_sync(0xffffffffu, m0 * myinv, j);
                        // L(ll, j), ll > j: compile-time panel offset + 6 ll
                        const double lv = (tid > j && tid < NU) ? sM[P_OFF(j / PB) - PB * PB * (j / PB) + j % PB + PB * ll] : 0.0;
                        m0 = (tid == j) ? xj : ((tid > j && tid < NU) ? fma(-lv, xj, m0) : m0);
                    }
                    if (tid < NU) { sv[FV_DZ + tid] = m0; vk[V_LROW + tid] = m0; }
                }
                __syncthreads();
                FMARK(17);
                if (tid < NX) {                                      // p = m_x - Lxu lu
                    double s0 = sv[FV_QT + NU + tid], s1 = 0.0;
                    const double* row = sM;                          // panel-packed: gather per panel
#pragma unroll
                    for (int jb = 0; jb < 5; jb++) {
                        double lv[6];
                        load6(row + P_OFF(jb) + (NU + tid - PB * jb) * PB, lv);
#pragma unroll
                        for (int c = 0; c < 6; c += 2) { s0 -= lv[c] * sv[FV_DZ + PB * jb + c]; s1 -= lv[c + 1] * sv[FV_DZ + PB * jb + c + 1]; }
                    }
                    const double pv = s0 + s1;
                    sv[FV_PN + tid] = pv; vk[V_P + tid] = pv;
                }
            } else if (tid < NX) {
                const double pv = sv[FV_QT + NU + tid];
                sv[FV_PN + tid] = pv; vk[V_P + tid] = pv;
            }
            __syncthreads();
            FMARK(18);
            mbar_wait(&bar, phase); phase ^= 1;                      // P_k is in sP
            __syncthreads();      // nobody may still be polling this phase when thread 0 re-arms the barrier
            FMARK(19);
        }

        FMARK(11);
        // ================= corrector: forward sweep with multiplier directions =================
        a_p = 1.0; a_d = 1.0;
        cs[0] = 0.0; cs[1] = 0.0; cs[2] = 0.0;
        if (tid < NX) sv[FV_DX + tid] = 0.0;
        __syncthreads();
        for (int k = 0; k <= N; k++) {
            const int tid = fq_opaque(tid_k);        // per-stage copy: thread-invariant addresses are not hoisted (and spilled)
            const bool term = (k == N);
            double* vk = vb + (size_t)k * V_STAGE;
            if (tid == 0) {
                const uint32_t bytes = (SR_SZ + (term ? 0 : LP_UPANELS) + (k > 0 ? NX * P_LD : 0)) * 8;
                mbar_arrive_expect_tx(&bar, bytes);
                bulk_g2s(S, SRb + (size_t)k * SR_SZ, SR_SZ * 8, &bar);
                if (!term) bulk_g2s(sM, FRb + (size_t)k * FR_SZ + FR_MU, LP_UPANELS * 8, &bar);
                if (k > 0) bulk_g2s(sP, FRb + (size_t)k * FR_SZ + FR_P, NX * P_LD * 8, &bar);
            }
            if (tid < NX) {
                sv[FV_RB + tid] = term ? 0.0 : vk[V_RB + tid];
                sv[FV_PN + tid] = vk[V_P + tid];
            }
            __syncthreads();
            mbar_wait(&bar, phase); phase ^= 1;
            if (!term && tid < NU) sM[lp_idx(NZ, tid)] = vk[V_LROW + tid];      // the corrector's lu replaces the affine one
            __syncthreads();
            fq_forward_stage<true>(sM, sP, S, sv, vk, k > 0 ? vb + (size_t)(k - 1) * V_STAGE + V_DPI : nullptr,
                                   qlamb + (size_t)k * 2 * NG, qtb + (size_t)k * 2 * NG, term, tid, dt, c2, a_p, a_d, cs);
        }
        alpha_p = fq_block_min(a_p, red, tid);
        alpha_d = fq_block_min(a_d, red, tid);
        if (cpass == 1 || !a.cond_pred_corr) break;
        {
            const double s0 = fq_block_sum(cs[0], red, tid);
            const double s1 = fq_block_sum(cs[1], red, tid);
            const double s2 = fq_block_sum(cs[2], red, tid);
            const double al = fmin(alpha_p, alpha_d);
            const double mu_c = nc_total > 0 ? fma(al, fma(al, s2, s1), s0) / nc_total : 0.0;
            if (!(mu_c > 2.0 * mu_aff)) break;
            ncond++;
        }
        }
        FMARK(12);
    }

#ifdef NMPC_TIMING
    if (bl == 0 && tid == 0) {
#pragma unroll
        for (int i = 0; i < 24; i++) g_fq_cycles[i] += f_acc[i];
    }
#endif
    // ---- statistics of this QP ----
    if (tid == 0) {
        bulk_store_wait_all();
        a.qp_iter[b] = iters;
        a.qp_status[b] = status;
        a.qp_ncond[b] = ncond;
    }
#undef SRb
#undef FRb
#undef vb
#undef qz
#undef qpi
#undef qlamb
#undef qtb
}
