// nmpc_qp_layout.cuh -- shared layouts of the OCP-QP interior point: panel-packed factor storage, the per-stage
// vector slots of the workspace, the argument block handed from the SQP driver to the QP kernels, and small
// block-level helpers.  The interior point itself is nmpc_qp_fused.cuh (one persistent CTA per OCP); the
// generic dense split-kernel IPM of the first round was removed (an unstructured external QP is refused with
// NMPC_ENOSYS instead of being routed to kernels no test covered).
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

// per-stage record of the per-problem vector workspace `vec` (doubles).  Grouped so that every pass of the
// interior point fetches what it needs with one or two bulk copies (all offsets and spans are multiples of 16 B):
//   F  (factorisation sweep)   QT GAM KFF RB   one span of 186
//   A1 / D1 (forward chain)    KFF RB          72
//   A2 / D2 (forward bulk)     DZ RD RM        240
//   R  (update + residuals)    DZ | DT DLAM | DPI of the previous stage
//   C1 (corrector bulk)        RG PRB | DT DLAM | RD
//   C2 (corrector chain)       M0
#define V_QT 0                           // 72  q~ = rg + C' gamma: linear term of the stage before G' w
#define V_GAM (V_QT + NZ)                // 42  Gamma = lam_l / t_l + lam_u / t_u
#define V_KFF (V_GAM + NG)               // 30  feed-forward input step  Luu^-T lu
#define V_RB (V_KFF + NU)                // 42  dynamics residual
#define V_DZ (V_RB + NX)                 // 72  primal direction
#define V_RD (V_DZ + NZ)                 // 84  slack residual
#define V_RM (V_RD + 2 * NG)             // 84  complementarity residual (affine, then corrector)
#define V_DT (V_RM + 2 * NG)             // 84
#define V_DLAM (V_DT + 2 * NG)           // 84
#define V_RG (V_DLAM + 2 * NG)           // 72  stationarity residual
#define V_PRB (V_RG + NZ)                // 42  P_{k+1} rb_k (kept from the factorisation for the corrector)
#define V_M0 (V_PRB + NX)                // 72  corrector: rg + C' gamma_c + G' (P+ rb)
#define V_P (V_M0 + NZ)                  // 42  corrector: linear term p_k of the cost-to-go
#define V_DPI (V_P + NX)                 // 42  direction of the dynamics multiplier pi_k
// time-opt block (enable_time_opt): everything the passes keep per stage about the node duration dt_k, its two
// bounds and its elimination from the stage (P~ = P+ - (P+ b)(P+ b)' / d)
#define V_X (V_DPI + NX)                 // 864
#define VX_PB 0                          // 42  P+ b
#define VX_DINV 42                       // 1 / d,  d = h_tt + Gamma_t + b' P+ b
#define VX_S 43                          // q~_t + b' p+ of the direction being solved for
#define VX_DTAU 44                       // direction of dt
#define VX_RG 45                         // stationarity residual of dt
#define VX_QT 46                         // q~_t = rg_t + gamma_t (affine)
#define VX_M0 47                         // corrector: rg_t + gamma_c
#define VX_GAM 48                        // Gamma_t
#define VX_RD 50                         // [2] slack residuals of the bounds
#define VX_RM 52                         // [2]
#define VX_DT 54                         // [2] slack directions
#define VX_DL 56                         // [2] multiplier directions
#define VX_SZ 58
#define V_STAGE (V_X + VX_SZ)            // 922
// QP iterate of dt per stage: [tau, -, lam_l, lam_u, t_l, t_u]
#define QX_SZ 6
static_assert(V_GAM % 2 == 0 && V_KFF % 2 == 0 && V_RB % 2 == 0 && V_DZ % 2 == 0 && V_RD % 2 == 0 && V_RM % 2 == 0 &&
              V_DT % 2 == 0 && V_DLAM % 2 == 0 && V_RG % 2 == 0 && V_PRB % 2 == 0 && V_M0 % 2 == 0 && V_P % 2 == 0 &&
              V_DPI % 2 == 0 && V_X % 2 == 0 && VX_SZ % 2 == 0 && V_STAGE % 2 == 0, "16-byte alignment of bulk copies");

#define POS_INF_BITS 0x7ff0000000000000ULL
#define ONE_BITS 0x3ff0000000000000ULL

struct QpArgs {
    double* qz; double* qpi; double* qlam; double* qt;   // QP iterate, global problem index
    double* qx;                     // time-opt: QP iterate of dt [B][N][QX_SZ]
    int32_t* qp_iter; int32_t* qp_status; int32_t* qp_ncond; double* qp_res;
    const double* dx0_ext;          // [B][NX] or null -> x0 - x[0]
    const double* x0;
    double* x; double* u; double* pi; double* lam;       // SQP iterate (updated by qp_apply)
    const int32_t* sqp_done; int32_t* sqp_iter;
    int N, b0, max_iter, warm, cond_pred_corr;
    double tol;
};

__device__ __forceinline__ void load6(const double* p, double* v)
{
    const double2 b0 = *reinterpret_cast<const double2*>(p);
    const double2 b1 = *reinterpret_cast<const double2*>(p + 2);
    const double2 b2 = *reinterpret_cast<const double2*>(p + 4);
    v[0] = b0.x; v[1] = b0.y; v[2] = b1.x; v[3] = b1.y; v[4] = b2.x; v[5] = b2.y;
}

