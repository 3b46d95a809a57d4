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

#define POS_INF_BITS 0x7ff0000000000000ULL
#define ONE_BITS 0x3ff0000000000000ULL

struct QpArgs {
    double* qz; double* qpi; double* qlam; double* qt;   // QP iterate, global problem index
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

