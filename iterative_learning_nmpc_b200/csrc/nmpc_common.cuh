// nmpc_common.cuh -- dimensions, device-side layouts and sm_100a primitives (mbarrier + bulk TMA).
//
// Device layouts (all fp64, one contiguous record per problem so a CTA streams its own OCP):
//   iterate      x [B][(N+1)*42]  u [B][N*30]  pi [B][N*42]  lam [B][(N+1)*84]
//   problem data x0 [B][42]  p [B][(N+1)*44]  yref [B][N*76]  yref_e [B][52]  W [B][76]  We [B][52]
//   QP stage block (per problem, per node; "k-major" so rank-k updates read contiguous rows):
//     G  [42][74]   row l = [B(l,0:30) | A(l,0:42) | b(l) | pad]        (BAbt^T, ld 74)
//     M  [73][73]+1 full symmetric Hessian, row 72 = gradient             (RSQrq,  ld 73)
//     C  [42][72]   row j = constraint j over z = [u;x]                  (DCt^T,  ld 72)
//     lg [42] ug [42] mask [84]
//   factor block  L [2880] panel-packed lower Cholesky factor of M_k + row 72 = L^-1 m
//                 (twelve 6-column panels, nmpc_qp_layout.cuh)
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/nmpc_b200.h"

#define NX NMPC_NX
#define NU NMPC_NU
#define NZ NMPC_NZ
#define NPAR NMPC_NP
#define NY NMPC_NY
#define NYE NMPC_NYE
#define NG NMPC_NG
#define NQ 18
#define NF 4

// z = [u; x] offsets
#define IA 0
#define IFRC 18
#define IQ 30
#define IV 48
#define IH 66
// constraint row blocks
#define GR_FRIC 0
#define GR_CNT 20
#define GR_PATCH 32
#define GR_CONS 36
// per-foot parameter offsets
#define P_ACTIVE 0
#define P_POINT 1
#define P_NORMAL 4
#define P_GAIN 7
#define P_RESTRICT 8
#define P_RADIUS 9
#define P_PEAK 10

// stage-block geometry (doubles)
#define G_LD 74
#define G_SZ (NX * G_LD)            // 3108
#define M_LD 73
#define M_SZ (73 * 73 + 1)          // 5330 (padded to a multiple of 16 bytes)
#define C_LD 72
#define C_SZ (NG * C_LD)            // 3024
#define D_SZ (NG + NG + 2 * NG)     // lg | ug | mask = 168
#define L_SZ M_SZ

// IPM constants (HPIPM `speed` defaults as recalled in SURVEY.md 8a)
#define IPM_MU0 1e1
#define IPM_THR0 1e-1
#define IPM_REG_PRIM 1e-15
#define IPM_LAM_MIN 1e-16
#define IPM_T_MIN 1e-16
#define IPM_ALPHA_MIN 1e-12
#define IPM_TAU_MIN 1e-16


// ---------------------------------------------------------------------------------------------
// mbarrier + bulk async copy (TMA 1-D form; SASS: UBLKCP / SYNCS)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

__device__ __forceinline__ void fence_barrier_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

// make generic-proxy smem writes visible to the async proxy (before a bulk store reads them)
__device__ __forceinline__ void fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem),
                 "r"(smem_u32(src_smem)), "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

// wait until the bulk stores issued by this thread have finished READING shared memory
__device__ __forceinline__ void bulk_store_wait_read()
{
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

__device__ __forceinline__ void bulk_store_wait_all()
{
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// Spin on the mbarrier phase with a bound: a mismatch between expect_tx and the bytes actually
// copied would otherwise hang the GPU; trap instead so the error surfaces as a launch failure.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    uint32_t done = 0;
    for (uint32_t spin = 0; spin < (1u << 28); ++spin) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
        if (done) return;
    }
    __trap();
}

// ---------------------------------------------------------------------------------------------
// small block-level helpers
// ---------------------------------------------------------------------------------------------
// max that keeps a NaN once it has seen one (fmax drops NaNs; the norms must report them)
__device__ __forceinline__ double nanmax(double a, double b) { return (b > a || b != b) ? b : a; }
__device__ __forceinline__ double warp_max(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = nanmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_min(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// tangent basis of a contact plane from its (not necessarily unit) normal; same construction as
// the oracle: t1 = normalised projection of e_x (e_y if the normal is nearly along x), t2 = n x t1
__device__ __forceinline__ void tangent_basis(const double* n_in, double* n, double* t1, double* t2)
{
    double nn = sqrt(n_in[0] * n_in[0] + n_in[1] * n_in[1] + n_in[2] * n_in[2]);
    if (nn < 1e-12) { n[0] = 0; n[1] = 0; n[2] = 1; }
    else { n[0] = n_in[0] / nn; n[1] = n_in[1] / nn; n[2] = n_in[2] / nn; }
    double e0 = 1, e1 = 0, e2 = 0;
    if (fabs(n[0]) > 0.9) { e0 = 0; e1 = 1; }
    double d = e0 * n[0] + e1 * n[1] + e2 * n[2];
    t1[0] = e0 - d * n[0]; t1[1] = e1 - d * n[1]; t1[2] = e2 - d * n[2];
    double tn = sqrt(t1[0] * t1[0] + t1[1] * t1[1] + t1[2] * t1[2]);
    t1[0] /= tn; t1[1] /= tn; t1[2] /= tn;
    t2[0] = n[1] * t1[2] - n[2] * t1[1];
    t2[1] = n[2] * t1[0] - n[0] * t1[2];
    t2[2] = n[0] * t1[1] - n[1] * t1[0];
}
