// nmpc_linearize.cuh -- stage linearisation: generated model code + QP stage assembly.
//
// Replaces, for every node of every problem, what acados does per SQP iteration with the
// CasADi-generated functions of the reference's model (SURVEY.md 3.2c "linearise all N+1
// stages"; model declared at mpc_controller/utils/dynamics.py:109-134 and in the missing
// contact_tamp FloatingBaseDynamics/PointContact): multiple-shooting defect and its Jacobians,
// Gauss-Newton Hessian/gradient of the least-squares cost, constraint values and Jacobians.
//
// Two kernels:
//  1. nmpc_model_kernel: ONE THREAD per (stage instance, model part) runs the generated
//     straight-line kinematics (feet, CoM, centroidal momentum and their Jacobians; 21k flops per
//     instance in 8 independent parts).  A block is 128 consecutive instances of one part, so
//     control flow is uniform and the record is written structure-of-arrays, blocked by 128
//     instances: rec[(inst/128)][slot][inst%128] -- every store of a warp is one 256-byte line.
//  2. nmpc_linearize_kernel: one CTA per stage instance gathers its record and scatters the
//     structured non-zeros of the dense stage blocks G, M, C, D the Riccati kernel streams.
// Formulation: DESIGN.md "The OCP".
#pragma once
#include "nmpc_common.cuh"

#define REC_BLK 128                                   // instances per SoA block of the record
#define NMPC_HD __device__ __noinline__
#define NMPC_REC(i) rec[(size_t)(i) * REC_BLK]
#define NMPC_WITH_RNEA 1
#include "nmpc_gen_model.h"

#define LIN_THREADS 256
#define MODEL_THREADS REC_BLK

struct ModelArgs {
    const double* x;          // [B][(N+1)*42]
    double* rec;              // [ceil(n_inst/128)][NMPC_REC_SIZE][128]
    const int32_t* done;      // [B] or null
    int n_inst, N, b0;
};

__global__ void __launch_bounds__(MODEL_THREADS, 2) nmpc_model_kernel(ModelArgs a)
{
    const int i = blockIdx.x * MODEL_THREADS + threadIdx.x;
    if (i >= a.n_inst) return;
    const int bl = i / (a.N + 1);
    if (a.done && a.done[a.b0 + bl]) return;
    const double* xp = a.x + ((size_t)a.b0 * (a.N + 1) + i) * NX;
    double q[NQ], v[NQ];
#pragma unroll
    for (int j = 0; j < NQ; j++) { q[j] = xp[j]; v[j] = xp[NQ + j]; }
    double* rec = a.rec + (size_t)blockIdx.x * NMPC_REC_SIZE * REC_BLK + threadIdx.x;
    nmpc_model_part(blockIdx.y, q, v, rec);
}

struct LinArgs {
    const double* x; const double* u; const double* p;
    const double* yref; const double* yref_e; const double* W; const double* We;
    double* G; double* M; double* C; double* D;
    double* cost_stage;      // [Bc][N+1]
    const double* rec;       // model record, blocked SoA (nmpc_model_kernel)
    const int32_t* done;     // [B] or null
    int N; int b0;
    double dt, reg, reg_e, mu_c;
    // enable_time_opt (stage-record kernel only): per-node durations [B][N], weight, bounds
    const double* dtn; int topt; double w_dt, dt_min, dt_max;
};

__global__ void __launch_bounds__(LIN_THREADS, 4) nmpc_linearize_kernel(LinArgs a)
{
    const int k = blockIdx.x, bl = blockIdx.y, b = a.b0 + bl;
    const int tid = threadIdx.x;
    if (a.done && a.done[b]) return;
    const int N = a.N;
    const bool term = (k == N);

    __shared__ double rec[NMPC_REC_SIZE];
    __shared__ double s_x[NX], s_u[NU], s_xn[NX], s_p[NPAR], s_yr[NY], s_W[NY];
    __shared__ double s_n[NF][3], s_t1[NF][3], s_t2[NF][3], s_r[NF][3];
    __shared__ double s_cost[LIN_THREADS / 32];

    const int ny = term ? NYE : NY;
    for (int i = tid; i < NX; i += LIN_THREADS) {
        s_x[i] = a.x[((size_t)b * (N + 1) + k) * NX + i];
        s_xn[i] = term ? 0.0 : a.x[((size_t)b * (N + 1) + k + 1) * NX + i];
    }
    for (int i = tid; i < NU; i += LIN_THREADS) s_u[i] = term ? 0.0 : a.u[((size_t)b * N + k) * NU + i];
    for (int i = tid; i < NPAR; i += LIN_THREADS) s_p[i] = a.p[((size_t)b * (N + 1) + k) * NPAR + i];
    for (int i = tid; i < ny; i += LIN_THREADS) {
        s_yr[i] = term ? a.yref_e[(size_t)b * NYE + i] : a.yref[((size_t)b * N + k) * NY + i];
        s_W[i] = term ? a.We[(size_t)b * NYE + i] : a.W[(size_t)b * NY + i];
    }
    __syncthreads();

    double* Gp = term ? nullptr : a.G + ((size_t)bl * N + k) * G_SZ;
    double* Mp = a.M + ((size_t)bl * (N + 1) + k) * M_SZ;
    double* Cp = a.C + ((size_t)bl * (N + 1) + k) * C_SZ;
    double* Dp = a.D + ((size_t)bl * (N + 1) + k) * D_SZ;

    {
        const int inst = bl * (N + 1) + k;
        const double* rp = a.rec + (size_t)(inst / REC_BLK) * NMPC_REC_SIZE * REC_BLK + (inst % REC_BLK);
        for (int i = tid; i < NMPC_REC_SIZE; i += LIN_THREADS) rec[i] = rp[(size_t)i * REC_BLK];
        if (tid < NF) tangent_basis(&s_p[tid * NMPC_NP_FOOT + P_NORMAL], s_n[tid], s_t1[tid], s_t2[tid]);
        // dense blocks: zero-fill with 16-byte stores, then scatter the structured non-zeros
        double2 z2 = make_double2(0.0, 0.0);
        if (Gp) for (int i = tid; i < G_SZ / 2; i += LIN_THREADS) reinterpret_cast<double2*>(Gp)[i] = z2;
        for (int i = tid; i < M_SZ / 2; i += LIN_THREADS) reinterpret_cast<double2*>(Mp)[i] = z2;
        for (int i = tid; i < C_SZ / 2; i += LIN_THREADS) reinterpret_cast<double2*>(Cp)[i] = z2;
        for (int i = tid; i < D_SZ / 2; i += LIN_THREADS) reinterpret_cast<double2*>(Dp)[i] = z2;
    }
    __syncthreads();

    const double* q = s_x;
    const double* v = s_x + NQ;
    const double* hh = s_x + 2 * NQ;
    const double* acc_u = s_u;
    const double* f = s_u + NQ;
    const double* FP = rec + NMPC_REC_FOOT_P;
    const double* FJ = rec + NMPC_REC_FOOT_J;
    const double* FV = rec + NMPC_REC_FOOT_V;
    const double* FVQ = rec + NMPC_REC_FOOT_VQ;
    const double* COM = rec + NMPC_REC_COM;
    const double* CJ = rec + NMPC_REC_COM_J;
    const double* HM = rec + NMPC_REC_H;
    const double* HQ = rec + NMPC_REC_H_Q;
    const double* HV = rec + NMPC_REC_H_V;
    const double dt = a.dt;

    if (tid < NF * 3) s_r[tid / 3][tid % 3] = FP[tid] - COM[tid % 3];
    __syncthreads();

    // ---------------- dynamics block G = [B A b] ----------------
    if (!term) {
        for (int i = tid; i < NQ; i += LIN_THREADS) {
            double* r0 = Gp + i * G_LD;
            r0[IA + i] = 0.5 * dt * dt; r0[IQ + i] = 1.0; r0[IV + i] = dt;
            r0[NZ] = q[i] + dt * v[i] + 0.5 * dt * dt * acc_u[i] - s_xn[i];
            double* r1 = Gp + (NQ + i) * G_LD;
            r1[IA + i] = dt; r1[IV + i] = 1.0;
            r1[NZ] = v[i] + dt * acc_u[i] - s_xn[NQ + i];
        }
        if (tid < 6) {
            const int i = tid;
            double hd = (i == 2) ? NMPC_MASS * (-9.81) : 0.0;
            for (int ft = 0; ft < NF; ft++) {
                const double act = s_p[ft * NMPC_NP_FOOT + P_ACTIVE];
                const double* fi = f + 3 * ft;
                const double* r = s_r[ft];
                double val;
                if (i < 3) val = fi[i];
                else if (i == 3) val = r[1] * fi[2] - r[2] * fi[1];
                else if (i == 4) val = r[2] * fi[0] - r[0] * fi[2];
                else val = r[0] * fi[1] - r[1] * fi[0];
                hd += act * val;
            }
            double* row = Gp + (2 * NQ + i) * G_LD;
            row[IH + i] = 1.0;
            row[NZ] = hh[i] + dt * hd - s_xn[2 * NQ + i];
        }
        // d h+/d f
        for (int e = tid; e < NF * 9; e += LIN_THREADS) {
            const int ft = e / 9, i = (e % 9) / 3, j = e % 3;
            const double act = s_p[ft * NMPC_NP_FOOT + P_ACTIVE];
            const double* r = s_r[ft];
            if (i == j) Gp[(2 * NQ + i) * G_LD + IFRC + 3 * ft + i] = dt * act;
            // [r]_x
            double rx = 0.0;
            if (i == 0 && j == 1) rx = -r[2]; else if (i == 0 && j == 2) rx = r[1];
            else if (i == 1 && j == 0) rx = r[2]; else if (i == 1 && j == 2) rx = -r[0];
            else if (i == 2 && j == 0) rx = -r[1]; else if (i == 2 && j == 1) rx = r[0];
            Gp[(2 * NQ + 3 + i) * G_LD + IFRC + 3 * ft + j] = dt * act * rx;
        }
        // d (angular part)/d q = dt sum_ft act * (-[f]_x (Jp - Jc))
        for (int e = tid; e < 3 * NQ; e += LIN_THREADS) {
            const int i = e / NQ, j = e % NQ;
            double s = 0.0;
            for (int ft = 0; ft < NF; ft++) {
                const double act = s_p[ft * NMPC_NP_FOOT + P_ACTIVE];
                const double* fi = f + 3 * ft;
                const double d0 = FJ[(3 * ft + 0) * NQ + j] - CJ[0 * NQ + j];
                const double d1 = FJ[(3 * ft + 1) * NQ + j] - CJ[1 * NQ + j];
                const double d2 = FJ[(3 * ft + 2) * NQ + j] - CJ[2 * NQ + j];
                // -[f]_x d = d x f
                double val;
                if (i == 0) val = d1 * fi[2] - d2 * fi[1];
                else if (i == 1) val = d2 * fi[0] - d0 * fi[2];
                else val = d0 * fi[1] - d1 * fi[0];
                s += act * val;
            }
            Gp[(2 * NQ + 3 + i) * G_LD + IQ + j] = dt * s;
        }
    }

    // ---------------- Gauss-Newton Hessian / gradient block M ----------------
    const double reg = term ? a.reg_e : a.reg;
    const int yb = 0, yj = 12;                         // base, joint offsets (same in y and y_e)
    const int ysw = term ? 36 : 48, ypos = term ? 40 : 64;
    double cost = 0.0;
    // u diagonal + gradient
    for (int i = tid; i < NU; i += LIN_THREADS) {
        double w = 0.0, e = 0.0;
        if (!term) {
            if (i >= 6 && i < NQ) { w = s_W[36 + i - 6]; e = acc_u[i] - s_yr[36 + i - 6]; }
            else if (i >= NQ) { w = s_W[52 + i - NQ]; e = s_u[i] - s_yr[52 + i - NQ]; }
        }
        Mp[i * M_LD + i] = w + reg;
        Mp[NZ * M_LD + i] = w * e;
        cost += 0.5 * w * e * e;
    }
    // v and h diagonal + gradient
    for (int i = tid; i < NQ + 6; i += LIN_THREADS) {
        if (i < NQ) {
            const int yi = (i < 6) ? yb + 6 + i : yj + 12 + (i - 6);
            const double w = s_W[yi], e = v[i] - s_yr[yi];
            Mp[(IV + i) * M_LD + IV + i] = w + reg;
            Mp[NZ * M_LD + IV + i] = w * e;
            cost += 0.5 * w * e * e;
        } else {
            const int i6 = i - NQ;
            Mp[(IH + i6) * M_LD + IH + i6] = reg;
        }
    }
    // q block (dense symmetric 18x18) and q gradient
    for (int e = tid; e < NQ * NQ; e += LIN_THREADS) {
        const int i = e / NQ, j = e % NQ;      // both triangles: the stage kernel multiplies by rows
        double s = 0.0;
        for (int ft = 0; ft < NF; ft++) {
            const double pk = s_p[ft * NMPC_NP_FOOT + P_PEAK];
            const double wsw = s_W[ysw + ft] * pk * pk;
            const double* Jz = FJ + (3 * ft + 2) * NQ;
            s += wsw * Jz[i] * Jz[j];
            for (int d = 0; d < 3; d++) {
                const double wp = s_W[ypos + 3 * ft + d];
                const double* Jd = FJ + (3 * ft + d) * NQ;
                s += wp * Jd[i] * Jd[j];
            }
        }
        if (i == j) {
            const int yi = (i < 6) ? yb + i : yj + (i - 6);
            s += s_W[yi] + reg;
        }
        Mp[(IQ + i) * M_LD + IQ + j] = s;
    }
    for (int j = tid; j < NQ; j += LIN_THREADS) {
        const int yi = (j < 6) ? yb + j : yj + (j - 6);
        const double e0 = q[j] - s_yr[yi];
        double g = s_W[yi] * e0;
        cost += 0.5 * s_W[yi] * e0 * e0;
        for (int ft = 0; ft < NF; ft++) {
            const double pk = s_p[ft * NMPC_NP_FOOT + P_PEAK];
            const double esw = pk * FP[3 * ft + 2] - s_yr[ysw + ft];
            g += s_W[ysw + ft] * esw * pk * FJ[(3 * ft + 2) * NQ + j];
            for (int d = 0; d < 3; d++) {
                const double ep = FP[3 * ft + d] - s_yr[ypos + 3 * ft + d];
                g += s_W[ypos + 3 * ft + d] * ep * FJ[(3 * ft + d) * NQ + j];
            }
        }
        Mp[NZ * M_LD + IQ + j] = g;
    }
    if (tid < NF) {
        const int ft = tid;
        const double pk = s_p[ft * NMPC_NP_FOOT + P_PEAK];
        const double esw = pk * FP[3 * ft + 2] - s_yr[ysw + ft];
        cost += 0.5 * s_W[ysw + ft] * esw * esw;
        for (int d = 0; d < 3; d++) {
            const double ep = FP[3 * ft + d] - s_yr[ypos + 3 * ft + d];
            cost += 0.5 * s_W[ypos + 3 * ft + d] * ep * ep;
        }
    }
    cost = warp_sum(cost);
    if ((tid & 31) == 0) s_cost[tid >> 5] = cost;

    // ---------------- constraints C, bounds, masks ----------------
    double* lg = Dp;
    double* ug = Dp + NG;
    double* ml = Dp + 2 * NG;
    double* mu = Dp + 3 * NG;
    if (!term && tid < NF * 5) {
        const int ft = tid / 5, r = tid % 5;
        const double* n = s_n[ft]; const double* t1 = s_t1[ft]; const double* t2 = s_t2[ft];
        const double* fi = f + 3 * ft;
        const bool on = s_p[ft * NMPC_NP_FOOT + P_ACTIVE] > 0.5;
        double cvec[3];
        for (int i = 0; i < 3; i++) {
            if (r == 0) cvec[i] = a.mu_c * n[i] - t1[i];
            else if (r == 1) cvec[i] = a.mu_c * n[i] + t1[i];
            else if (r == 2) cvec[i] = a.mu_c * n[i] - t2[i];
            else if (r == 3) cvec[i] = a.mu_c * n[i] + t2[i];
            else cvec[i] = n[i];
        }
        const int row = GR_FRIC + tid;
        double cv = 0.0;
        for (int i = 0; i < 3; i++) { Cp[row * C_LD + IFRC + 3 * ft + i] = cvec[i]; cv += cvec[i] * fi[i]; }
        lg[row] = -cv; ug[row] = -cv;
        ml[row] = on ? 1.0 : 0.0;
    }
    if (k > 0) {
        // stance kinematics rows: [n.(p - pp)], [t1.v_foot], [t2.v_foot]; state-only
        for (int e = tid; e < NF * 3 * NQ; e += LIN_THREADS) {
            const int ft = e / (3 * NQ), r = (e / NQ) % 3, j = e % NQ;
            const double* n = s_n[ft]; const double* t1 = s_t1[ft]; const double* t2 = s_t2[ft];
            const int row = GR_CNT + 3 * ft + r;
            if (r == 0) {
                double jn = 0.0;
                for (int i = 0; i < 3; i++) jn += n[i] * FJ[(3 * ft + i) * NQ + j];
                Cp[row * C_LD + IQ + j] = jn;
            } else {
                const double* t = (r == 1) ? t1 : t2;
                double vq = 0.0, vv = 0.0;
                for (int i = 0; i < 3; i++) {
                    vq += t[i] * FVQ[(3 * ft + i) * NQ + j];
                    vv += t[i] * FJ[(3 * ft + i) * NQ + j];
                }
                Cp[row * C_LD + IQ + j] = vq;
                Cp[row * C_LD + IV + j] = vv;
            }
        }
        if (tid < NF * 3) {
            const int ft = tid / 3, r = tid % 3;
            const double* pf = &s_p[ft * NMPC_NP_FOOT];
            const bool on = pf[P_ACTIVE] > 0.5;
            const double* n = s_n[ft]; const double* t1 = s_t1[ft]; const double* t2 = s_t2[ft];
            double cv;
            if (r == 0) {
                cv = 0.0;
                for (int i = 0; i < 3; i++) cv += n[i] * (FP[3 * ft + i] - pf[P_POINT + i]);
            } else {
                const double* t = (r == 1) ? t1 : t2;
                cv = t[0] * FV[3 * ft] + t[1] * FV[3 * ft + 1] + t[2] * FV[3 * ft + 2];
            }
            const int row = GR_CNT + tid;
            lg[row] = -cv; ug[row] = -cv;
            ml[row] = on ? 1.0 : 0.0; mu[row] = on ? 1.0 : 0.0;
        }
        // contact patch rows
        for (int e = tid; e < NF * NQ; e += LIN_THREADS) {
            const int ft = e / NQ, j = e % NQ;
            const double* pf = &s_p[ft * NMPC_NP_FOOT];
            const double rad = pf[P_RADIUS];
            const bool restricted = pf[P_ACTIVE] > 0.5 && pf[P_RESTRICT] > 0.5 && rad < 1e9;
            if (!restricted) continue;
            const double* t1 = s_t1[ft]; const double* t2 = s_t2[ft];
            double d1 = 0, d2 = 0, j1 = 0, j2 = 0;
            for (int i = 0; i < 3; i++) {
                const double d = FP[3 * ft + i] - pf[P_POINT + i];
                d1 += t1[i] * d; d2 += t2[i] * d;
                j1 += t1[i] * FJ[(3 * ft + i) * NQ + j];
                j2 += t2[i] * FJ[(3 * ft + i) * NQ + j];
            }
            const int row = GR_PATCH + ft;
            Cp[row * C_LD + IQ + j] = 2 * d1 * j1 + 2 * d2 * j2;
            if (j == 0) {
                const double cv = d1 * d1 + d2 * d2 - rad * rad;
                lg[row] = -cv; ug[row] = -cv; mu[row] = 1.0;
            }
        }
        // centroidal consistency rows h - h_model(q, v) = 0
        for (int e = tid; e < 6 * NQ; e += LIN_THREADS) {
            const int i = e / NQ, j = e % NQ;
            const int row = GR_CONS + i;
            Cp[row * C_LD + IQ + j] = -HQ[i * NQ + j];
            Cp[row * C_LD + IV + j] = -HV[i * NQ + j];
            if (j == 0) {
                Cp[row * C_LD + IH + i] = 1.0;
                const double cv = hh[i] - HM[i];
                lg[row] = -cv; ug[row] = -cv; ml[row] = 1.0; mu[row] = 1.0;
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        double c = 0.0;
        for (int w = 0; w < LIN_THREADS / 32; w++) c += s_cost[w];
        a.cost_stage[(size_t)bl * (N + 1) + k] = c;
    }
}

// ---------------------------------------------------------------------------------------------
// values-only kinematics: feet, CoM, centroidal momentum (update_pin / get_feet_position_w)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(64, 4) nmpc_kin0_kernel(const double* qv, int n, double* feet, double* com, double* hg)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double q[NQ], v[NQ], f[12], c[3], h[6];
    for (int j = 0; j < NQ; j++) { q[j] = qv[(size_t)i * 36 + j]; v[j] = qv[(size_t)i * 36 + NQ + j]; }
    nmpc_model_kin0(q, v, f, c, h);
    if (feet) for (int j = 0; j < 12; j++) feet[(size_t)i * 12 + j] = f[j];
    if (com) for (int j = 0; j < 3; j++) com[(size_t)i * 3 + j] = c[j];
    if (hg) for (int j = 0; j < 6; j++) hg[(size_t)i * 6 + j] = h[j];
}

// (launch bounds on the kernels that call the generated, non-inlined model functions: without a minimum-blocks hint
// ptxas gave them 32 registers and the straight-line code spilled 20-100 KB per thread)
// feed-forward joint torques (QuadrupedDynamics.id_torques, dynamics.py:136-163): thread per instance
__global__ void __launch_bounds__(64, 4) nmpc_rnea_kernel(const double* qva, const double* f, int n, double* tau)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double q[NQ], v[NQ], a[NQ], ff[12], t[NQ - 6];
    for (int j = 0; j < NQ; j++) {
        q[j] = qva[(size_t)i * 54 + j]; v[j] = qva[(size_t)i * 54 + NQ + j]; a[j] = qva[(size_t)i * 54 + 2 * NQ + j];
    }
    for (int j = 0; j < 12; j++) ff[j] = f[(size_t)i * 12 + j];
    nmpc_model_rnea(q, v, a, ff, t);
    for (int j = 0; j < NQ - 6; j++) tau[(size_t)i * (NQ - 6) + j] = t[j];
}

// x0 = [q, v, hg(q, v)]; also overwrites node 0 of the iterate (solver.py:179-192)
__global__ void __launch_bounds__(64, 4) nmpc_set_x0_kernel(const double* qv, int B, int N, double* x0, double* x)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    double q[NQ], v[NQ], f[12], c[3], h[6];
    for (int j = 0; j < NQ; j++) { q[j] = qv[(size_t)b * 36 + j]; v[j] = qv[(size_t)b * 36 + NQ + j]; }
    nmpc_model_kin0(q, v, f, c, h);
    double* o0 = x0 + (size_t)b * NX;
    double* o1 = x + (size_t)b * (N + 1) * NX;
    for (int j = 0; j < NQ; j++) { o0[j] = q[j]; o0[NQ + j] = v[j]; o1[j] = q[j]; o1[NQ + j] = v[j]; }
    for (int j = 0; j < 6; j++) { o0[2 * NQ + j] = h[j]; o1[2 * NQ + j] = h[j]; }
}

__global__ void nmpc_copy_x0_kernel(int B, int N, const double* x0, double* x)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * NX) return;
    const int b = i / NX, j = i % NX;
    x[(size_t)b * (N + 1) * NX + j] = x0[i];
}

// setup_initial_feet_pos (solver.py:194-210)
__global__ void __launch_bounds__(64, 4) nmpc_pin_feet_kernel(int B, int N, const double* x0, const int32_t* i_node, double* p)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    double q[NQ], v[NQ], feet[12], c[3], h[6];
    for (int j = 0; j < NQ; j++) { q[j] = x0[(size_t)b * NX + j]; v[j] = x0[(size_t)b * NX + NQ + j]; }
    nmpc_model_kin0(q, v, feet, c, h);
    double* pb = p + (size_t)b * (N + 1) * NPAR;
    for (int ft = 0; ft < NF; ft++) {
        if (i_node[b] == 0) pb[ft * NMPC_NP_FOOT + P_ACTIVE] = 1.0;
        if (pb[ft * NMPC_NP_FOOT + P_ACTIVE] != 0.0) {
            int next_swing = 0;
            double best = pb[ft * NMPC_NP_FOOT + P_ACTIVE];
            for (int k = 1; k <= N; k++) {
                const double av = pb[k * NPAR + ft * NMPC_NP_FOOT + P_ACTIVE];
                if (av < best) { best = av; next_swing = k; }
            }
            for (int k = 0; k < next_swing; k++)
                for (int i = 0; i < 3; i++) pb[k * NPAR + ft * NMPC_NP_FOOT + P_POINT + i] = feet[3 * ft + i];
        }
    }
}
