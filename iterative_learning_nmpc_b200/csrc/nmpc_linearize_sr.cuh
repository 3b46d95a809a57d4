// nmpc_linearize_sr.cuh -- linearisation straight into the structured stage record (SR) of the
// fused QP path (nmpc_qp_fused.cuh), plus the SR versions of the NLP residual and of the SQP step.
//
// Same arithmetic as nmpc_linearize_kernel (generated model record -> dynamics defect, momentum
// Jacobians, Gauss-Newton Hessian / gradient, constraint rows, bounds, masks), but only the 1692
// structured doubles of a stage are written instead of 11.6k doubles of mostly-zero dense blocks.
// The dense kernel stays for nmpc_export_qp and the generic dense QP path.
#pragma once
#include "nmpc_linearize.cuh"
#include "nmpc_qp_fused.cuh"
#include "nmpc_sqp.cuh"

// Small CTAs: the kernel is a chain of short, unevenly filled phases between barriers, so what counts is how many
// independent CTAs an SM holds (profiles/r1w_linearize.md: 256 threads x 4 CTAs 0.50 ms, 64 threads x 16 CTAs 0.30 ms).
#ifndef LIN_SR_THREADS
#define LIN_SR_THREADS 64
#endif
#ifndef LIN_MINB
#define LIN_MINB 16
#endif
static_assert(LIN_SR_THREADS >= 64 && LIN_SR_THREADS % 32 == 0, "warp 1 computes the lever arms");
__global__ void __launch_bounds__(LIN_SR_THREADS, LIN_MINB) nmpc_linearize_sr_kernel(LinArgs a, double* SR)
{
    const int k = blockIdx.x, bl = blockIdx.y, b = a.b0 + bl;
    const int tid = threadIdx.x;
    if (a.done && a.done[b]) return;
    const int N = a.N;
    const bool term = (k == N);

    __shared__ double rec[NMPC_REC_SIZE];
    __shared__ double s_x[NX], s_u[NU], s_xn[NX], s_p[NPAR], s_yr[NY], s_W[NY];
    __shared__ double s_n[NF][3], s_t1[NF][3], s_t2[NF][3], s_r[NF][3];
    __shared__ double s_cost[LIN_SR_THREADS / 32];

    const int ny = term ? NYE : NY;
    for (int i = tid; i < NX; i += LIN_SR_THREADS) {
        s_x[i] = a.x[((size_t)b * (N + 1) + k) * NX + i];
        s_xn[i] = term ? 0.0 : a.x[((size_t)b * (N + 1) + k + 1) * NX + i];
    }
    for (int i = tid; i < NU; i += LIN_SR_THREADS) s_u[i] = term ? 0.0 : a.u[((size_t)b * N + k) * NU + i];
    for (int i = tid; i < NPAR; i += LIN_SR_THREADS) s_p[i] = a.p[((size_t)b * (N + 1) + k) * NPAR + i];
    for (int i = tid; i < ny; i += LIN_SR_THREADS) {
        s_yr[i] = term ? a.yref_e[(size_t)b * NYE + i] : a.yref[((size_t)b * N + k) * NY + i];
        s_W[i] = term ? a.We[(size_t)b * NYE + i] : a.W[(size_t)b * NY + i];
    }
    double* S = SR + ((size_t)bl * (N + 1) + k) * SR_SZ;
    double* Dp = S + SR_D;
    {
        // the model record is gathered in the same round trip as the stage inputs
        const int inst = bl * (N + 1) + k;
        const double* rp = a.rec + (size_t)(inst / REC_BLK) * NMPC_REC_SIZE * REC_BLK + (inst % REC_BLK);
        for (int i = tid; i < NMPC_REC_SIZE; i += LIN_SR_THREADS) rec[i] = rp[(size_t)i * REC_BLK];
        // the record is small: zero it, then write the structured entries
        double2 z2 = make_double2(0.0, 0.0);
        for (int i = tid; i < SR_SZ / 2; i += LIN_SR_THREADS) reinterpret_cast<double2*>(S)[i] = z2;
    }
    __syncthreads();
    if (tid < NF) tangent_basis(&s_p[tid * NMPC_NP_FOOT + P_NORMAL], s_n[tid], s_t1[tid], s_t2[tid]);

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
    // enable_time_opt: the node duration is an optimisation variable with its own iterate
    const double dt = (a.topt && !term) ? a.dtn[(size_t)b * N + k] : a.dt;

    if (tid >= 32 && tid < 32 + NF * 3) s_r[(tid - 32) / 3][(tid - 32) % 3] = FP[tid - 32] - COM[(tid - 32) % 3];
    __syncthreads();

    // ---------------- dynamics block G = [B A b] ----------------
    if (!term) {
        for (int i = tid; i < NQ; i += LIN_SR_THREADS) {
            S[SR_B + i] = q[i] + dt * v[i] + 0.5 * dt * dt * acc_u[i] - s_xn[i];
            S[SR_B + NQ + i] = v[i] + dt * acc_u[i] - s_xn[NQ + i];
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
            S[SR_B + 2 * NQ + i] = hh[i] + dt * hd - s_xn[2 * NQ + i];
            if (a.topt) S[SR_X + SRX_B + 2 * NQ + i] = hd;              // d h+ / d dt
        }
        if (a.topt) {
            // d q+ / d dt = v + dt a,  d v+ / d dt = a; cost weight (dt - dt_nodes)^2 / 2; bounds dt_min <= dt <= dt_max
            for (int i = tid; i < NQ; i += LIN_SR_THREADS) {
                S[SR_X + SRX_B + i] = v[i] + dt * acc_u[i];
                S[SR_X + SRX_B + NQ + i] = acc_u[i];
            }
            if (tid == 0) {
                S[SR_X + SRX_H] = a.w_dt + a.reg;
                S[SR_X + SRX_G] = a.w_dt * (dt - a.dt);
                S[SR_X + SRX_LB] = a.dt_min - dt;
                S[SR_X + SRX_UB] = a.dt_max - dt;
                S[SR_X + SRX_DT] = dt;
                S[SR_X + SRX_C2] = 0.5 * dt * dt;
            }
        }
        // d h+/d f
        for (int e = tid; e < NF * 9; e += LIN_SR_THREADS) {
            const int ft = e / 9, i = (e % 9) / 3, j = e % 3;
            const double act = s_p[ft * NMPC_NP_FOOT + P_ACTIVE];
            const double* r = s_r[ft];
            if (i == j) S[SR_HF + 12 * i + 3 * ft + i] = dt * act;
            // [r]_x
            double rx = 0.0;
            if (i == 0 && j == 1) rx = -r[2]; else if (i == 0 && j == 2) rx = r[1];
            else if (i == 1 && j == 0) rx = r[2]; else if (i == 1 && j == 2) rx = -r[0];
            else if (i == 2 && j == 0) rx = -r[1]; else if (i == 2 && j == 1) rx = r[0];
            S[SR_HF + 12 * (3 + i) + 3 * ft + j] = dt * act * rx;
        }
        // d (angular part)/d q = dt sum_ft act * (-[f]_x (Jp - Jc))
        for (int e = tid; e < 3 * NQ; e += LIN_SR_THREADS) {
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
            S[SR_HQ + NQ * (3 + i) + j] = dt * s;
        }
    }

    // ---------------- Gauss-Newton Hessian / gradient block M ----------------
    const double reg = term ? a.reg_e : a.reg;
    const int yb = 0, yj = 12;                         // base, joint offsets (same in y and y_e)
    const int ysw = term ? 36 : 48, ypos = term ? 40 : 64;
    double cost = 0.0;
    // u diagonal + gradient
    for (int i = tid; i < NU; i += LIN_SR_THREADS) {
        double w = 0.0, e = 0.0;
        if (!term) {
            if (i >= 6 && i < NQ) { w = s_W[36 + i - 6]; e = acc_u[i] - s_yr[36 + i - 6]; }
            else if (i >= NQ) { w = s_W[52 + i - NQ]; e = s_u[i] - s_yr[52 + i - NQ]; }
        }
        S[SR_RD + i] = w + reg;
        S[SR_GR + i] = w * e;
        cost += 0.5 * w * e * e;
    }
    // v and h diagonal + gradient
    for (int i = tid; i < NQ + 6; i += LIN_SR_THREADS) {
        if (i < NQ) {
            const int yi = (i < 6) ? yb + 6 + i : yj + 12 + (i - 6);
            const double w = s_W[yi], e = v[i] - s_yr[yi];
            S[SR_QV + i] = w + reg;
            S[SR_GR + IV + i] = w * e;
            cost += 0.5 * w * e * e;
        } else {
            const int i6 = i - NQ;
            S[SR_QH + i6] = reg;
        }
    }
    // q block (dense symmetric 18x18) and q gradient
    for (int e = tid; e < NQ * NQ; e += LIN_SR_THREADS) {
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
        S[SR_QQ + NQ * i + j] = s;
    }
    for (int j = tid; j < NQ; j += LIN_SR_THREADS) {
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
        S[SR_GR + IQ + j] = g;
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
    if (a.topt && !term && tid == 0) cost += 0.5 * a.w_dt * (dt - a.dt) * (dt - a.dt);
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
        for (int i = 0; i < 3; i++) { S[SR_CF + 3 * row + i] = cvec[i]; cv += cvec[i] * fi[i]; }
        lg[row] = -cv; ug[row] = -cv;
        ml[row] = on ? 1.0 : 0.0;
    }
    if (k > 0) {
        // stance kinematics rows: [n.(p - pp)], [t1.v_foot], [t2.v_foot]; state-only
        for (int e = tid; e < NF * 3 * NQ; e += LIN_SR_THREADS) {
            const int ft = e / (3 * NQ), r = (e / NQ) % 3, j = e % NQ;
            const double* n = s_n[ft]; const double* t1 = s_t1[ft]; const double* t2 = s_t2[ft];
            const int row = GR_CNT + 3 * ft + r;
            if (r == 0) {
                double jn = 0.0;
                for (int i = 0; i < 3; i++) jn += n[i] * FJ[(3 * ft + i) * NQ + j];
                S[SR_CX + SR_CXLD * (row - NGF) + j] = jn;
            } else {
                const double* t = (r == 1) ? t1 : t2;
                double vq = 0.0, vv = 0.0;
                for (int i = 0; i < 3; i++) {
                    vq += t[i] * FVQ[(3 * ft + i) * NQ + j];
                    vv += t[i] * FJ[(3 * ft + i) * NQ + j];
                }
                S[SR_CX + SR_CXLD * (row - NGF) + j] = vq;
                S[SR_CX + SR_CXLD * (row - NGF) + NQ + j] = vv;
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
        for (int e = tid; e < NF * NQ; e += LIN_SR_THREADS) {
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
            S[SR_CX + SR_CXLD * (row - NGF) + j] = 2 * d1 * j1 + 2 * d2 * j2;
            if (j == 0) {
                const double cv = d1 * d1 + d2 * d2 - rad * rad;
                lg[row] = -cv; ug[row] = -cv; mu[row] = 1.0;
            }
        }
        // centroidal consistency rows h - h_model(q, v) = 0
        for (int e = tid; e < 6 * NQ; e += LIN_SR_THREADS) {
            const int i = e / NQ, j = e % NQ;
            const int row = GR_CONS + i;
            S[SR_CX + SR_CXLD * (row - NGF) + j] = -HQ[i * NQ + j];
            S[SR_CX + SR_CXLD * (row - NGF) + NQ + j] = -HV[i * NQ + j];
            if (j == 0) {
                const double cv = hh[i] - HM[i];
                lg[row] = -cv; ug[row] = -cv; ml[row] = 1.0; mu[row] = 1.0;
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        double c = 0.0;
        for (int w = 0; w < LIN_SR_THREADS / 32; w++) c += s_cost[w];
        a.cost_stage[(size_t)bl * (N + 1) + k] = c;
    }
}


// ---------------------------------------------------------------------------------------------
// inf-norms of the KKT residuals of the NLP at the current linearisation, from the stage record
// (SR twin of nmpc_residual_kernel): one CTA per node
// ---------------------------------------------------------------------------------------------
struct ResSrArgs {
    const double* SR;
    const double* x; const double* x0; const double* pi; const double* lam;
    unsigned long long* res_acc;
    const int32_t* done;
    int N, b0;
    double dt;
    const double* lamt; int topt;       // enable_time_opt: multipliers of the dt bounds [B][N][2]
};

__global__ void __launch_bounds__(RES_THREADS, 4) nmpc_residual_sr_kernel(ResSrArgs a)
{
    const int k = blockIdx.x, bl = blockIdx.y, b = a.b0 + bl, tid = threadIdx.x;
    if (a.done && a.done[b]) return;
    const int N = a.N;
    const bool term = (k == N);
    __shared__ double s_pik[NX], s_pin[NX], s_dl[NG], s_red[4][RES_THREADS / 32];
    const double* S = a.SR + ((size_t)bl * (N + 1) + k) * SR_SZ;
    const double* lam = a.lam + ((size_t)b * (N + 1) + k) * 2 * NG;
    const double dt = (a.topt && !term) ? S[SR_X + SRX_DT] : a.dt, c2 = 0.5 * dt * dt;
    double r_s = 0, r_e = 0, r_i = 0, r_c = 0;
    for (int i = tid; i < NX; i += RES_THREADS) {
        s_pik[i] = k > 0 ? a.pi[((size_t)b * N + k - 1) * NX + i] : 0.0;
        s_pin[i] = term ? 0.0 : a.pi[((size_t)b * N + k) * NX + i];
        if (k == 0) r_e = nanmax(r_e, fabs(a.x[(size_t)b * (N + 1) * NX + i] - a.x0[(size_t)b * NX + i]));
        if (!term) r_e = nanmax(r_e, fabs(S[SR_B + i]));
    }
    for (int j = tid; j < NG; j += RES_THREADS) {
        const double ml = S[SR_D + 2 * NG + j], mu = S[SR_D + 3 * NG + j];
        const double c = -S[SR_D + j];
        double dl = 0.0;
        if (ml > 0.5) { dl += lam[j]; r_i = nanmax(r_i, -c); r_c = nanmax(r_c, fabs(lam[j] * c)); }
        if (mu > 0.5) { dl -= lam[NG + j]; r_i = nanmax(r_i, c); r_c = nanmax(r_c, fabs(lam[NG + j] * c)); }
        s_dl[j] = dl;
    }
    __syncthreads();
    for (int c = tid; c < NZ; c += RES_THREADS) {
        if ((k == 0 && c >= NU) || (term && c < NU)) continue;
        double s = S[SR_GR + c];
        if (!term) s += fq_gt(S, s_pin, c, dt, c2);
        if (k > 0 && c >= NU) s -= s_pik[c - NU];
        s -= fq_ct(S, s_dl, c);
        r_s = nanmax(r_s, fabs(s));
    }
    if (a.topt && !term && tid == RES_THREADS - 1) {
        // the node duration: stationarity g + b' pi_k - (lam_l - lam_u), its two bounds, complementarity
        const double* lt = a.lamt + ((size_t)b * N + k) * 2;
        double s = S[SR_X + SRX_G] - (lt[0] - lt[1]);
        for (int l = 0; l < NX; l++) s += S[SR_X + SRX_B + l] * s_pin[l];
        r_s = nanmax(r_s, fabs(s));
        const double cl = -S[SR_X + SRX_LB], cu = -S[SR_X + SRX_UB];       // dt - dt_min, dt - dt_max
        r_i = nanmax(r_i, -cl); r_i = nanmax(r_i, cu);
        r_c = nanmax(r_c, fabs(lt[0] * cl)); r_c = nanmax(r_c, fabs(lt[1] * cu));
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

