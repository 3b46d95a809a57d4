// nmpc_rollout.cuh -- batched rigid-body rollout: closes the loop around the resident MPC plans without MuJoCo
// (SURVEY.md 8f rank 4).
//
// Reference loop (one rollout at a time, in MuJoCo): every simulation step the controller turns the current plan
// into joint torques and the simulator advances the state --
//   LocomotionMPC.compute_torques_dof             mpc.py:503-599   (feed-forward inverse dynamics + PD on the plan)
//   Rollout_MPC / Rollout_combined_controller     Behavior_Cloning/utils/Rollout_MPC.py:465,
//                                                 DAgger/utils/Rollout_combined_controller.py:484-499
// Here one thread owns one rollout and runs `n_steps` simulation steps between two replanning ticks:
//   1. plan sample (Hermite (q, v), held (a, f)) at the step, torques tau = rnea(q, v, a_plan) - J' f_plan
//      + Kp (q_plan_next - q) + Kd (v_plan_next - v)                           (nmpc_plan.cuh, same arithmetic)
//   2. forward dynamics in the solver's Euler coordinates:  M(q) a = S' tau + sum_feet J_f' f_c - bias(q, v)
//      with the generated mass matrix / bias / foot Jacobians (nmpc_gen_sim.h) and a dense 18 x 18 Cholesky solve
//   3. ground contact: compliant point feet on the plane z = ground_z -- normal spring-damper (unilateral), viscous
//      tangential force capped by the friction cone.  MuJoCo's soft-constraint contact solver is NOT restated:
//      this is a penalty model with its own parameters.
//   4. semi-implicit Euler:  v += a dt,  q += v dt.
// The CPU checker (oracle/nmpc_oracle.c, oracle_sim_steps) runs the same arithmetic from the same generated code.
#pragma once
#include "nmpc_common.cuh"
#include "nmpc_plan.cuh"

struct SimParams {
    double sim_dt;          // simulation step [s]
    double kp, kd;          // PD gains of the low-level controller (MPCOptConfig.Kp / Kd)
    double k_n, d_n;        // normal contact stiffness [N/m] and damping [N s/m]
    double d_t;             // tangential (viscous) contact gain [N s/m]
    double mu;              // friction coefficient of the cone cap
    double ground_z;        // height of the ground plane
    double tau_max;         // joint torque limit (0: none)
};

struct SimArgs {
    const double* x; const double* u; const double* dtn;     // resident plans [B][(N+1)*42], [B][N*30], [B][N] or null
    double* qv;                 // [B][36] simulation state, advanced in place
    double* out_qv;             // [B][n_steps][36] state BEFORE each step (the measured state the controller saw) or null
    double* out_tau;            // [B][n_steps][12] applied torques or null
    double* out_fc;             // [B][n_steps][12] contact forces or null
    int32_t* diverged;          // [B] set when the state stops being finite
    int B, N, n_interp, step0, n_steps;
    double dt_nodes;
    SimParams p;
};

#define SIM_NQ 18

// in-place Cholesky solve of the packed lower-triangular SPD matrix M (18 x 18): M a = r
__device__ __forceinline__ void nmpc_sim_solve(double* M, double* r)
{
    for (int j = 0; j < SIM_NQ; j++) {
        double d = M[j * (j + 1) / 2 + j];
        for (int l = 0; l < j; l++) d -= M[j * (j + 1) / 2 + l] * M[j * (j + 1) / 2 + l];
        d = sqrt(d);
        M[j * (j + 1) / 2 + j] = d;
        const double inv = 1.0 / d;
        for (int i = j + 1; i < SIM_NQ; i++) {
            double s = M[i * (i + 1) / 2 + j];
            for (int l = 0; l < j; l++) s -= M[i * (i + 1) / 2 + l] * M[j * (j + 1) / 2 + l];
            M[i * (i + 1) / 2 + j] = s * inv;
        }
    }
    for (int i = 0; i < SIM_NQ; i++) {
        double s = r[i];
        for (int l = 0; l < i; l++) s -= M[i * (i + 1) / 2 + l] * r[l];
        r[i] = s / M[i * (i + 1) / 2 + i];
    }
    for (int i = SIM_NQ - 1; i >= 0; i--) {
        double s = r[i];
        for (int l = i + 1; l < SIM_NQ; l++) s -= M[l * (l + 1) / 2 + i] * r[l];
        r[i] = s / M[i * (i + 1) / 2 + i];
    }
}

// contact force of one point foot: position p, velocity v -> f (world frame)
__device__ __forceinline__ void nmpc_sim_contact(const SimParams& sp, const double* p, const double* v, double* f)
{
    f[0] = 0.0; f[1] = 0.0; f[2] = 0.0;
    const double pen = sp.ground_z - p[2];
    if (!(pen > 0.0)) return;
    double fn = sp.k_n * pen - sp.d_n * v[2];
    if (!(fn > 0.0)) return;                       // unilateral: the ground does not pull
    double fx = -sp.d_t * v[0], fy = -sp.d_t * v[1];
    const double ft = sqrt(fx * fx + fy * fy), cap = sp.mu * fn;
    if (ft > cap) { const double sc = cap / ft; fx *= sc; fy *= sc; }
    f[0] = fx; f[1] = fy; f[2] = fn;
}

// one physics step from joint torques tau[12]: q, v advanced in place; fc[12] = contact forces used
__device__ __forceinline__ void nmpc_sim_physics(const SimParams& sp, double* q, double* v, const double* tau, double* fc)
{
    double M[SIM_NQ * (SIM_NQ + 1) / 2], r[SIM_NQ], fp[12], fv[12], fJ[12 * SIM_NQ];
    nmpc_sim_mass(q, M);
    nmpc_sim_bias(q, v, r);
    nmpc_sim_feet(q, v, fp, fv, fJ);
    for (int i = 0; i < SIM_NQ; i++) r[i] = -r[i];
    for (int j = 0; j < 12; j++) r[6 + j] += tau[j];
    for (int ft = 0; ft < 4; ft++) {
        nmpc_sim_contact(sp, fp + 3 * ft, fv + 3 * ft, fc + 3 * ft);
        for (int c = 0; c < 3; c++) {
            const double fcc = fc[3 * ft + c];
            if (fcc != 0.0)
                for (int i = 0; i < SIM_NQ; i++) r[i] += fJ[(3 * ft + c) * SIM_NQ + i] * fcc;
        }
    }
    nmpc_sim_solve(M, r);
    for (int i = 0; i < SIM_NQ; i++) { v[i] += r[i] * sp.sim_dt; q[i] += v[i] * sp.sim_dt; }
}

#ifdef __CUDACC__
__global__ void __launch_bounds__(64, 2) nmpc_sim_kernel(SimArgs a)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= a.B) return;
    const int N = a.N, n = a.n_interp;
    const double* xb = a.x + (size_t)b * (N + 1) * NX;
    const double* ub = a.u + (size_t)b * N * NU;
    const double* dtn = a.dtn ? a.dtn + (size_t)b * N : nullptr;
    double q[NQ], v[NQ];
    for (int c = 0; c < NQ; c++) { q[c] = a.qv[(size_t)b * 36 + c]; v[c] = a.qv[(size_t)b * 36 + NQ + c]; }
    bool bad = false;
    for (int s = 0; s < a.n_steps; s++) {
        const int ip = a.step0 + s;
        double qr[NQ], vr[NQ], acc[NQ], ff[12], tq[12], fc[12];
        if (a.out_qv) for (int c = 0; c < NQ; c++) { a.out_qv[((size_t)b * a.n_steps + s) * 36 + c] = q[c]; a.out_qv[((size_t)b * a.n_steps + s) * 36 + NQ + c] = v[c]; }
        // controller: LocomotionMPC.compute_torques_dof on the measured state (see nmpc_plan_torque_kernel)
        nmpc_plan_sample(xb, ub, dtn, a.dt_nodes, N, n, ip + 1 < n ? ip + 1 : n - 1, qr, vr);
        const double* ur = ub + (size_t)nmpc_plan_hold_node(N, n, ip < n ? ip : n - 1) * NU;
        for (int c = 0; c < NQ; c++) acc[c] = ur[c];
        for (int c = 0; c < 12; c++) ff[c] = ur[NQ + c];
        nmpc_model_rnea(q, v, acc, ff, tq);
        for (int c = 0; c < 12; c++) {
            double t = tq[c] + a.p.kp * (qr[6 + c] - q[6 + c]) + a.p.kd * (vr[6 + c] - v[6 + c]);
            if (a.p.tau_max > 0.0) t = fmin(fmax(t, -a.p.tau_max), a.p.tau_max);
            tq[c] = t;
        }
        nmpc_sim_physics(a.p, q, v, tq, fc);
        if (a.out_tau) for (int c = 0; c < 12; c++) a.out_tau[((size_t)b * a.n_steps + s) * 12 + c] = tq[c];
        if (a.out_fc) for (int c = 0; c < 12; c++) a.out_fc[((size_t)b * a.n_steps + s) * 12 + c] = fc[c];
        for (int c = 0; c < NQ; c++) if (!isfinite(q[c]) || !isfinite(v[c])) bad = true;
        if (bad) break;
    }
    for (int c = 0; c < NQ; c++) { a.qv[(size_t)b * 36 + c] = q[c]; a.qv[(size_t)b * 36 + NQ + c] = v[c]; }
    if (bad && a.diverged) a.diverged[b] = 1;
}
#endif
