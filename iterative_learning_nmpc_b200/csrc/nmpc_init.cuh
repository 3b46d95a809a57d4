// nmpc_init.cuh -- QuadrupedAcadosSolver.init on the device (8f rank 2).
//
// In the reference, `init` expands a handful of per-problem planner outputs -- base references, the
// contact mask, swing peaks, optional contact locations -- into per-node dictionaries and pushes them
// into acados node by node (`update_solver`, solver.py:344-353: ~40 KB per solve through Python).
// Here the compact planner outputs cross PCIe (0.5 KB per problem instead of 41 KB) and the expansion
// into the node-major parameter / reference records is done by the kernels below:
//   setup_reference           solver.py:153-177   -> yref [B,N,76], yref_e [B,52]
//   init_contacts_parameters  solver.py:212-225 } -> p [B,N+1,44]
//   setup_cnt_status          solver.py:227-252 }
//   setup_contact_loc         solver.py:254-275   -> plane_point in p, foot targets in yref / yref_e
// plus the cold first guess the batched front-end uses (x0 held over the horizon, the robot's weight
// shared by the stance feet of each node).
#pragma once
#include "nmpc_common.cuh"

// y layout (include/nmpc_b200.h): base 12 | joint 24 | acc 12 | swing 4 | f_reg 12 | foot pos 12
#define Y_BASE 0
#define Y_JOINT 12
#define Y_SWING 48
#define Y_POS 64
#define YE_BASE 0
#define YE_JOINT 12
#define YE_SWING 36
#define YE_POS 40

struct TickArgs {
    const double* base_ref;      // [B,12]
    const double* base_ref_e;    // [B,12]
    const double* joint_ref;     // [B,12] or [12]
    const double* step_height;   // [B] or [1]
    const signed char* cnt;      // [B,4,N+1]
    const signed char* peak;     // [B,4,N+1] or nullptr (opt_peak off)
    const double* loc;           // [B,4,N+1,3] or nullptr
    double p_gain[NF];
    double radius, height_offset;
    int restrict_cnt, joint_per_problem, height_per_problem;
    double* p; double* yref; double* yref_e;
    int B, N;
};

__global__ void __launch_bounds__(128) nmpc_init_tick_kernel(TickArgs a)
{
    const int b = blockIdx.x, tid = threadIdx.x, N = a.N;
    const double* br = a.base_ref + (size_t)b * 12;
    const double* bre = a.base_ref_e + (size_t)b * 12;
    const double* jr = a.joint_ref + (a.joint_per_problem ? (size_t)b * 12 : 0);
    const double sh = a.step_height[a.height_per_problem ? b : 0];
    const signed char* cnt = a.cnt + (size_t)b * NF * (N + 1);
    const signed char* peak = a.peak ? a.peak + (size_t)b * NF * (N + 1) : nullptr;
    const double* loc = a.loc ? a.loc + (size_t)b * NF * (N + 1) * 3 : nullptr;
    const bool use_loc = a.restrict_cnt && loc;

    // ---- running references ----
    double* yr = a.yref + (size_t)b * N * NY;
    for (int e = tid; e < N * NY; e += 128) {
        const int k = e / NY, c = e % NY;
        double v = 0.0;
        if (c < Y_JOINT) v = br[c];
        else if (c < Y_JOINT + 12) v = jr[c - Y_JOINT];
        else if (c >= Y_SWING && c < Y_SWING + NF) v = sh;
        else if (c >= Y_POS && loc) { const int f = (c - Y_POS) / 3, i = (c - Y_POS) % 3; v = loc[((size_t)f * (N + 1) + k + 1) * 3 + i]; }
        yr[e] = v;
    }
    // ---- terminal references ----
    double* ye = a.yref_e + (size_t)b * NYE;
    for (int c = tid; c < NYE; c += 128) {
        double v = 0.0;
        if (c < YE_JOINT) v = bre[c];
        else if (c < YE_JOINT + 12) v = jr[c - YE_JOINT];
        else if (c >= YE_SWING && c < YE_SWING + NF) v = sh;
        else if (c >= YE_POS && loc) { const int f = (c - YE_POS) / 3, i = (c - YE_POS) % 3; v = loc[((size_t)f * (N + 1) + N) * 3 + i]; }
        ye[c] = v;
    }
    // ---- per-node, per-foot parameters ----
    double* pb = a.p + (size_t)b * (N + 1) * NPAR;
    for (int e = tid; e < (N + 1) * NF; e += 128) {
        const int k = e / NF, f = e % NF;
        double* o = pb + (size_t)k * NPAR + f * 11;
        const int act = cnt[f * (N + 1) + k];
        o[P_ACTIVE] = (double)act;
        o[P_POINT] = use_loc ? loc[((size_t)f * (N + 1) + k) * 3] : 0.0;
        o[P_POINT + 1] = use_loc ? loc[((size_t)f * (N + 1) + k) * 3 + 1] : 0.0;
        o[P_POINT + 2] = use_loc ? loc[((size_t)f * (N + 1) + k) * 3 + 2] : a.height_offset;
        o[P_NORMAL] = 0.0; o[P_NORMAL + 1] = 0.0; o[P_NORMAL + 2] = 1.0;
        o[P_GAIN] = a.p_gain[f];
        int made = 0;                                                   // 1 on the node where the foot touches down
        if (a.restrict_cnt && k > 0) { made = act - (int)cnt[f * (N + 1) + k - 1]; if (made < 0) made = 0; }
        o[P_RESTRICT] = (double)made;
        o[P_RADIUS] = a.restrict_cnt ? a.radius : 1.0e10;
        o[P_PEAK] = peak ? (double)peak[f * (N + 1) + k] : 0.0;
    }
}

// cold first guess: x_k = x0, a_k = 0, vertical forces = weight / number of stance feet of node k
__global__ void __launch_bounds__(128) nmpc_cold_guess_kernel(int N, const double* x0, const double* p, double* x, double* u)
{
    const int b = blockIdx.x, tid = threadIdx.x;
    const double* x0b = x0 + (size_t)b * NX;
    const double* pb = p + (size_t)b * (N + 1) * NPAR;
    double* xb = x + (size_t)b * (N + 1) * NX;
    double* ub = u + (size_t)b * N * NU;
    for (int e = tid; e < (N + 1) * NX; e += 128) xb[e] = x0b[e % NX];
    for (int e = tid; e < N * NU; e += 128) {
        const int k = e / NU, c = e % NU;
        double v = 0.0;
        if (c >= IFRC && (c - IFRC) % 3 == 2) {
            const double* pk = pb + (size_t)k * NPAR;
            const double n_act = pk[P_ACTIVE] + pk[11 + P_ACTIVE] + pk[22 + P_ACTIVE] + pk[33 + P_ACTIVE];
            v = pk[((c - IFRC) / 3) * 11 + P_ACTIVE] * (NMPC_MASS * 9.81) / (n_act > 1.0 ? n_act : 1.0);
        }
        ub[e] = v;
    }
}
