/* nmpc_oracle.c -- CPU restatement (plain C, fp64) of the hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * PARITY UNPINNED: the arithmetic of the reference's path lives in dependencies that are absent
 * from /root/reference (contact_tamp @ 06601f49, acados HEAD, HPIPM, BLASFEO 0.1.4.1, CasADi
 * 3.6.7, Pinocchio 3.3.1) and the reference holds no golden vectors for it.  This
 * file restates the published algorithms -- Gauss-Newton SQP / SQP_RTI with full steps (acados
 * ocp_nlp_sqp[_rti]), Mehrotra predictor-corrector interior point on the OCP-QP with a
 * square-root backward Riccati factorisation (HPIPM d_ocp_qp_ipm, mode `speed`) -- for the OCP
 * documented in spec/FORMULATION.md, and follows the in-tree call sites:
 *   mpc_controller/utils/solver.py:108-141 (weights), :153-177 (references), :179-192 (x0),
 *   :194-275 (per-node contact parameters), :290-342 (shift warm start), :397-429 (solution),
 *   mpc_controller/utils/dynamics.py:121-134 (cost expressions), :48-50 (h0).
 * It is checked algorithm-independently in tests/ (finite differences, dense KKT solves, KKT
 * residuals).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference arm may
 * load it.  Everything is dense and written for obviousness, not speed.
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/nmpc_b200.h"

/* Two builds of this file: the default one (nu = 30) and, with -DNMPC_ORACLE_TOPT, the `enable_time_opt` variant
 * (solver.py:43-44, 124-125, 166-167, 189-190, 325-336, 424-427): the node duration dt_k is a 31st input with the
 * cost weight `time_opt` on (dt - dt_nodes), the bounds [dt_min, dt_max] (config_abstract.py:49-52, 80-88; one more
 * two-sided constraint row) and it multiplies the dynamics.  Same algorithms, dense, nothing specialised. */
#define NX NMPC_NX
#define NP NMPC_NP
#define NY NMPC_NY
#define NYE NMPC_NYE
#ifdef NMPC_ORACLE_TOPT
#define NU (NMPC_NU + 1)
#define NZ (NMPC_NZ + 1)
#define NG (NMPC_NG + 1)
#else
#define NU NMPC_NU
#define NZ NMPC_NZ
#define NG NMPC_NG
#endif
#define NQ 18
#define NF 4
#define NMAX NMPC_MAX_NODES

#define NMPC_HD static
#define NMPC_REC(i) rec[(i)]
#define NMPC_WITH_RNEA 1
#include "../iterative_learning_nmpc_b200/csrc/nmpc_gen_model.h"

/* z = [u; x] index helpers */
#define IA 0          /* a   in u */
#define IFRC 18       /* f   in u */
#define IDT 30        /* dt  in u (time-opt build only) */
#define IQ NU         /* q   in z */
#define IV (NU + 18)  /* v   in z */
#define IH (NU + 36)  /* h   in z */
/* constraint row blocks */
#define GR_FRIC 0     /* 4 feet x 5 */
#define GR_CNT 20     /* 4 feet x 3 */
#define GR_PATCH 32   /* 4 feet x 1 */
#define GR_CONS 36    /* 6 */
#define GR_DT 42      /* dt_min <= dt <= dt_max (time-opt build only) */
/* parameter offsets inside one foot block (solver.py:212-252) */
#define P_ACTIVE 0
#define P_POINT 1
#define P_NORMAL 4
#define P_GAIN 7
#define P_RESTRICT 8
#define P_RADIUS 9
#define P_PEAK 10

/* IPM constants: HPIPM `speed` mode defaults as listed in SURVEY.md 8a (from upstream memory) */
#define IPM_MU0 1e1
#define IPM_THR0 1e-1
#define IPM_REG_PRIM 1e-15
#define IPM_LAM_MIN 1e-16
#define IPM_T_MIN 1e-16
#define IPM_ALPHA_MIN 1e-12
#define IPM_TAU_MIN 1e-16
static const double GRAV[3] = {0.0, 0.0, -9.81};
/* running inf-norm that keeps a NaN once it has seen one (plain > comparisons drop NaNs) */
#define NORM_ACC(r, a) do { const double a_ = fabs(a); if (a_ > (r) || a_ != a_) (r) = a_; } while (0)

typedef struct {
    double A[NX][NX], B[NX][NU], b[NX];
    double H[NZ][NZ], g[NZ];
    double C[NG][NZ], c[NG];   /* constraint values at the linearisation point */
    double cl[NG], cu[NG];     /* value minus the lower / upper bound (= c: the bounds are 0, except on the dt row) */
    double ml[NG], mu[NG];     /* row masks: lower / upper bound present */
    double lgq[NG], ugq[NG];   /* QP-level bounds on C dz: lg - c, ug - c */
    double cost;
} stage_lin;

typedef struct {
    double L[NZ][NZ];    /* Cholesky factor of M_k */
    double l[NZ];        /* L^-1 m */
    double p[NX];        /* linear term of the value function */
} stage_fac;

/* per-thread workspace */
typedef struct {
    stage_lin lin[NMAX + 1];
    stage_fac fac[NMAX + 1];
    double Ht[NMAX + 1][NZ][NZ];      /* H + C^T Gamma C */
    double qt[NMAX + 1][NZ];
    double rg[NMAX + 1][NZ], rb[NMAX][NX], rd[NMAX + 1][2 * NG], rm[NMAX + 1][2 * NG];
    double dz[NMAX + 1][NZ], dpi[NMAX][NX], dlam[NMAX + 1][2 * NG], dt[NMAX + 1][2 * NG];
    double dlam_a[NMAX + 1][2 * NG], dt_a[NMAX + 1][2 * NG];
} workspace;

/* ------------------------------------------------------------------------------------------ */
/* tangent basis of the contact plane from its normal                                         */
/* ------------------------------------------------------------------------------------------ */
static void tangent_basis(const double* n_in, double* n, double* t1, double* t2)
{
    double nn = sqrt(n_in[0] * n_in[0] + n_in[1] * n_in[1] + n_in[2] * n_in[2]);
    if (nn < 1e-12) { n[0] = 0; n[1] = 0; n[2] = 1; } else { for (int i = 0; i < 3; i++) n[i] = n_in[i] / nn; }
    /* t1 = normalised projection of e_x on the plane (e_y if the normal is along x) */
    double e[3] = {1, 0, 0};
    if (fabs(n[0]) > 0.9) { e[0] = 0; e[1] = 1; }
    double d = e[0] * n[0] + e[1] * n[1] + e[2] * n[2];
    for (int i = 0; i < 3; i++) t1[i] = e[i] - d * n[i];
    double tn = sqrt(t1[0] * t1[0] + t1[1] * t1[1] + t1[2] * t1[2]);
    for (int i = 0; i < 3; i++) t1[i] /= tn;
    t2[0] = n[1] * t1[2] - n[2] * t1[1];
    t2[1] = n[2] * t1[0] - n[0] * t1[2];
    t2[2] = n[0] * t1[1] - n[1] * t1[0];
}

/* ------------------------------------------------------------------------------------------ */
/* linearise one stage: dynamics defect, Gauss-Newton cost, constraints (spec/FORMULATION.md)  */
/* ------------------------------------------------------------------------------------------ */
static void linearize_stage(const nmpc_opts* o, int k, int N, const double* x, const double* u,
                            const double* xnext, const double* p, const double* yref,
                            const double* W, stage_lin* s)
{
    const int term = (k == N);
    const double* q = x;
    const double* v = x + NQ;
    const double* h = x + 2 * NQ;
    static const double zero_u[NU] = {0};
    if (term) u = zero_u;
#ifdef NMPC_ORACLE_TOPT
    const double dt = term ? o->dt_nodes : u[IDT];
#else
    const double dt = o->dt_nodes;
#endif
    const double* a = u;
    const double* f = u + NQ;
    double rec[NMPC_REC_SIZE];
    nmpc_model_kin(q, v, rec);
    const double* FP = rec + NMPC_REC_FOOT_P;
    const double* FJ = rec + NMPC_REC_FOOT_J;
    const double* FV = rec + NMPC_REC_FOOT_V;
    const double* FVQ = rec + NMPC_REC_FOOT_VQ;
    const double* COM = rec + NMPC_REC_COM;
    const double* CJ = rec + NMPC_REC_COM_J;
    const double* HM = rec + NMPC_REC_H;
    const double* HQ = rec + NMPC_REC_H_Q;
    const double* HV = rec + NMPC_REC_H_V;

    memset(s, 0, sizeof(*s));

    /* ---- dynamics x+ = phi(x,u,p): q+ = q + dt v + dt^2/2 a ; v+ = v + dt a ;
     *      h+ = h + dt [ m g + sum_i act_i f_i ; sum_i act_i (p_i(q) - c(q)) x f_i ] ---- */
    if (!term) {
        double phi[NX];
        for (int i = 0; i < NQ; i++) {
            phi[i] = q[i] + dt * v[i] + 0.5 * dt * dt * a[i];
            phi[NQ + i] = v[i] + dt * a[i];
            s->A[i][i] = 1.0;
            s->A[i][NQ + i] = dt;
            s->A[NQ + i][NQ + i] = 1.0;
            s->B[i][IA + i] = 0.5 * dt * dt;
            s->B[NQ + i][IA + i] = dt;
        }
        double hd[6] = {NMPC_MASS * GRAV[0], NMPC_MASS * GRAV[1], NMPC_MASS * GRAV[2], 0, 0, 0};
        for (int i = 0; i < 6; i++) s->A[2 * NQ + i][2 * NQ + i] = 1.0;
        for (int ft = 0; ft < NF; ft++) {
            const double act = p[ft * NMPC_NP_FOOT + P_ACTIVE];
            const double* fi = f + 3 * ft;
            double r[3];
            for (int i = 0; i < 3; i++) r[i] = FP[3 * ft + i] - COM[i];
            hd[0] += act * fi[0]; hd[1] += act * fi[1]; hd[2] += act * fi[2];
            hd[3] += act * (r[1] * fi[2] - r[2] * fi[1]);
            hd[4] += act * (r[2] * fi[0] - r[0] * fi[2]);
            hd[5] += act * (r[0] * fi[1] - r[1] * fi[0]);
            for (int i = 0; i < 3; i++) s->B[2 * NQ + i][IFRC + 3 * ft + i] = dt * act;
            /* d(r x f)/df = [r]_x */
            double rx[3][3] = {{0, -r[2], r[1]}, {r[2], 0, -r[0]}, {-r[1], r[0], 0}};
            for (int i = 0; i < 3; i++)
                for (int j = 0; j < 3; j++) s->B[2 * NQ + 3 + i][IFRC + 3 * ft + j] = dt * act * rx[i][j];
            /* d(r x f)/dq = -[f]_x (Jp - Jc) */
            double fx[3][3] = {{0, -fi[2], fi[1]}, {fi[2], 0, -fi[0]}, {-fi[1], fi[0], 0}};
            for (int i = 0; i < 3; i++)
                for (int j = 0; j < NQ; j++) {
                    double acc = 0;
                    for (int l = 0; l < 3; l++) acc -= fx[i][l] * (FJ[(3 * ft + l) * NQ + j] - CJ[l * NQ + j]);
                    s->A[2 * NQ + 3 + i][j] += dt * act * acc;
                }
        }
        for (int i = 0; i < 6; i++) phi[2 * NQ + i] = h[i] + dt * hd[i];
#ifdef NMPC_ORACLE_TOPT
        for (int i = 0; i < NQ; i++) { s->B[i][IDT] = v[i] + dt * a[i]; s->B[NQ + i][IDT] = a[i]; }
        for (int i = 0; i < 6; i++) s->B[2 * NQ + i][IDT] = hd[i];
#endif
        for (int i = 0; i < NX; i++) s->b[i] = phi[i] - xnext[i];
    }

    /* ---- Gauss-Newton least-squares cost 1/2 (y - yref)^T W (y - yref) ---- */
    {
        const int ny = term ? NYE : NY;
        double J[NY][NZ];
        double y[NY];
        memset(J, 0, sizeof(J));
        int r = 0;
        /* base_cost = [q[:3], q[3:6], v[:6]]  (dynamics.py:121-124) */
        for (int i = 0; i < 6; i++, r++) { y[r] = q[i]; J[r][IQ + i] = 1; }
        for (int i = 0; i < 6; i++, r++) { y[r] = v[i]; J[r][IV + i] = 1; }
        /* joint_cost = [q[6:], v[6:]]  (dynamics.py:126-127) */
        for (int i = 6; i < NQ; i++, r++) { y[r] = q[i]; J[r][IQ + i] = 1; }
        for (int i = 6; i < NQ; i++, r++) { y[r] = v[i]; J[r][IV + i] = 1; }
        /* acc_cost = a[6:]  (dynamics.py:129-130), running only */
        if (!term) for (int i = 6; i < NQ; i++, r++) { y[r] = a[i]; J[r][IA + i] = 1; }
        /* sw_cost = [peak_i * foot_i.z]  (dynamics.py:132-134) */
        for (int ft = 0; ft < NF; ft++, r++) {
            const double pk = p[ft * NMPC_NP_FOOT + P_PEAK];
            y[r] = pk * FP[3 * ft + 2];
            for (int j = 0; j < NQ; j++) J[r][IQ + j] = pk * FJ[(3 * ft + 2) * NQ + j];
        }
        /* f_reg_i = f_i, running only */
        if (!term) for (int i = 0; i < 12; i++, r++) { y[r] = f[i]; J[r][IFRC + i] = 1; }
        /* pos_cost_i = foot_i position (solver.py:126-131,273-275) */
        for (int ft = 0; ft < NF; ft++)
            for (int i = 0; i < 3; i++, r++) {
                y[r] = FP[3 * ft + i];
                for (int j = 0; j < NQ; j++) J[r][IQ + j] = FJ[(3 * ft + i) * NQ + j];
            }
        double cost = 0;
        for (int i = 0; i < ny; i++) {
            double e = y[i] - yref[i];
            cost += 0.5 * W[i] * e * e;
            for (int c = 0; c < NZ; c++) {
                if (J[i][c] == 0.0) continue;
                s->g[c] += J[i][c] * W[i] * e;
                for (int d = 0; d < NZ; d++) s->H[c][d] += J[i][c] * W[i] * J[i][d];
            }
        }
#ifdef NMPC_ORACLE_TOPT
        if (!term) {        /* W["dt"] (dt - dt_nodes)^2 / 2  (solver.py:124-125, 166-167) */
            const double e = dt - o->dt_nodes;
            cost += 0.5 * o->time_opt_weight * e * e;
            s->g[IDT] += o->time_opt_weight * e;
            s->H[IDT][IDT] += o->time_opt_weight;
        }
#endif
        s->cost = cost;
        const double reg = term ? o->reg_eps_e : o->reg_eps;
        for (int c = 0; c < NZ; c++) s->H[c][c] += reg;
    }

    /* ---- constraints: every row is g_j(x,u,p) with bound 0 on the masked side(s) ---- */
    for (int ft = 0; ft < NF; ft++) {
        const double* pf = p + ft * NMPC_NP_FOOT;
        const double act = pf[P_ACTIVE];
        const int on = act > 0.5;
        double n[3], t1[3], t2[3];
        tangent_basis(pf + P_NORMAL, n, t1, t2);
        const double* fi = f + 3 * ft;
        const double mu_c = o->mu_contact;
        /* friction pyramid + unilateral force, rows use f only; none at the terminal node */
        if (!term) {
            const double fn = n[0] * fi[0] + n[1] * fi[1] + n[2] * fi[2];
            const double f1 = t1[0] * fi[0] + t1[1] * fi[1] + t1[2] * fi[2];
            const double f2 = t2[0] * fi[0] + t2[1] * fi[1] + t2[2] * fi[2];
            const int r0 = GR_FRIC + 5 * ft;
            s->c[r0 + 0] = mu_c * fn - f1;
            s->c[r0 + 1] = mu_c * fn + f1;
            s->c[r0 + 2] = mu_c * fn - f2;
            s->c[r0 + 3] = mu_c * fn + f2;
            s->c[r0 + 4] = fn;
            for (int i = 0; i < 3; i++) {
                s->C[r0 + 0][IFRC + 3 * ft + i] = mu_c * n[i] - t1[i];
                s->C[r0 + 1][IFRC + 3 * ft + i] = mu_c * n[i] + t1[i];
                s->C[r0 + 2][IFRC + 3 * ft + i] = mu_c * n[i] - t2[i];
                s->C[r0 + 3][IFRC + 3 * ft + i] = mu_c * n[i] + t2[i];
                s->C[r0 + 4][IFRC + 3 * ft + i] = n[i];
            }
            for (int j = 0; j < 5; j++) s->ml[r0 + j] = on ? 1.0 : 0.0;
        }
        /* stance kinematics: foot on the plane, no tangential velocity; state-only, nodes >= 1 */
        if (k > 0) {
            const int r0 = GR_CNT + 3 * ft;
            double d[3];
            for (int i = 0; i < 3; i++) d[i] = FP[3 * ft + i] - pf[P_POINT + i];
            s->c[r0 + 0] = n[0] * d[0] + n[1] * d[1] + n[2] * d[2];
            s->c[r0 + 1] = t1[0] * FV[3 * ft] + t1[1] * FV[3 * ft + 1] + t1[2] * FV[3 * ft + 2];
            s->c[r0 + 2] = t2[0] * FV[3 * ft] + t2[1] * FV[3 * ft + 1] + t2[2] * FV[3 * ft + 2];
            for (int j = 0; j < NQ; j++) {
                double jn = 0, v1q = 0, v2q = 0, v1v = 0, v2v = 0;
                for (int i = 0; i < 3; i++) {
                    jn += n[i] * FJ[(3 * ft + i) * NQ + j];
                    v1q += t1[i] * FVQ[(3 * ft + i) * NQ + j];
                    v2q += t2[i] * FVQ[(3 * ft + i) * NQ + j];
                    v1v += t1[i] * FJ[(3 * ft + i) * NQ + j];
                    v2v += t2[i] * FJ[(3 * ft + i) * NQ + j];
                }
                s->C[r0 + 0][IQ + j] = jn;
                s->C[r0 + 1][IQ + j] = v1q; s->C[r0 + 1][IV + j] = v1v;
                s->C[r0 + 2][IQ + j] = v2q; s->C[r0 + 2][IV + j] = v2v;
            }
            for (int j = 0; j < 3; j++) { s->ml[r0 + j] = on; s->mu[r0 + j] = on; }
            /* contact patch: |tangential offset|^2 - radius^2 <= 0 where restricted */
            const int rp = GR_PATCH + ft;
            const double rad = pf[P_RADIUS];
            const int restricted = on && pf[P_RESTRICT] > 0.5 && rad < 1e9;
            const double d1 = t1[0] * d[0] + t1[1] * d[1] + t1[2] * d[2];
            const double d2 = t2[0] * d[0] + t2[1] * d[1] + t2[2] * d[2];
            if (restricted) {
                s->c[rp] = d1 * d1 + d2 * d2 - rad * rad;
                for (int j = 0; j < NQ; j++) {
                    double j1 = 0, j2 = 0;
                    for (int i = 0; i < 3; i++) {
                        j1 += t1[i] * FJ[(3 * ft + i) * NQ + j];
                        j2 += t2[i] * FJ[(3 * ft + i) * NQ + j];
                    }
                    s->C[rp][IQ + j] = 2 * d1 * j1 + 2 * d2 * j2;
                }
                s->mu[rp] = 1.0;
            }
        }
    }
    /* centroidal consistency h - h_model(q,v) = 0, nodes >= 1 */
    if (k > 0) {
        for (int i = 0; i < 6; i++) {
            const int r = GR_CONS + i;
            s->c[r] = h[i] - HM[i];
            for (int j = 0; j < NQ; j++) {
                s->C[r][IQ + j] = -HQ[i * NQ + j];
                s->C[r][IV + j] = -HV[i * NQ + j];
            }
            s->C[r][IH + i] = 1.0;
            s->ml[r] = 1.0; s->mu[r] = 1.0;
        }
    }
    for (int j = 0; j < NG; j++) { s->cl[j] = s->c[j]; s->cu[j] = s->c[j]; }
#ifdef NMPC_ORACLE_TOPT
    if (!term) {            /* dt_min <= dt <= dt_max */
        s->C[GR_DT][IDT] = 1.0;
        s->c[GR_DT] = dt; s->cl[GR_DT] = dt - o->dt_min; s->cu[GR_DT] = dt - o->dt_max;
        s->ml[GR_DT] = 1.0; s->mu[GR_DT] = 1.0;
    }
#endif
    for (int j = 0; j < NG; j++) { s->lgq[j] = -s->cl[j]; s->ugq[j] = -s->cu[j]; }
}

/* ------------------------------------------------------------------------------------------ */
/* dense helpers                                                                              */
/* ------------------------------------------------------------------------------------------ */
/* in-place lower Cholesky of the leading n x n block of M (ld NZ); non-positive pivots zero the
 * column (BLASFEO potrf convention).  Returns the number of bad pivots. */
static int chol_lower(double M[NZ][NZ], int off, int n)
{
    int bad = 0;
    for (int j = 0; j < n; j++) {
        double d = M[off + j][off + j];
        for (int l = 0; l < j; l++) d -= M[off + j][off + l] * M[off + j][off + l];
        double inv;
        if (d > 0.0) { d = sqrt(d); inv = 1.0 / d; } else { d = 0.0; inv = 0.0; bad++; }
        M[off + j][off + j] = d;
        for (int i = j + 1; i < n; i++) {
            double s = M[off + i][off + j];
            for (int l = 0; l < j; l++) s -= M[off + i][off + l] * M[off + j][off + l];
            M[off + i][off + j] = s * inv;
        }
        for (int i = 0; i < j; i++) M[off + i][off + j] = 0.0;
    }
    return bad;
}

/* ------------------------------------------------------------------------------------------ */
/* OCP-QP interior point (Mehrotra predictor-corrector, square-root Riccati)                  */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
    int iter, status;           /* HPIPM numbering 0 ok, 1 max-iter, 2 min-step, 3 NaN */
    double res[4];              /* stat, eq, ineq, comp inf-norms at exit */
    double mu;
    int ncond;                  /* iterations in which the conditional predictor-corrector fell back to centring */
} qp_info;

static int stage_off(int k) { return k == 0 ? 1 : 0; }   /* unused placeholder for clarity */

/* factorise (do_fact) or only re-solve the vector recursion; fills w->dz, w->dpi */
static int riccati(workspace* w, int N, const double* dx0, int do_fact)
{
    int bad = 0;
    (void)stage_off;
    /* terminal stage: P_N = Ht_N[xx], p_N = qt_N[x] */
    stage_fac* fN = &w->fac[N];
    if (do_fact) {
        memset(fN->L, 0, sizeof(fN->L));
        for (int i = 0; i < NX; i++)
            for (int j = 0; j <= i; j++) fN->L[NU + i][NU + j] = w->Ht[N][NU + i][NU + j];
        bad += chol_lower(fN->L, NU, NX);
    }
    for (int i = 0; i < NX; i++) fN->p[i] = w->qt[N][NU + i];

    for (int k = N - 1; k >= 0; k--) {
        stage_lin* s = &w->lin[k];
        stage_fac* fk = &w->fac[k];
        stage_fac* fn = &w->fac[k + 1];
        /* G = [B A] (NX x NZ) */
        if (do_fact) {
            /* Wm = Lxx_{k+1}^T G */
            static __thread double Wm[NX][NZ];
            for (int i = 0; i < NX; i++)
                for (int c = 0; c < NZ; c++) {
                    double acc = 0;
                    for (int l = i; l < NX; l++) {
                        double gl = c < NU ? s->B[l][c] : s->A[l][c - NU];
                        acc += fn->L[NU + l][NU + i] * gl;
                    }
                    Wm[i][c] = acc;
                }
            for (int i = 0; i < NZ; i++)
                for (int j = 0; j <= i; j++) {
                    double acc = w->Ht[k][i][j];
                    for (int l = 0; l < NX; l++) acc += Wm[l][i] * Wm[l][j];
                    fk->L[i][j] = acc;
                }
            for (int i = 0; i < NZ; i++) fk->L[i][i] += IPM_REG_PRIM;
            bad += chol_lower(fk->L, 0, NZ);
        }
        /* m = qt_k + G^T (P_{k+1} rb_k + p_{k+1}),  P = Lxx Lxx^T */
        double tmp[NX], tmp2[NX];
        for (int i = 0; i < NX; i++) {
            double acc = 0;
            for (int l = i; l < NX; l++) acc += fn->L[NU + l][NU + i] * w->rb[k][l];
            tmp[i] = acc;
        }
        for (int i = 0; i < NX; i++) {
            double acc = fn->p[i];
            for (int l = 0; l <= i; l++) acc += fn->L[NU + i][NU + l] * tmp[l];
            tmp2[i] = acc;
        }
        double m[NZ];
        for (int c = 0; c < NZ; c++) {
            double acc = w->qt[k][c];
            for (int l = 0; l < NX; l++) acc += (c < NU ? s->B[l][c] : s->A[l][c - NU]) * tmp2[l];
            m[c] = acc;
        }
        /* l = L^-1 m */
        for (int i = 0; i < NZ; i++) {
            double acc = m[i];
            for (int l = 0; l < i; l++) acc -= fk->L[i][l] * fk->l[l];
            double d = fk->L[i][i];
            fk->l[i] = d != 0.0 ? acc / d : 0.0;
        }
        /* p_k = Lxx lx */
        for (int i = 0; i < NX; i++) {
            double acc = 0;
            for (int l = 0; l <= i; l++) acc += fk->L[NU + i][NU + l] * fk->l[NU + l];
            fk->p[i] = acc;
        }
    }
    /* forward sweep */
    double dx[NX];
    memcpy(dx, dx0, sizeof(dx));
    for (int k = 0; k < N; k++) {
        stage_lin* s = &w->lin[k];
        stage_fac* fk = &w->fac[k];
        stage_fac* fn = &w->fac[k + 1];
        /* du = -Luu^-T (lu + Lxu^T dx) */
        double rhs[NU], du[NU];
        for (int i = 0; i < NU; i++) {
            double acc = fk->l[i];
            for (int l = 0; l < NX; l++) acc += fk->L[NU + l][i] * dx[l];
            rhs[i] = acc;
        }
        for (int i = NU - 1; i >= 0; i--) {
            double acc = -rhs[i];
            for (int l = i + 1; l < NU; l++) acc -= fk->L[l][i] * du[l];
            double d = fk->L[i][i];
            du[i] = d != 0.0 ? acc / d : 0.0;
        }
        for (int i = 0; i < NU; i++) w->dz[k][i] = du[i];
        for (int i = 0; i < NX; i++) w->dz[k][NU + i] = dx[i];
        double dxn[NX];
        for (int i = 0; i < NX; i++) {
            double acc = w->rb[k][i];
            for (int l = 0; l < NX; l++) acc += s->A[i][l] * dx[l];
            for (int l = 0; l < NU; l++) acc += s->B[i][l] * du[l];
            dxn[i] = acc;
        }
        /* dpi_{k+1} = P_{k+1} dx+ + p_{k+1} */
        double tmp[NX];
        for (int i = 0; i < NX; i++) {
            double acc = 0;
            for (int l = i; l < NX; l++) acc += fn->L[NU + l][NU + i] * dxn[l];
            tmp[i] = acc;
        }
        for (int i = 0; i < NX; i++) {
            double acc = fn->p[i];
            for (int l = 0; l <= i; l++) acc += fn->L[NU + i][NU + l] * tmp[l];
            w->dpi[k][i] = acc;
        }
        memcpy(dx, dxn, sizeof(dx));
    }
    for (int i = 0; i < NU; i++) w->dz[N][i] = 0.0;
    for (int i = 0; i < NX; i++) w->dz[N][NU + i] = dx[i];
    return bad;
}

/* QP of the SQP step at the current linearisation held in w->lin.
 * in/out: z [N+1][NZ], pi [N][NX], lam/t [N+1][2NG] (lower | upper).  dx0 = x0 - xbar_0 is imposed on stage 0.
 * warm = HPIPM's `warm_start` level (d_ocp_qp_ipm.c, INIT_VAR):
 *   0 cold: z = 0, pi = 0, t = max(slack(z), thr0), lam = mu0 / t
 *   1 primal: z kept, otherwise as 0 (pi is zeroed as well)
 *   2 primal and dual ("hot"): z, pi, lam, t kept; lam and t of the masked rows clipped from below at thr0
 *     (rows whose stored lam or t is not positive -- never initialised, e.g. after reset_duals -- are initialised as in 1) */
static void qp_ipm(workspace* w, const nmpc_opts* o, int N, const double* dx0, int warm,
                   double (*z)[NZ], double (*pi)[NX], double (*lam)[2 * NG], double (*t)[2 * NG],
                   qp_info* info)
{
    const double tol = o->qp_tol;
    if (warm == 0) memset(z, 0, sizeof(double) * (N + 1) * NZ);
    if (warm < 2) memset(pi, 0, sizeof(double) * N * NX);
    for (int i = 0; i < NX; i++) z[0][NU + i] = dx0[i];
    for (int i = 0; i < NU; i++) z[N][i] = 0.0;
    int nc = 0;
    for (int k = 0; k <= N; k++) {
        stage_lin* s = &w->lin[k];
        for (int j = 0; j < NG; j++) {
            double cz = 0;
            for (int c = 0; c < NZ; c++) cz += s->C[j][c] * z[k][c];
            if (s->ml[j] > 0.5) {
                nc++;
                if (warm < 2 || !(t[k][j] > 0.0) || !(lam[k][j] > 0.0)) { double tl = cz - s->lgq[j]; t[k][j] = tl > IPM_THR0 ? tl : IPM_THR0; lam[k][j] = IPM_MU0 / t[k][j]; }
                else { if (t[k][j] < IPM_THR0) t[k][j] = IPM_THR0; if (lam[k][j] < IPM_THR0) lam[k][j] = IPM_THR0; }
            } else { t[k][j] = 1.0; lam[k][j] = 0.0; }
            if (s->mu[j] > 0.5) {
                nc++;
                if (warm < 2 || !(t[k][NG + j] > 0.0) || !(lam[k][NG + j] > 0.0)) { double tu = s->ugq[j] - cz; t[k][NG + j] = tu > IPM_THR0 ? tu : IPM_THR0; lam[k][NG + j] = IPM_MU0 / t[k][NG + j]; }
                else { if (t[k][NG + j] < IPM_THR0) t[k][NG + j] = IPM_THR0; if (lam[k][NG + j] < IPM_THR0) lam[k][NG + j] = IPM_THR0; }
            } else { t[k][NG + j] = 1.0; lam[k][NG + j] = 0.0; }
        }
    }
    const double inv_nc = nc > 0 ? 1.0 / nc : 0.0;
    info->status = 1;
    info->ncond = 0;
    int it = 0;
    for (;; it++) {
        /* ---- residuals ---- */
        double res_g = 0, res_b = 0, res_d = 0, res_m = 0, mu = 0;
        for (int k = 0; k <= N; k++) {
            stage_lin* s = &w->lin[k];
            double cz[NG];
            for (int j = 0; j < NG; j++) {
                double acc = 0;
                for (int c = 0; c < NZ; c++) acc += s->C[j][c] * z[k][c];
                cz[j] = acc;
            }
            for (int j = 0; j < NG; j++) {
                w->rd[k][j] = s->ml[j] * (cz[j] - t[k][j] - s->lgq[j]);
                w->rd[k][NG + j] = s->mu[j] * (-cz[j] - t[k][NG + j] + s->ugq[j]);
                w->rm[k][j] = s->ml[j] * lam[k][j] * t[k][j];
                w->rm[k][NG + j] = s->mu[j] * lam[k][NG + j] * t[k][NG + j];
                mu += w->rm[k][j] + w->rm[k][NG + j];
                for (int e = 0; e < 2; e++) {
                    NORM_ACC(res_d, w->rd[k][e * NG + j]);
                    NORM_ACC(res_m, w->rm[k][e * NG + j]);
                }
            }
            for (int c = 0; c < NZ; c++) {
                double acc = s->g[c];
                for (int d = 0; d < NZ; d++) acc += s->H[c][d] * z[k][d];
                if (k < N) for (int l = 0; l < NX; l++) acc += (c < NU ? s->B[l][c] : s->A[l][c - NU]) * pi[k][l];
                if (k > 0 && c >= NU) acc -= pi[k - 1][c - NU];
                for (int j = 0; j < NG; j++) acc -= s->C[j][c] * (lam[k][j] - lam[k][NG + j]);
                if ((k == 0 && c >= NU) || (k == N && c < NU)) acc = 0.0;   /* fixed x0 / no u_N */
                w->rg[k][c] = acc;
                NORM_ACC(res_g, acc);
            }
            if (k < N) for (int i = 0; i < NX; i++) {
                double acc = s->b[i] - z[k + 1][NU + i];
                for (int l = 0; l < NX; l++) acc += s->A[i][l] * z[k][NU + l];
                for (int l = 0; l < NU; l++) acc += s->B[i][l] * z[k][l];
                w->rb[k][i] = acc;
                NORM_ACC(res_b, acc);
            }
        }
        mu *= inv_nc;
        info->res[0] = res_g; info->res[1] = res_b; info->res[2] = res_d; info->res[3] = res_m; info->mu = mu;
        if (!(res_g == res_g) || !(res_b == res_b) || !(res_d == res_d) || !(res_m == res_m)) { info->status = 3; break; }
        if (res_g < tol && res_b < tol && res_d < tol && res_m < tol) { info->status = 0; break; }
        if (it >= o->max_qp_iter) { info->status = 1; break; }

        /* ---- condensed Hessian / gradient (affine direction, sigma = 0) ---- */
        for (int k = 0; k <= N; k++) {
            stage_lin* s = &w->lin[k];
            double Gam[NG], gam[NG];
            for (int j = 0; j < NG; j++) {
                Gam[j] = s->ml[j] * lam[k][j] / t[k][j] + s->mu[j] * lam[k][NG + j] / t[k][NG + j];
                gam[j] = s->ml[j] * (w->rm[k][j] + lam[k][j] * w->rd[k][j]) / t[k][j]
                       - s->mu[j] * (w->rm[k][NG + j] + lam[k][NG + j] * w->rd[k][NG + j]) / t[k][NG + j];
            }
            for (int c = 0; c < NZ; c++) {
                double acc = w->rg[k][c];
                for (int j = 0; j < NG; j++) acc += s->C[j][c] * gam[j];
                w->qt[k][c] = acc;
                for (int d = 0; d <= c; d++) {
                    double h = s->H[c][d];
                    for (int j = 0; j < NG; j++) h += s->C[j][c] * Gam[j] * s->C[j][d];
                    w->Ht[k][c][d] = h;
                }
            }
        }
        double zero0[NX] = {0};
        riccati(w, N, zero0, 1);
        /* ---- affine slack / multiplier directions, step length ---- */
        double alpha_aff = 1.0;
        for (int k = 0; k <= N; k++) {
            stage_lin* s = &w->lin[k];
            for (int j = 0; j < NG; j++) {
                double cdz = 0;
                for (int c = 0; c < NZ; c++) cdz += s->C[j][c] * w->dz[k][c];
                for (int e = 0; e < 2; e++) {
                    const int id = e * NG + j;
                    const double msk = e ? s->mu[j] : s->ml[j];
                    if (msk < 0.5) { w->dt_a[k][id] = 0; w->dlam_a[k][id] = 0; continue; }
                    double dtv = (e ? -cdz : cdz) + w->rd[k][id];
                    double dl = -(w->rm[k][id] + lam[k][id] * dtv) / t[k][id];
                    w->dt_a[k][id] = dtv; w->dlam_a[k][id] = dl;
                    if (dtv < 0 && -t[k][id] / dtv < alpha_aff) alpha_aff = -t[k][id] / dtv;
                    if (dl < 0 && -lam[k][id] / dl < alpha_aff) alpha_aff = -lam[k][id] / dl;
                }
            }
        }
        double mu_aff = 0;
        for (int k = 0; k <= N; k++)
            for (int id = 0; id < 2 * NG; id++) {
                const double msk = id < NG ? w->lin[k].ml[id] : w->lin[k].mu[id - NG];
                if (msk > 0.5) mu_aff += (lam[k][id] + alpha_aff * w->dlam_a[k][id]) * (t[k][id] + alpha_aff * w->dt_a[k][id]);
            }
        mu_aff *= inv_nc;
        double sigma = mu > 0 ? mu_aff / mu : 0.0;
        sigma = sigma * sigma * sigma;
        /* ---- corrector: rm <- lam*t + dt_a*dlam_a - sigma*mu, same factorisation (HPIPM
         *      COMPUTE_CENTERING_CORRECTION_QP); pass 1 only if the conditional predictor-corrector
         *      (arg->cond_pred_corr, on in every HPIPM mode) rejects that direction: when the complementarity
         *      it predicts, mu(alpha), exceeds twice the affine prediction mu_aff, the centring direction
         *      rm <- lam*t - sigma*mu (COMPUTE_CENTERING_QP) is solved for instead ---- */
        double sigma_mu = sigma * mu;
        if (sigma_mu < IPM_TAU_MIN) sigma_mu = IPM_TAU_MIN;
        double alpha_p = 1.0, alpha_d = 1.0;
        for (int pass = 0; pass < 2; pass++) {
            for (int k = 0; k <= N; k++) {
                stage_lin* s = &w->lin[k];
                double gam[NG];
                for (int j = 0; j < NG; j++) {
                    for (int e = 0; e < 2; e++) {
                        const int id = e * NG + j;
                        const double msk = e ? s->mu[j] : s->ml[j];
                        if (msk > 0.5)
                            w->rm[k][id] = lam[k][id] * t[k][id] + (pass == 0 ? w->dt_a[k][id] * w->dlam_a[k][id] : 0.0) - sigma_mu;
                    }
                    gam[j] = s->ml[j] * (w->rm[k][j] + lam[k][j] * w->rd[k][j]) / t[k][j]
                           - s->mu[j] * (w->rm[k][NG + j] + lam[k][NG + j] * w->rd[k][NG + j]) / t[k][NG + j];
                }
                for (int c = 0; c < NZ; c++) {
                    double acc = w->rg[k][c];
                    for (int j = 0; j < NG; j++) acc += s->C[j][c] * gam[j];
                    w->qt[k][c] = acc;
                }
            }
            riccati(w, N, zero0, 0);
            alpha_p = 1.0; alpha_d = 1.0;
            for (int k = 0; k <= N; k++) {
                stage_lin* s = &w->lin[k];
                for (int j = 0; j < NG; j++) {
                    double cdz = 0;
                    for (int c = 0; c < NZ; c++) cdz += s->C[j][c] * w->dz[k][c];
                    for (int e = 0; e < 2; e++) {
                        const int id = e * NG + j;
                        const double msk = e ? s->mu[j] : s->ml[j];
                        if (msk < 0.5) { w->dt[k][id] = 0; w->dlam[k][id] = 0; continue; }
                        double dtv = (e ? -cdz : cdz) + w->rd[k][id];
                        double dl = -(w->rm[k][id] + lam[k][id] * dtv) / t[k][id];
                        w->dt[k][id] = dtv; w->dlam[k][id] = dl;
                        if (dtv < 0 && -t[k][id] / dtv < alpha_p) alpha_p = -t[k][id] / dtv;
                        if (dl < 0 && -lam[k][id] / dl < alpha_d) alpha_d = -lam[k][id] / dl;
                    }
                }
            }
            if (pass == 1 || !o->cond_pred_corr) break;
            const double alpha_c = alpha_p < alpha_d ? alpha_p : alpha_d;
            double mu_c = 0;
            for (int k = 0; k <= N; k++)
                for (int id = 0; id < 2 * NG; id++) {
                    const double msk = id < NG ? w->lin[k].ml[id] : w->lin[k].mu[id - NG];
                    if (msk > 0.5) mu_c += (lam[k][id] + alpha_c * w->dlam[k][id]) * (t[k][id] + alpha_c * w->dt[k][id]);
                }
            mu_c *= inv_nc;
            if (!(mu_c > 2.0 * mu_aff)) break;
            info->ncond++;
        }
        /* keep strictly inside the cone (HPIPM's step scaling) */
        alpha_p *= (1.0 - alpha_p) * 0.99 + alpha_p * 0.9999999;
        alpha_d *= (1.0 - alpha_d) * 0.99 + alpha_d * 0.9999999;
        if (alpha_p < IPM_ALPHA_MIN && alpha_d < IPM_ALPHA_MIN) { info->status = 2; break; }
        /* ---- update (split step: primal with alpha_p, dual with alpha_d) ---- */
        for (int k = 0; k <= N; k++) {
            for (int c = 0; c < NZ; c++) z[k][c] += alpha_p * w->dz[k][c];
            if (k < N) for (int i = 0; i < NX; i++) pi[k][i] += alpha_d * w->dpi[k][i];
            for (int id = 0; id < 2 * NG; id++) {
                const double msk = id < NG ? w->lin[k].ml[id] : w->lin[k].mu[id - NG];
                if (msk < 0.5) continue;
                t[k][id] += alpha_p * w->dt[k][id];
                lam[k][id] += alpha_d * w->dlam[k][id];
                if (t[k][id] < IPM_T_MIN) t[k][id] = IPM_T_MIN;
                if (lam[k][id] < IPM_LAM_MIN) lam[k][id] = IPM_LAM_MIN;
            }
        }
    }
    info->iter = it;
}

#ifndef NMPC_ORACLE_TOPT
#include "nmpc_oracle_fast.h"
#else
typedef struct { int unused; } fast_ws;       /* the structured variant exists for nu = 30 only */
#endif

/* ------------------------------------------------------------------------------------------ */
/* batch state ("handle")                                                                     */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
    nmpc_opts o;
    int B, N;
    int fast;                            /* 1: structured interior point (nmpc_oracle_fast.h), the timed CPU baseline */
    double *x, *u, *pi, *lam;            /* iterate */
    double *x0, *p, *yref, *yref_e, *W, *We;
    double *qz, *qpi, *qlam, *qt;        /* stored QP solution (warm start) */
    int32_t *status, *sqp_iter, *qp_iter, *qp_status, *qp_ncond;
    double *res, *qp_res, *cost;
} oracle_handle;

#define ALLOC(ptr, n) do { (ptr) = calloc((size_t)(n), sizeof(*(ptr))); } while (0)

oracle_handle* oracle_create(const nmpc_opts* o, int32_t B)
{
#ifdef NMPC_ORACLE_TOPT
    if (!o || B <= 0 || o->n_nodes < 1 || o->n_nodes > NMAX || !o->enable_time_opt) return NULL;
#else
    if (!o || B <= 0 || o->n_nodes < 1 || o->n_nodes > NMAX || o->enable_time_opt) return NULL;
#endif
    oracle_handle* h = calloc(1, sizeof(*h));
    h->o = *o; h->B = B; h->N = o->n_nodes;
    const size_t N = h->N;
    ALLOC(h->x, B * (N + 1) * NX); ALLOC(h->u, B * N * NU); ALLOC(h->pi, B * N * NX);
    ALLOC(h->lam, B * (N + 1) * 2 * NG);
    ALLOC(h->x0, B * NX); ALLOC(h->p, B * (N + 1) * NP); ALLOC(h->yref, B * N * NY);
    ALLOC(h->yref_e, B * NYE); ALLOC(h->W, B * NY); ALLOC(h->We, B * NYE);
    ALLOC(h->qz, B * (N + 1) * NZ); ALLOC(h->qpi, B * N * NX);
    ALLOC(h->qlam, B * (N + 1) * 2 * NG); ALLOC(h->qt, B * (N + 1) * 2 * NG);
    ALLOC(h->status, B); ALLOC(h->sqp_iter, B); ALLOC(h->qp_iter, B); ALLOC(h->qp_status, B); ALLOC(h->qp_ncond, B);
    ALLOC(h->res, B * 4); ALLOC(h->qp_res, B * 4); ALLOC(h->cost, B);
    return h;
}

void oracle_destroy(oracle_handle* h)
{
    if (!h) return;
    free(h->x); free(h->u); free(h->pi); free(h->lam); free(h->x0); free(h->p); free(h->yref);
    free(h->yref_e); free(h->W); free(h->We); free(h->qz); free(h->qpi); free(h->qlam); free(h->qt);
    free(h->status); free(h->sqp_iter); free(h->qp_iter); free(h->qp_status); free(h->qp_ncond);
    free(h->res); free(h->qp_res); free(h->cost);
    free(h);
}

int oracle_set_opts(oracle_handle* h, const nmpc_opts* o)
{
#ifdef NMPC_ORACLE_TOPT
    if (!h || !o || o->n_nodes != h->N || !o->enable_time_opt) return NMPC_EINVAL;
#else
    if (!h || !o || o->n_nodes != h->N || o->enable_time_opt) return NMPC_EINVAL;
#endif
    h->o = *o;
    return 0;
}

/* 0: dense restatement (the checker), 1: structured variant (timed CPU baseline; same iterates to rounding) */
int oracle_set_fast(oracle_handle* h, int32_t fast)
{
    if (!h) return NMPC_EINVAL;
#ifdef NMPC_ORACLE_TOPT
    if (fast) return NMPC_ENOSYS;
#endif
    h->fast = fast != 0;
    return 0;
}

int oracle_kinematics(const double* qv, int32_t n, double* feet, double* com, double* hg)
{
    for (int i = 0; i < n; i++) {
        double f[12], c[3], hh[6];
        nmpc_model_kin0(qv + 36 * i, qv + 36 * i + 18, f, c, hh);
        if (feet) memcpy(feet + 12 * i, f, sizeof(f));
        if (com) memcpy(com + 3 * i, c, sizeof(c));
        if (hg) memcpy(hg + 6 * i, hh, sizeof(hh));
    }
    return 0;
}

int oracle_inverse_dynamics(const double* qva, const double* f, int32_t n, double* tau)
{
    for (int i = 0; i < n; i++)
        nmpc_model_rnea(qva + 54 * i, qva + 54 * i + 18, qva + 54 * i + 36, f + 12 * i, tau + 12 * i);
    return 0;
}

/* full kinematics record (tests compare it with finite differences of an independent model) */
int oracle_model_record(const double* qv, double* rec_out)
{
    double* rec = rec_out;
    nmpc_model_kin(qv, qv + 18, rec);
    return NMPC_REC_SIZE;
}

int oracle_set_x0(oracle_handle* h, const double* qv)
{
    for (int b = 0; b < h->B; b++) {
        double* x0 = h->x0 + (size_t)b * NX;
        memcpy(x0, qv + 36 * b, 36 * sizeof(double));
        double f[12], c[3];
        nmpc_model_kin0(x0, x0 + 18, f, c, x0 + 36);
        memcpy(h->x + (size_t)b * (h->N + 1) * NX, x0, NX * sizeof(double));
    }
    return 0;
}

int oracle_set_x0_full(oracle_handle* h, const double* x0)
{
    memcpy(h->x0, x0, sizeof(double) * h->B * NX);
    for (int b = 0; b < h->B; b++) memcpy(h->x + (size_t)b * (h->N + 1) * NX, x0 + (size_t)b * NX, NX * sizeof(double));
    return 0;
}

int oracle_set_yref(oracle_handle* h, const double* yref, const double* yref_e)
{
    if (yref) memcpy(h->yref, yref, sizeof(double) * h->B * h->N * NY);
    if (yref_e) memcpy(h->yref_e, yref_e, sizeof(double) * h->B * NYE);
    return 0;
}

int oracle_set_weights(oracle_handle* h, const double* W, const double* We, int32_t per_problem)
{
    for (int b = 0; b < h->B; b++) {
        if (W) memcpy(h->W + (size_t)b * NY, W + (per_problem ? (size_t)b * NY : 0), NY * sizeof(double));
        if (We) memcpy(h->We + (size_t)b * NYE, We + (per_problem ? (size_t)b * NYE : 0), NYE * sizeof(double));
    }
    return 0;
}

int oracle_set_params(oracle_handle* h, const double* p)
{
    memcpy(h->p, p, sizeof(double) * h->B * (h->N + 1) * NP);
    return 0;
}

/* solver.py:194-210 */
int oracle_pin_stance_feet(oracle_handle* h, const int32_t* i_node)
{
    const int N = h->N;
    for (int b = 0; b < h->B; b++) {
        double* p = h->p + (size_t)b * (N + 1) * NP;
        const double* x0 = h->x0 + (size_t)b * NX;
        double feet[12], c[3], hh[6];
        nmpc_model_kin0(x0, x0 + 18, feet, c, hh);
        for (int ft = 0; ft < NF; ft++) {
            if (i_node[b] == 0) p[ft * NMPC_NP_FOOT + P_ACTIVE] = 1.0;
            if (p[ft * NMPC_NP_FOOT + P_ACTIVE] != 0.0) {
                /* next_swing = argmin(active[0, :]) : first node with the smallest value */
                int next_swing = 0;
                double best = p[ft * NMPC_NP_FOOT + P_ACTIVE];
                for (int k = 1; k <= N; k++) {
                    double a = p[k * NP + ft * NMPC_NP_FOOT + P_ACTIVE];
                    if (a < best) { best = a; next_swing = k; }
                }
                for (int k = 0; k < next_swing; k++)
                    for (int i = 0; i < 3; i++) p[k * NP + ft * NMPC_NP_FOOT + P_POINT + i] = feet[3 * ft + i];
            }
        }
    }
    return 0;
}

int oracle_set_guess(oracle_handle* h, const double* x, const double* u)
{
    if (x) memcpy(h->x, x, sizeof(double) * h->B * (h->N + 1) * NX);
    if (u) memcpy(h->u, u, sizeof(double) * h->B * h->N * NU);
    return 0;
}

int oracle_reset_duals(oracle_handle* h)
{
    const size_t B = h->B, N = h->N;
    memset(h->pi, 0, sizeof(double) * B * N * NX);
    memset(h->lam, 0, sizeof(double) * B * (N + 1) * 2 * NG);
    memset(h->qz, 0, sizeof(double) * B * (N + 1) * NZ);
    memset(h->qpi, 0, sizeof(double) * B * N * NX);
    memset(h->qlam, 0, sizeof(double) * B * (N + 1) * 2 * NG);
    memset(h->qt, 0, sizeof(double) * B * (N + 1) * 2 * NG);
    return 0;
}

/* solver.py:290-342: shift by start_node; states 1..n_ws from start+1.., inputs 0..n_ws-1 from
 * start.., force tail zero (or last), rest left as it was; duals shifted the same way. */
int oracle_shift_warm_start(oracle_handle* h, const int32_t* start_node, int32_t repeat_last)
{
    const int N = h->N;
    for (int b = 0; b < h->B; b++) {
        const int s = start_node[b];
        if (s <= 0 || s > N) continue;
        const int nws = N - s;
        double* x = h->x + (size_t)b * (N + 1) * NX;
        double* u = h->u + (size_t)b * N * NU;
        double* pi = h->pi + (size_t)b * N * NX;
        double* lam = h->lam + (size_t)b * (N + 1) * 2 * NG;
        for (int k = 1; k <= nws; k++) memcpy(x + k * NX, x + (k + s) * NX, NX * sizeof(double));
        for (int k = 0; k < nws; k++) memcpy(u + k * NU, u + (k + s) * NU, NU * sizeof(double));
        for (int k = nws; k < N; k++)
            for (int i = 0; i < 12; i++) u[k * NU + IFRC + i] = repeat_last ? u[(N - 1) * NU + IFRC + i] : 0.0;
#ifdef NMPC_ORACLE_TOPT
        for (int k = nws; k < N; k++) u[k * NU + IDT] = h->o.dt_nodes;      /* solver.py:334-336: nominal dt on the tail */
#endif
        if (repeat_last && nws < N) {
            for (int k = nws; k <= N; k++) memcpy(x + k * NX, x + N * NX, NX * sizeof(double));
            for (int k = nws; k < N; k++) memcpy(u + k * NU, u + (N - 1) * NU, NQ * sizeof(double));
        }
        for (int k = 0; k < nws; k++) memcpy(pi + k * NX, pi + (k + s) * NX, NX * sizeof(double));
        for (int k = 0; k <= nws; k++) memcpy(lam + k * 2 * NG, lam + (k + s) * 2 * NG, 2 * NG * sizeof(double));
    }
    return 0;
}

/* linearise all stages of problem b at its current iterate into w->lin */
static void linearize_all(const oracle_handle* h, int b, workspace* w)
{
    const int N = h->N;
    const double* x = h->x + (size_t)b * (N + 1) * NX;
    const double* u = h->u + (size_t)b * N * NU;
    const double* p = h->p + (size_t)b * (N + 1) * NP;
    for (int k = 0; k <= N; k++) {
        const int term = k == N;
        linearize_stage(&h->o, k, N, x + k * NX, term ? NULL : u + k * NU, term ? NULL : x + (k + 1) * NX,
                        p + k * NP, term ? h->yref_e + (size_t)b * NYE : h->yref + ((size_t)b * N + k) * NY,
                        term ? h->We + (size_t)b * NYE : h->W + (size_t)b * NY, &w->lin[k]);
    }
}

/* NLP residual inf-norms at the linearisation in w->lin with multipliers pi, lam */
static void nlp_residuals(const oracle_handle* h, int b, const workspace* w, double res[4], double* cost)
{
    const int N = h->N;
    const double* pi = h->pi + (size_t)b * N * NX;
    const double* lam = h->lam + (size_t)b * (N + 1) * 2 * NG;
    const double* x = h->x + (size_t)b * (N + 1) * NX;
    const double* x0 = h->x0 + (size_t)b * NX;
    double rs = 0, re = 0, ri = 0, rc = 0, cst = 0;
    for (int i = 0; i < NX; i++) NORM_ACC(re, x[i] - x0[i]);
    for (int k = 0; k <= N; k++) {
        const stage_lin* s = &w->lin[k];
        cst += s->cost;
        for (int c = 0; c < NZ; c++) {
            if ((k == 0 && c >= NU) || (k == N && c < NU)) continue;
            double acc = s->g[c];
            if (k < N) for (int l = 0; l < NX; l++) acc += (c < NU ? s->B[l][c] : s->A[l][c - NU]) * pi[k * NX + l];
            if (k > 0 && c >= NU) acc -= pi[(k - 1) * NX + c - NU];
            for (int j = 0; j < NG; j++) acc -= s->C[j][c] * (s->ml[j] * lam[k * 2 * NG + j] - s->mu[j] * lam[k * 2 * NG + NG + j]);
            NORM_ACC(rs, acc);
        }
        if (k < N) for (int i = 0; i < NX; i++) NORM_ACC(re, s->b[i]);
        for (int j = 0; j < NG; j++) {
            if (s->ml[j] > 0.5) {
                if (-s->cl[j] > ri || s->cl[j] != s->cl[j]) ri = -s->cl[j];
                NORM_ACC(rc, lam[k * 2 * NG + j] * s->cl[j]);
            }
            if (s->mu[j] > 0.5) {
                if (s->cu[j] > ri || s->cu[j] != s->cu[j]) ri = s->cu[j];
                NORM_ACC(rc, lam[k * 2 * NG + NG + j] * s->cu[j]);
            }
        }
    }
    res[0] = rs; res[1] = re; res[2] = ri; res[3] = rc; *cost = cst;
}

/* one QP on the linearisation in w->lin, dense or structured */
static void qp_solve(const oracle_handle* h, workspace* w, fast_ws* f, int N, const double* dx0, int warm,
                     double (*z)[NZ], double (*pi)[NX], double (*lam)[2 * NG], double (*t)[2 * NG], qp_info* info)
{
#ifndef NMPC_ORACLE_TOPT
    if (h->fast && f) {
        int bad = 0;
        for (int k = 0; k <= N; k++) bad |= sr_from_lin(&w->lin[k], k == N, h->o.dt_nodes, &f->sr[k]);
        if (!bad) { qp_ipm_fast(w, f, &h->o, N, dx0, warm, z, pi, lam, t, info); return; }
    }
#else
    (void)f;
#endif
    qp_ipm(w, &h->o, N, dx0, warm, z, pi, lam, t, info);
}

static void solve_one(oracle_handle* h, int b, workspace* w, fast_ws* f)
{
    const int N = h->N;
    const nmpc_opts* o = &h->o;
    double* x = h->x + (size_t)b * (N + 1) * NX;
    double* u = h->u + (size_t)b * N * NU;
    double* pi = h->pi + (size_t)b * N * NX;
    double* lam = h->lam + (size_t)b * (N + 1) * 2 * NG;
    double (*qz)[NZ] = (double (*)[NZ])(h->qz + (size_t)b * (N + 1) * NZ);
    double (*qpi)[NX] = (double (*)[NX])(h->qpi + (size_t)b * N * NX);
    double (*qlam)[2 * NG] = (double (*)[2 * NG])(h->qlam + (size_t)b * (N + 1) * 2 * NG);
    double (*qt)[2 * NG] = (double (*)[2 * NG])(h->qt + (size_t)b * (N + 1) * 2 * NG);
    const double* x0 = h->x0 + (size_t)b * NX;
    int status = NMPC_STATUS_MAXITER;
    int it = 0;
    const int max_iter = o->real_time_it ? 1 : o->max_iter;
    h->qp_iter[b] = 0; h->qp_status[b] = 0; h->qp_ncond[b] = 0;
    for (;; it++) {
        linearize_all(h, b, w);
        nlp_residuals(h, b, w, h->res + 4 * b, h->cost + b);
        const double* r = h->res + 4 * b;
        if (!(r[0] == r[0]) || !(r[1] == r[1]) || !(r[2] == r[2]) || !(r[3] == r[3])) { status = NMPC_STATUS_NAN; break; }
        if (!o->real_time_it && r[0] < o->nlp_tol && r[1] < o->nlp_tol && r[2] < o->nlp_tol && r[3] < o->nlp_tol) {
            status = NMPC_STATUS_SUCCESS; break;
        }
        if (it >= max_iter) { status = NMPC_STATUS_MAXITER; break; }
        double dx0[NX];
        for (int i = 0; i < NX; i++) dx0[i] = x0[i] - x[i];
        qp_info info;
        /* acados ocp_nlp_sqp: "(typically) no warm start at first iteration" -- the QP solver's warm_start option
         * is forced to 0 for the first QP of a solve unless opts->warm_start_first_qp, then restored */
        const int warm = (it == 0 && !o->warm_start_first_qp) ? 0 : o->warm_start_qp;
        qp_solve(h, w, f, N, dx0, warm, qz, qpi, qlam, qt, &info);
        h->qp_iter[b] = info.iter; h->qp_status[b] = info.status; h->qp_ncond[b] = info.ncond;
        memcpy(h->qp_res + 4 * b, info.res, sizeof(info.res));
        if (info.status != 0 && info.status != 1) { status = NMPC_STATUS_QP_FAILURE; it++; break; }
        /* full step; multipliers are replaced by the QP's */
        for (int k = 0; k <= N; k++) {
            for (int i = 0; i < NX; i++) x[k * NX + i] += qz[k][NU + i];
            if (k < N) for (int i = 0; i < NU; i++) u[k * NU + i] += qz[k][i];
        }
        memcpy(pi, qpi, sizeof(double) * N * NX);
        for (int k = 0; k <= N; k++)
            for (int j = 0; j < NG; j++) {
                lam[k * 2 * NG + j] = w->lin[k].ml[j] > 0.5 ? qlam[k][j] : 0.0;
                lam[k * 2 * NG + NG + j] = w->lin[k].mu[j] > 0.5 ? qlam[k][NG + j] : 0.0;
            }
        /* SQP_RTI: one linearisation, one QP, one full step, no convergence test */
        if (o->real_time_it) { status = NMPC_STATUS_SUCCESS; it++; break; }
    }
    h->status[b] = status;
    h->sqp_iter[b] = it;
}

int oracle_solve(oracle_handle* h, int32_t* status, int32_t n_threads)
{
    if (!h) return NMPC_EINVAL;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#pragma omp parallel
#endif
    {
        workspace* w = malloc(sizeof(workspace));
        fast_ws* f = h->fast ? malloc(sizeof(fast_ws)) : NULL;
#ifdef _OPENMP
#pragma omp for schedule(dynamic)
#endif
        for (int b = 0; b < h->B; b++) solve_one(h, b, w, f);
        free(w);
        free(f);
    }
    if (status) memcpy(status, h->status, sizeof(int32_t) * h->B);
    return 0;
}

int oracle_get_solution(oracle_handle* h, double* x, double* u, double* pi, double* lam)
{
    const size_t B = h->B, N = h->N;
    if (x) memcpy(x, h->x, sizeof(double) * B * (N + 1) * NX);
    if (u) memcpy(u, h->u, sizeof(double) * B * N * NU);
    if (pi) memcpy(pi, h->pi, sizeof(double) * B * N * NX);
    if (lam) memcpy(lam, h->lam, sizeof(double) * B * (N + 1) * 2 * NG);
    return 0;
}

int oracle_get_stats(oracle_handle* h, nmpc_stats* s)
{
    const size_t B = h->B;
    if (s->sqp_iter) memcpy(s->sqp_iter, h->sqp_iter, sizeof(int32_t) * B);
    if (s->qp_iter) memcpy(s->qp_iter, h->qp_iter, sizeof(int32_t) * B);
    if (s->qp_status) memcpy(s->qp_status, h->qp_status, sizeof(int32_t) * B);
    if (s->res) memcpy(s->res, h->res, sizeof(double) * B * 4);
    if (s->qp_res) memcpy(s->qp_res, h->qp_res, sizeof(double) * B * 4);
    if (s->cost) memcpy(s->cost, h->cost, sizeof(double) * B);
    if (s->qp_ncond) memcpy(s->qp_ncond, h->qp_ncond, sizeof(int32_t) * B);
    return 0;
}

int oracle_get_qp_step(oracle_handle* h, double* dz)
{
    memcpy(dz, h->qz, sizeof(double) * h->B * (h->N + 1) * NZ);
    return 0;
}

/* export the dense QP of the current iterate in the product ABI's layouts (nmpc_export_qp) */
int oracle_export_qp(oracle_handle* h, double* BAbt, double* RSQrq, double* DCt, double* lg,
                     double* ug, double* mask)
{
    const int N = h->N;
    workspace* w = malloc(sizeof(workspace));
    for (int b = 0; b < h->B; b++) {
        linearize_all(h, b, w);
        for (int k = 0; k <= N; k++) {
            const stage_lin* s = &w->lin[k];
            if (BAbt && k < N) {
                double* o = BAbt + ((size_t)b * N + k) * (NZ + 1) * NX;
                for (int c = 0; c < NZ; c++)
                    for (int i = 0; i < NX; i++) o[c * NX + i] = c < NU ? s->B[i][c] : s->A[i][c - NU];
                for (int i = 0; i < NX; i++) o[NZ * NX + i] = s->b[i];
            }
            if (RSQrq) {
                double* o = RSQrq + ((size_t)b * (N + 1) + k) * (NZ + 1) * NZ;
                for (int c = 0; c < NZ; c++)
                    for (int d = 0; d < NZ; d++) o[c * NZ + d] = s->H[c][d];
                for (int d = 0; d < NZ; d++) o[NZ * NZ + d] = s->g[d];
            }
            if (DCt) {
                double* o = DCt + ((size_t)b * (N + 1) + k) * NZ * NG;
                for (int c = 0; c < NZ; c++)
                    for (int j = 0; j < NG; j++) o[c * NG + j] = s->C[j][c];
            }
            for (int j = 0; j < NG; j++) {
                const size_t id = ((size_t)b * (N + 1) + k) * NG + j;
                if (lg) lg[id] = -s->cl[j];
                if (ug) ug[id] = -s->cu[j];
                if (mask) {
                    mask[((size_t)b * (N + 1) + k) * 2 * NG + j] = s->ml[j];
                    mask[((size_t)b * (N + 1) + k) * 2 * NG + NG + j] = s->mu[j];
                }
            }
        }
    }
    free(w);
    return 0;
}

/* solve externally supplied QPs (layouts of nmpc_solve_qp), cold start */
int oracle_solve_qp(oracle_handle* h, const double* BAbt, const double* RSQrq, const double* DCt,
                    const double* lg, const double* ug, const double* mask, const double* dx0,
                    double* dz, double* pi, double* lam, double* t, int32_t* qp_iter, int32_t* qp_status,
                    double* qp_res)
{
    const int N = h->N;
    workspace* w = malloc(sizeof(workspace));
    fast_ws* f = h->fast ? malloc(sizeof(fast_ws)) : NULL;
    for (int b = 0; b < h->B; b++) {
        for (int k = 0; k <= N; k++) {
            stage_lin* s = &w->lin[k];
            memset(s, 0, sizeof(*s));
            if (k < N) {
                const double* o = BAbt + ((size_t)b * N + k) * (NZ + 1) * NX;
                for (int c = 0; c < NZ; c++)
                    for (int i = 0; i < NX; i++) { if (c < NU) s->B[i][c] = o[c * NX + i]; else s->A[i][c - NU] = o[c * NX + i]; }
                for (int i = 0; i < NX; i++) s->b[i] = o[NZ * NX + i];
            }
            const double* o = RSQrq + ((size_t)b * (N + 1) + k) * (NZ + 1) * NZ;
            for (int c = 0; c < NZ; c++)
                for (int d = 0; d <= c; d++) { s->H[c][d] = o[c * NZ + d]; s->H[d][c] = o[c * NZ + d]; }
            for (int d = 0; d < NZ; d++) s->g[d] = o[NZ * NZ + d];
            const double* dc = DCt + ((size_t)b * (N + 1) + k) * NZ * NG;
            for (int c = 0; c < NZ; c++)
                for (int j = 0; j < NG; j++) s->C[j][c] = dc[c * NG + j];
            for (int j = 0; j < NG; j++) {
                const size_t id = ((size_t)b * (N + 1) + k) * NG + j;
                s->ml[j] = mask[((size_t)b * (N + 1) + k) * 2 * NG + j];
                s->mu[j] = mask[((size_t)b * (N + 1) + k) * 2 * NG + NG + j];
                s->lgq[j] = lg[id]; s->ugq[j] = ug[id];
            }
        }
        qp_info info;
        qp_solve(h, w, f, N, dx0 + (size_t)b * NX, 0,
                 (double (*)[NZ])(dz + (size_t)b * (N + 1) * NZ), (double (*)[NX])(pi + (size_t)b * N * NX),
                 (double (*)[2 * NG])(lam + (size_t)b * (N + 1) * 2 * NG),
                 (double (*)[2 * NG])(t + (size_t)b * (N + 1) * 2 * NG), &info);
        if (qp_iter) qp_iter[b] = info.iter;
        if (qp_status) qp_status[b] = info.status;
        if (qp_res) memcpy(qp_res + 4 * b, info.res, sizeof(info.res));
    }
    free(w);
    free(f);
    return 0;
}

double oracle_model_mass(void) { return NMPC_MASS; }

/* ------------------------------------------------------------------------------------------ */
/* batched rigid-body rollout: CPU restatement of nmpc_rollout.cuh (checker of nmpc_sim_steps) */
/* ------------------------------------------------------------------------------------------ */
#include "../iterative_learning_nmpc_b200/csrc/nmpc_gen_sim.h"

/* sample ip of the interpolated plan: scipy CubicHermiteSpline of (q, v) and (v, a') on the node grid, evaluated the
 * way mpc.py:388-414 samples it (linspace(0, T, n + 1)[1:]); dtn: per-node durations */
static void plan_sample(const double* x, const double* u, const double* dtn, int N, int n, int ip, double* q, double* v)
{
    double T = 0.0;
    for (int k = 0; k < N; k++) T += dtn[k];
    const double step = T / n;
    const int j = ip + 1;
    const double t = (j >= n) ? T : j * step;
    int k = 0;
    double tk = 0.0, tk1 = dtn[0];
    while (k < N - 1 && t >= tk1) { k++; tk = tk1; tk1 += dtn[k]; }
    const double hh = tk1 - tk, s = (t - tk) / hh, s2 = s * s, s3 = s2 * s;
    const double h00 = 2.0 * s3 - 3.0 * s2 + 1.0, h10 = s3 - 2.0 * s2 + s, h01 = -2.0 * s3 + 3.0 * s2, h11 = s3 - s2;
    const double* x0 = x + (size_t)k * NX;
    const double* x1 = x0 + NX;
    const double* a0 = u + (size_t)(k > 0 ? k - 1 : 0) * NU;
    const double* a1 = u + (size_t)k * NU;
    for (int c = 0; c < NQ; c++) {
        q[c] = h00 * x0[c] + h10 * hh * x0[NQ + c] + h01 * x1[c] + h11 * hh * x1[NQ + c];
        v[c] = h00 * x0[NQ + c] + h10 * hh * a0[c] + h01 * x1[NQ + c] + h11 * hh * a1[c];
    }
}

typedef struct { double sim_dt, kp, kd, k_n, d_n, d_t, mu, ground_z, tau_max; } sim_params;

static void sim_physics(const sim_params* sp, double* q, double* v, const double* tau, double* fc)
{
    double M[NQ * (NQ + 1) / 2], r[NQ], fp[12], fv[12], fJ[12 * NQ];
    nmpc_sim_mass(q, M);
    nmpc_sim_bias(q, v, r);
    nmpc_sim_feet(q, v, fp, fv, fJ);
    for (int i = 0; i < NQ; i++) r[i] = -r[i];
    for (int j = 0; j < 12; j++) r[6 + j] += tau[j];
    for (int ft = 0; ft < NF; ft++) {
        double* f = fc + 3 * ft;
        f[0] = f[1] = f[2] = 0.0;
        const double pen = sp->ground_z - fp[3 * ft + 2];
        if (pen > 0.0) {
            const double fn = sp->k_n * pen - sp->d_n * fv[3 * ft + 2];
            if (fn > 0.0) {
                double fx = -sp->d_t * fv[3 * ft], fy = -sp->d_t * fv[3 * ft + 1];
                const double ftn = sqrt(fx * fx + fy * fy), cap = sp->mu * fn;
                if (ftn > cap) { const double sc = cap / ftn; fx *= sc; fy *= sc; }
                f[0] = fx; f[1] = fy; f[2] = fn;
            }
        }
        for (int c = 0; c < 3; c++)
            if (f[c] != 0.0)
                for (int i = 0; i < NQ; i++) r[i] += fJ[(3 * ft + c) * NQ + i] * f[c];
    }
    /* dense Cholesky solve M a = r on the packed lower triangle */
    for (int j = 0; j < NQ; j++) {
        double d = M[j * (j + 1) / 2 + j];
        for (int l = 0; l < j; l++) d -= M[j * (j + 1) / 2 + l] * M[j * (j + 1) / 2 + l];
        d = sqrt(d);
        M[j * (j + 1) / 2 + j] = d;
        const double inv = 1.0 / d;
        for (int i = j + 1; i < NQ; i++) {
            double s = M[i * (i + 1) / 2 + j];
            for (int l = 0; l < j; l++) s -= M[i * (i + 1) / 2 + l] * M[j * (j + 1) / 2 + l];
            M[i * (i + 1) / 2 + j] = s * inv;
        }
    }
    for (int i = 0; i < NQ; i++) {
        double s = r[i];
        for (int l = 0; l < i; l++) s -= M[i * (i + 1) / 2 + l] * r[l];
        r[i] = s / M[i * (i + 1) / 2 + i];
    }
    for (int i = NQ - 1; i >= 0; i--) {
        double s = r[i];
        for (int l = i + 1; l < NQ; l++) s -= M[l * (l + 1) / 2 + i] * r[l];
        r[i] = s / M[i * (i + 1) / 2 + i];
    }
    for (int i = 0; i < NQ; i++) { v[i] += r[i] * sp->sim_dt; q[i] += v[i] * sp->sim_dt; }
}

/* forward-dynamics terms of the model (tests: M a + bias against the generated inverse dynamics) */
int oracle_sim_terms(const double* qv, double* M_full, double* bias)
{
    double M[NQ * (NQ + 1) / 2];
    nmpc_sim_mass(qv, M);
    nmpc_sim_bias(qv, qv + NQ, bias);
    for (int i = 0; i < NQ; i++)
        for (int j = 0; j <= i; j++) { M_full[i * NQ + j] = M[i * (i + 1) / 2 + j]; M_full[j * NQ + i] = M[i * (i + 1) / 2 + j]; }
    return 0;
}

/* physics only: n_steps steps from joint torques tau [n_steps][12] (or a constant tau when hold != 0) */
int oracle_sim_physics(const double* params, double* qv, const double* tau, int32_t n_steps, int32_t hold, double* out_qv,
                       double* out_fc)
{
    sim_params sp;
    memcpy(&sp, params, sizeof(sp));
    double* q = qv;
    double* v = qv + NQ;
    for (int s = 0; s < n_steps; s++) {
        double fc[12];
        if (out_qv) memcpy(out_qv + (size_t)s * 36, qv, 36 * sizeof(double));
        sim_physics(&sp, q, v, tau + (hold ? 0 : (size_t)s * 12), fc);
        if (out_fc) memcpy(out_fc + (size_t)s * 12, fc, sizeof(fc));
    }
    return 0;
}

/* closed loop on the plan held by the handle: the controller of LocomotionMPC.compute_torques_dof + physics */
int oracle_sim_steps(oracle_handle* h, const double* params, int32_t n_interp, int32_t step0, int32_t n_steps, double* qv,
                     double* out_qv, double* out_tau, double* out_fc)
{
    const int N = h->N, n = n_interp;
    sim_params sp;
    memcpy(&sp, params, sizeof(sp));
    for (int b = 0; b < h->B; b++) {
        const double* x = h->x + (size_t)b * (N + 1) * NX;
        const double* u = h->u + (size_t)b * N * NU;
        double dtn[NMAX];
        for (int k = 0; k < N; k++) {
#ifdef NMPC_ORACLE_TOPT
            dtn[k] = u[k * NU + IDT];
#else
            dtn[k] = h->o.dt_nodes;
#endif
        }
        double* q = qv + (size_t)b * 36;
        double* v = q + NQ;
        for (int s = 0; s < n_steps; s++) {
            const int ip = step0 + s;
            double qr[NQ], vr[NQ], tq[12], fc[12];
            if (out_qv) memcpy(out_qv + ((size_t)b * n_steps + s) * 36, q, 36 * sizeof(double));
            plan_sample(x, u, dtn, N, n, ip + 1 < n ? ip + 1 : n - 1, qr, vr);
            const int ih = ip < n ? ip : n - 1;
            const double frac = (ih >= n - 1) ? 1.0 : ih * (1.0 / (n - 1));
            const double* ur = u + (size_t)((int)(frac * (N - 1))) * NU;
            nmpc_model_rnea(q, v, ur, ur + NQ, tq);
            for (int c = 0; c < 12; c++) {
                double t = tq[c] + sp.kp * (qr[6 + c] - q[6 + c]) + sp.kd * (vr[6 + c] - v[6 + c]);
                if (sp.tau_max > 0.0) t = fmin(fmax(t, -sp.tau_max), sp.tau_max);
                tq[c] = t;
            }
            sim_physics(&sp, q, v, tq, fc);
            if (out_tau) memcpy(out_tau + ((size_t)b * n_steps + s) * 12, tq, sizeof(tq));
            if (out_fc) memcpy(out_fc + ((size_t)b * n_steps + s) * 12, fc, sizeof(fc));
        }
    }
    return 0;
}

int oracle_dims(int32_t* out)
{
    out[0] = NX; out[1] = NU; out[2] = NP; out[3] = NY; out[4] = NYE; out[5] = NG; out[6] = NMPC_REC_SIZE;
    return 7;
}
