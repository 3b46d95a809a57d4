/* nmpc_oracle_fast.h -- structured variant of the OCP-QP interior point of nmpc_oracle.c.
 * TEST INFRASTRUCTURE ONLY (included by nmpc_oracle.c): it is (1) the timed CPU baseline of bench.py -- a fairer
 * opponent than the dense restatement: it skips the structural zeros of the stage matrices, factors only the
 * 30 x 30 input block per stage (classical Riccati recursion on P instead of a 72-pivot square-root factorisation)
 * and solves the corrector with stored gains -- and (2) the CPU statement of exactly the arithmetic organisation the
 * CUDA kernel uses (csrc/nmpc_qp_fused.cuh), so that a formula can be checked on the host first.
 * tests/test_oracle_qp.py asserts that it reproduces the dense restatement (qp_ipm) to rounding.
 *
 * Same algorithm, same iterates in exact arithmetic: HPIPM `speed` Mehrotra predictor-corrector with conditional
 * predictor-corrector, split step, warm-start levels 0/1/2 (see qp_ipm).  Structure used (this OCP family):
 *   dynamics  q+ = q + dt v + c2 a, v+ = v + dt a, h+ = h + Hq q + Hf f + b            (c2 = dt^2/2)
 *   Hessian   diagonal in (a, f, v, h), dense in the 18 x 18 q block
 *   C         friction rows touch one foot's force (CF), the other rows touch (q, v) (CX) plus an identity in h
 *             for the six centroidal rows
 * Per stage the Riccati step keeps K = Muu^-1 Mux (as Kt = K', 42 x 30), Linv = Luu^-1 and P, so that both forward
 * sweeps and the corrector's backward sweep are chains of small matrix-vector products (du = -(K dx + k),
 * p = m_x - K' m_u, k = Linv' Linv m_u) with no triangular substitution on them.
 */

typedef struct {
    double Hq[6][NQ], Hf[6][12], b[NX];
    double Ru[NU], Qv[NQ], Qh[6], Qqq[NQ][NQ], g[NZ];
    double CF[20][3], CX[22][36];
    double lg[NG], ug[NG], ml[NG], mu[NG];
} stage_sr;

typedef struct {
    stage_sr sr[NMAX + 1];
    double P[NMAX + 1][NX][NX], p[NMAX + 1][NX];
    double Kt[NMAX][NX][NU];        /* Kt[i][j] = K[j][i],  K = Muu^-1 Mux */
    double kv[NMAX][NU];            /* Muu^-1 m_u of the current right-hand side */
    double Linv[NMAX][NU][NU];      /* Luu^-1, lower triangular */
    double Prb[NMAX][NX];           /* P_{k+1} rb_k (the same for the affine and the corrector solve) */
    double Gam[NMAX + 1][NG];
} fast_ws;

/* stage record from the dense linearisation; returns 0 if the stage has the structure described above */
static int sr_from_lin(const stage_lin* s, int term, double dt, stage_sr* r)
{
    int bad = 0;
    const double c2 = 0.5 * dt * dt;
    memset(r, 0, sizeof(*r));
    if (!term) {
        for (int i = 0; i < 6; i++) {
            for (int j = 0; j < NQ; j++) r->Hq[i][j] = s->A[2 * NQ + i][j];
            for (int j = 0; j < 12; j++) r->Hf[i][j] = s->B[2 * NQ + i][NQ + j];
        }
        for (int l = 0; l < NX; l++) r->b[l] = s->b[l];
        for (int l = 0; l < NX; l++)
            for (int c = 0; c < NZ; c++) {
                const double v = c < NU ? s->B[l][c] : s->A[l][c - NU];
                double ex = 0.0;
                if (l < NQ) { if (c == IA + l) ex = c2; else if (c == IQ + l) ex = 1.0; else if (c == IV + l) ex = dt; }
                else if (l < 2 * NQ) { const int i = l - NQ; if (c == IA + i) ex = dt; else if (c == IV + i) ex = 1.0; }
                else {
                    const int i = l - 2 * NQ;
                    if (c >= IFRC && c < IV) ex = v;
                    else if (c == IH + i) ex = 1.0;
                }
                if (v != ex) bad = 1;
            }
    }
    for (int c = 0; c < NU; c++) r->Ru[c] = s->H[c][c];
    for (int i = 0; i < NQ; i++) r->Qv[i] = s->H[IV + i][IV + i];
    for (int i = 0; i < 6; i++) r->Qh[i] = s->H[IH + i][IH + i];
    for (int i = 0; i < NQ; i++)
        for (int j = 0; j < NQ; j++) r->Qqq[i][j] = s->H[IQ + i][IQ + j];
    for (int c = 0; c < NZ; c++) r->g[c] = s->g[c];
    for (int i = 0; i < NZ; i++)
        for (int j = 0; j < NZ; j++) {
            const int in_q = (i >= IQ && i < IV && j >= IQ && j < IV);
            if (i != j && !in_q && s->H[i][j] != 0.0) bad = 1;
        }
    for (int j = 0; j < 20; j++)
        for (int i = 0; i < 3; i++) r->CF[j][i] = s->C[j][IFRC + 3 * (j / 5) + i];
    for (int j = 0; j < 22; j++)
        for (int c = 0; c < 36; c++) r->CX[j][c] = s->C[20 + j][IQ + c];
    for (int j = 0; j < NG; j++) {
        r->lg[j] = s->lgq[j]; r->ug[j] = s->ugq[j]; r->ml[j] = s->ml[j]; r->mu[j] = s->mu[j];
        if (!(s->ml[j] > 0.5 || s->mu[j] > 0.5)) continue;          /* a row without bounds contributes nothing */
        for (int c = 0; c < NZ; c++) {
            double ex;
            if (j < 20) ex = (c >= IFRC + 3 * (j / 5) && c < IFRC + 3 * (j / 5) + 3) ? s->C[j][c] : 0.0;
            else if (c >= IQ && c < IH) ex = s->C[j][c];
            else ex = (j >= GR_CONS && c == IH + j - GR_CONS) ? 1.0 : 0.0;
            if (s->C[j][c] != ex) bad = 1;
        }
    }
    return bad;
}

/* ---- structured products ---- */
static inline double f_cz(const stage_sr* r, const double* z, int j)
{
    if (j < 20) { const double* f = z + IFRC + 3 * (j / 5); return r->CF[j][0] * f[0] + r->CF[j][1] * f[1] + r->CF[j][2] * f[2]; }
    double s = 0.0;
    const double* cx = r->CX[j - 20];
    for (int c = 0; c < 36; c++) s += cx[c] * z[IQ + c];
    if (j >= GR_CONS) s += z[IH + j - GR_CONS];
    return s;
}
static inline double f_ct(const stage_sr* r, const double* g, int c)       /* (C' g)[c] */
{
    if (c < IFRC) return 0.0;
    double s = 0.0;
    if (c < IQ) { const int ft = (c - IFRC) / 3, i = (c - IFRC) % 3; for (int q = 0; q < 5; q++) s += r->CF[5 * ft + q][i] * g[5 * ft + q]; return s; }
    if (c < IH) { for (int q = 0; q < 22; q++) s += r->CX[q][c - IQ] * g[20 + q]; return s; }
    return g[GR_CONS + c - IH];
}
static inline double f_gt(const stage_sr* r, const double* w, int c, double dt, double c2)   /* (G' w)[c] */
{
    if (c < IFRC) return c2 * w[c] + dt * w[NQ + c];
    double s = 0.0;
    if (c < IQ) { for (int h = 0; h < 6; h++) s += r->Hf[h][c - IFRC] * w[2 * NQ + h]; return s; }
    if (c < IV) { s = w[c - IQ]; for (int h = 0; h < 6; h++) s += r->Hq[h][c - IQ] * w[2 * NQ + h]; return s; }
    if (c < IH) return dt * w[c - IV] + w[NQ + c - IV];
    return w[2 * NQ + c - IH];
}
static inline double f_g(const stage_sr* r, const double* z, int l, double dt, double c2)    /* (G z)[l] */
{
    if (l < NQ) return z[IQ + l] + dt * z[IV + l] + c2 * z[IA + l];
    if (l < 2 * NQ) return z[IV + l - NQ] + dt * z[IA + l - NQ];
    const int i = l - 2 * NQ;
    double s = z[IH + i];
    for (int j = 0; j < NQ; j++) s += r->Hq[i][j] * z[IQ + j];
    for (int j = 0; j < 12; j++) s += r->Hf[i][j] * z[IFRC + j];
    return s;
}
static inline double f_h(const stage_sr* r, const double* z, int c)        /* (H z)[c] */
{
    if (c < IQ) return r->Ru[c] * z[c];
    if (c < IV) { double s = 0.0; for (int j = 0; j < NQ; j++) s += r->Qqq[c - IQ][j] * z[IQ + j]; return s; }
    if (c < IH) return r->Qv[c - IV] * z[c];
    return r->Qh[c - IH] * z[c];
}

/* Riccati step of stage k < N: P_k, p_k, Kt, kv, Linv, Prb from P_{k+1}, p_{k+1}, Gamma, q tilde, rb */
static void fast_factor_stage(fast_ws* f, const workspace* w, int k, double dt)
{
    const double c2 = 0.5 * dt * dt;
    const stage_sr* r = &f->sr[k];
    double (*Pn)[NX] = f->P[k + 1];
    const double* Gam = f->Gam[k];
    static __thread double W[NX][NZ], M[NZ][NZ];
    /* W = P+ G, column by column of G */
    for (int l = 0; l < NX; l++) {
        const double* pr = Pn[l];
        for (int c = 0; c < NQ; c++) {
            W[l][IA + c] = c2 * pr[c] + dt * pr[NQ + c];
            W[l][IV + c] = dt * pr[c] + pr[NQ + c];
            double s = pr[c];
            for (int h = 0; h < 6; h++) s += pr[2 * NQ + h] * r->Hq[h][c];
            W[l][IQ + c] = s;
        }
        for (int j = 0; j < 12; j++) {
            double s = 0.0;
            for (int h = 0; h < 6; h++) s += pr[2 * NQ + h] * r->Hf[h][j];
            W[l][IFRC + j] = s;
        }
        for (int h = 0; h < 6; h++) W[l][IH + h] = pr[2 * NQ + h];
    }
    /* M = G' W (lower triangle) */
    for (int c = 0; c < NZ; c++)
        for (int d = 0; d <= c; d++) {
            double s;
            if (c < IFRC) s = c2 * W[c][d] + dt * W[NQ + c][d];
            else if (c < IQ) { s = 0.0; for (int h = 0; h < 6; h++) s += r->Hf[h][c - IFRC] * W[2 * NQ + h][d]; }
            else if (c < IV) { s = W[c - IQ][d]; for (int h = 0; h < 6; h++) s += r->Hq[h][c - IQ] * W[2 * NQ + h][d]; }
            else if (c < IH) s = dt * W[c - IV][d] + W[NQ + c - IV][d];
            else s = W[2 * NQ + c - IH][d];
            M[c][d] = s;
        }
    /* + H + C' Gamma C + reg */
    for (int c = 0; c < NU; c++) M[c][c] += r->Ru[c];
    for (int i = 0; i < NQ; i++) {
        M[IV + i][IV + i] += r->Qv[i];
        for (int j = 0; j <= i; j++) M[IQ + i][IQ + j] += r->Qqq[i][j];
    }
    for (int i = 0; i < 6; i++) M[IH + i][IH + i] += r->Qh[i];
    for (int c = 0; c < NZ; c++) M[c][c] += IPM_REG_PRIM;
    for (int j = 0; j < 20; j++) {
        const double gm = Gam[j];
        if (gm == 0.0) continue;
        const int c0 = IFRC + 3 * (j / 5);
        for (int a = 0; a < 3; a++)
            for (int b = 0; b <= a; b++) M[c0 + a][c0 + b] += gm * r->CF[j][a] * r->CF[j][b];
    }
    for (int q = 0; q < 22; q++) {
        const double gm = Gam[20 + q];
        if (gm == 0.0) continue;
        const double* cx = r->CX[q];
        for (int a = 0; a < 36; a++) {
            const double v = gm * cx[a];
            if (v == 0.0) continue;
            for (int b = 0; b <= a; b++) M[IQ + a][IQ + b] += v * cx[b];
        }
        if (q >= 16) {            /* centroidal rows: identity in h */
            const int hi = IH + q - 16;
            for (int b = 0; b < 36; b++) M[hi][IQ + b] += gm * cx[b];
            M[hi][hi] += gm;
        }
    }
    /* w = P+ rb + p+,  m = q tilde + G' w */
    double wv[NX], m[NZ];
    for (int l = 0; l < NX; l++) {
        double s = 0.0;
        for (int j = 0; j < NX; j++) s += Pn[l][j] * w->rb[k][j];
        f->Prb[k][l] = s;
        wv[l] = s + f->p[k + 1][l];
    }
    for (int c = 0; c < NZ; c++) m[c] = w->qt[k][c] + f_gt(r, wv, c, dt, c2);
    /* Cholesky of the input block, Lxu, lu (row 72) */
    double (*Li)[NU] = f->Linv[k];
    double L[NZ + 1][NU];
    for (int j = 0; j < NU; j++) {
        double d = M[j][j];
        for (int l = 0; l < j; l++) d -= L[j][l] * L[j][l];
        double inv;
        if (d > 0.0) { d = sqrt(d); inv = 1.0 / d; } else { d = 0.0; inv = 0.0; }
        L[j][j] = d;
        for (int i = j + 1; i <= NZ; i++) {
            double s = i < NZ ? M[i][j] : m[j];
            for (int l = 0; l < j; l++) s -= L[i][l] * L[j][l];
            L[i][j] = s * inv;
        }
    }
    /* Linv = Luu^-1 (lower) */
    for (int j = 0; j < NU; j++) {
        for (int i = 0; i < j; i++) Li[i][j] = 0.0;
        Li[j][j] = L[j][j] != 0.0 ? 1.0 / L[j][j] : 0.0;
        for (int i = j + 1; i < NU; i++) {
            double s = 0.0;
            for (int l = j; l < i; l++) s -= L[i][l] * Li[l][j];
            Li[i][j] = L[i][i] != 0.0 ? s / L[i][i] : 0.0;
        }
    }
    /* Kt[i][j] = sum_{l >= j} Linv[l][j] Lxu[i][l];  kv = Linv' lu */
    for (int i = 0; i < NX; i++)
        for (int j = 0; j < NU; j++) {
            double s = 0.0;
            for (int l = j; l < NU; l++) s += Li[l][j] * L[NU + i][l];
            f->Kt[k][i][j] = s;
        }
    for (int j = 0; j < NU; j++) {
        double s = 0.0;
        for (int l = j; l < NU; l++) s += Li[l][j] * L[NZ][l];
        f->kv[k][j] = s;
    }
    /* P = Mxx - Lxu Lxu',  p = m_x - Lxu lu */
    for (int i = 0; i < NX; i++) {
        for (int j = 0; j <= i; j++) {
            double s = M[NU + i][NU + j];
            for (int l = 0; l < NU; l++) s -= L[NU + i][l] * L[NU + j][l];
            f->P[k][i][j] = s; f->P[k][j][i] = s;
        }
        double s = m[NU + i];
        for (int l = 0; l < NU; l++) s -= L[NU + i][l] * L[NZ][l];
        f->p[k][i] = s;
    }
}

/* terminal stage: P_N = Hxx + (C' Gamma C)xx, p_N = (q tilde)_x */
static void fast_factor_terminal(fast_ws* f, const workspace* w, int N)
{
    const stage_sr* r = &f->sr[N];
    const double* Gam = f->Gam[N];
    double (*P)[NX] = f->P[N];
    memset(P, 0, sizeof(double) * NX * NX);
    for (int i = 0; i < NQ; i++) {
        P[NQ + i][NQ + i] = r->Qv[i];
        for (int j = 0; j < NQ; j++) P[i][j] = r->Qqq[i][j];
    }
    for (int i = 0; i < 6; i++) P[2 * NQ + i][2 * NQ + i] = r->Qh[i];
    for (int q = 0; q < 22; q++) {
        const double gm = Gam[20 + q];
        if (gm == 0.0) continue;
        const double* cx = r->CX[q];
        for (int a = 0; a < 36; a++)
            for (int b = 0; b < 36; b++) P[a][b] += gm * cx[a] * cx[b];
        if (q >= 16) {
            const int hi = 2 * NQ + q - 16;
            for (int b = 0; b < 36; b++) { P[hi][b] += gm * cx[b]; P[b][hi] += gm * cx[b]; }
            P[hi][hi] += gm;
        }
    }
    for (int i = 0; i < NX; i++) f->p[N][i] = w->qt[N][NU + i];
}

/* vector-only backward sweep on the stored gains (corrector): p_k, kv_k from q tilde */
static void fast_backward_vectors(fast_ws* f, const workspace* w, int N, double dt)
{
    const double c2 = 0.5 * dt * dt;
    for (int i = 0; i < NX; i++) f->p[N][i] = w->qt[N][NU + i];
    for (int k = N - 1; k >= 0; k--) {
        const stage_sr* r = &f->sr[k];
        double wv[NX], m[NZ], t[NU];
        for (int l = 0; l < NX; l++) wv[l] = f->Prb[k][l] + f->p[k + 1][l];
        for (int c = 0; c < NZ; c++) m[c] = w->qt[k][c] + f_gt(r, wv, c, dt, c2);
        for (int i = 0; i < NU; i++) {
            double s = 0.0;
            for (int l = 0; l <= i; l++) s += f->Linv[k][i][l] * m[l];
            t[i] = s;
        }
        for (int j = 0; j < NU; j++) {
            double s = 0.0;
            for (int l = j; l < NU; l++) s += f->Linv[k][l][j] * t[l];
            f->kv[k][j] = s;
        }
        for (int i = 0; i < NX; i++) {
            double s = m[NU + i];
            for (int j = 0; j < NU; j++) s -= f->Kt[k][i][j] * m[j];
            f->p[k][i] = s;
        }
    }
}

/* forward sweep: dz; with dpi != NULL also dpi_{k-1} = P_k dx_k + p_k */
static void fast_forward(const fast_ws* f, workspace* w, int N, double dt, int with_dpi)
{
    const double c2 = 0.5 * dt * dt;
    double dx[NX] = {0};
    for (int k = 0; k <= N; k++) {
        double* dz = w->dz[k];
        if (k < N) {
            for (int j = 0; j < NU; j++) {
                double s = f->kv[k][j];
                for (int i = 0; i < NX; i++) s += f->Kt[k][i][j] * dx[i];
                dz[j] = -s;
            }
        } else for (int j = 0; j < NU; j++) dz[j] = 0.0;
        for (int i = 0; i < NX; i++) dz[NU + i] = dx[i];
        if (with_dpi && k > 0)
            for (int i = 0; i < NX; i++) {
                double s = f->p[k][i];
                for (int j = 0; j < NX; j++) s += f->P[k][i][j] * dx[j];
                w->dpi[k - 1][i] = s;
            }
        if (k < N) {
            double dxn[NX];
            for (int l = 0; l < NX; l++) dxn[l] = f_g(&f->sr[k], dz, l, dt, c2) + w->rb[k][l];
            memcpy(dx, dxn, sizeof(dx));
        }
    }
}

/* slack / multiplier directions of the current dz, step-length candidates, sum (lam + a dlam)(t + a dt) coefficients */
static void fast_slack_step(const fast_ws* f, workspace* w, int N, double (*lam)[2 * NG], double (*t)[2 * NG],
                            double (*dtv_out)[2 * NG], double (*dl_out)[2 * NG], double* a_p, double* a_d, double cs[3])
{
    *a_p = 1.0; *a_d = 1.0; cs[0] = cs[1] = cs[2] = 0.0;
    for (int k = 0; k <= N; k++) {
        const stage_sr* r = &f->sr[k];
        for (int j = 0; j < NG; j++) {
            const double cdz = f_cz(r, w->dz[k], j);
            for (int e = 0; e < 2; e++) {
                const int id = e * NG + j;
                const double msk = e ? r->mu[j] : r->ml[j];
                if (msk < 0.5) { dtv_out[k][id] = 0; dl_out[k][id] = 0; continue; }
                const double dtv = (e ? -cdz : cdz) + w->rd[k][id];
                const double dl = -(w->rm[k][id] + lam[k][id] * dtv) / t[k][id];
                dtv_out[k][id] = dtv; dl_out[k][id] = dl;
                if (dtv < 0 && -t[k][id] / dtv < *a_p) *a_p = -t[k][id] / dtv;
                if (dl < 0 && -lam[k][id] / dl < *a_d) *a_d = -lam[k][id] / dl;
                cs[0] += lam[k][id] * t[k][id]; cs[1] += lam[k][id] * dtv + t[k][id] * dl; cs[2] += dl * dtv;
            }
        }
    }
}

static void qp_ipm_fast(workspace* w, fast_ws* f, const nmpc_opts* o, int N, const double* dx0, int warm,
                        double (*z)[NZ], double (*pi)[NX], double (*lam)[2 * NG], double (*t)[2 * NG],
                        qp_info* info)
{
    const double tol = o->qp_tol, dt = o->dt_nodes, c2 = 0.5 * dt * dt;
    if (warm == 0) memset(z, 0, sizeof(double) * (N + 1) * NZ);
    if (warm < 2) memset(pi, 0, sizeof(double) * N * NX);
    for (int i = 0; i < NX; i++) z[0][NU + i] = dx0[i];
    for (int i = 0; i < NU; i++) z[N][i] = 0.0;
    int nc = 0;
    for (int k = 0; k <= N; k++) {
        const stage_sr* r = &f->sr[k];
        for (int j = 0; j < NG; j++) {
            const double cz = f_cz(r, z[k], j);
            for (int e = 0; e < 2; e++) {
                const int id = e * NG + j;
                const double msk = e ? r->mu[j] : r->ml[j];
                if (msk > 0.5) {
                    nc++;
                    if (warm < 2 || !(t[k][id] > 0.0) || !(lam[k][id] > 0.0)) {
                        const double sl = e ? r->ug[j] - cz : cz - r->lg[j];
                        t[k][id] = sl > IPM_THR0 ? sl : IPM_THR0; lam[k][id] = IPM_MU0 / t[k][id];
                    } else { if (t[k][id] < IPM_THR0) t[k][id] = IPM_THR0; if (lam[k][id] < IPM_THR0) lam[k][id] = IPM_THR0; }
                } else { t[k][id] = 1.0; lam[k][id] = 0.0; }
            }
        }
    }
    const double inv_nc = nc > 0 ? 1.0 / nc : 0.0;
    info->status = 1; info->ncond = 0;
    int it = 0;
    for (;; it++) {
        /* ---- residuals, Gamma, q tilde (stage-parallel) ---- */
        double res_g = 0, res_b = 0, res_d = 0, res_m = 0, mu = 0;
        for (int k = 0; k <= N; k++) {
            const stage_sr* r = &f->sr[k];
            double dl[NG], gam[NG];
            for (int j = 0; j < NG; j++) {
                const double cz = f_cz(r, z[k], j);
                w->rd[k][j] = r->ml[j] * (cz - t[k][j] - r->lg[j]);
                w->rd[k][NG + j] = r->mu[j] * (-cz - t[k][NG + j] + r->ug[j]);
                w->rm[k][j] = r->ml[j] * lam[k][j] * t[k][j];
                w->rm[k][NG + j] = r->mu[j] * lam[k][NG + j] * t[k][NG + j];
                mu += w->rm[k][j] + w->rm[k][NG + j];
                for (int e = 0; e < 2; e++) { NORM_ACC(res_d, w->rd[k][e * NG + j]); NORM_ACC(res_m, w->rm[k][e * NG + j]); }
                dl[j] = r->ml[j] * lam[k][j] - r->mu[j] * lam[k][NG + j];
                f->Gam[k][j] = r->ml[j] * lam[k][j] / t[k][j] + r->mu[j] * lam[k][NG + j] / t[k][NG + j];
                gam[j] = r->ml[j] * (w->rm[k][j] + lam[k][j] * w->rd[k][j]) / t[k][j]
                       - r->mu[j] * (w->rm[k][NG + j] + lam[k][NG + j] * w->rd[k][NG + j]) / t[k][NG + j];
            }
            for (int c = 0; c < NZ; c++) {
                double acc = r->g[c] + f_h(r, z[k], c);
                if (k < N) acc += f_gt(r, pi[k], c, dt, c2);
                if (k > 0 && c >= NU) acc -= pi[k - 1][c - NU];
                acc -= f_ct(r, dl, c);
                if ((k == 0 && c >= NU) || (k == N && c < NU)) acc = 0.0;
                w->rg[k][c] = acc;
                NORM_ACC(res_g, acc);
                w->qt[k][c] = acc + f_ct(r, gam, c);
            }
            if (k < N) for (int i = 0; i < NX; i++) {
                const double acc = f_g(r, z[k], i, dt, c2) + r->b[i] - z[k + 1][NU + i];
                w->rb[k][i] = acc;
                NORM_ACC(res_b, acc);
            }
        }
        mu *= inv_nc;
        info->res[0] = res_g; info->res[1] = res_b; info->res[2] = res_d; info->res[3] = res_m; info->mu = mu;
        if (!(res_g == res_g) || !(res_b == res_b) || !(res_d == res_d) || !(res_m == res_m)) { info->status = 3; break; }
        if (res_g < tol && res_b < tol && res_d < tol && res_m < tol) { info->status = 0; break; }
        if (it >= o->max_qp_iter) { info->status = 1; break; }

        /* ---- factorisation + affine direction ---- */
        fast_factor_terminal(f, w, N);
        for (int k = N - 1; k >= 0; k--) fast_factor_stage(f, w, k, dt);
        fast_forward(f, w, N, dt, 0);
        double a_p, a_d, cs[3];
        fast_slack_step(f, w, N, lam, t, w->dt_a, w->dlam_a, &a_p, &a_d, cs);
        const double alpha_aff = a_p < a_d ? a_p : a_d;
        const double mu_aff = (cs[0] + alpha_aff * (cs[1] + alpha_aff * cs[2])) * inv_nc;
        double sigma = mu > 0 ? mu_aff / mu : 0.0;
        sigma = sigma * sigma * sigma;
        double sigma_mu = sigma * mu;
        if (sigma_mu < IPM_TAU_MIN) sigma_mu = IPM_TAU_MIN;
        /* ---- corrector (pass 1: centring only, if the conditional predictor-corrector rejects pass 0) ---- */
        for (int pass = 0; pass < 2; pass++) {
            for (int k = 0; k <= N; k++) {
                const stage_sr* r = &f->sr[k];
                double gam[NG];
                for (int j = 0; j < NG; j++) {
                    for (int e = 0; e < 2; e++) {
                        const int id = e * NG + j;
                        const double msk = e ? r->mu[j] : r->ml[j];
                        if (msk > 0.5)
                            w->rm[k][id] = lam[k][id] * t[k][id] + (pass == 0 ? w->dt_a[k][id] * w->dlam_a[k][id] : 0.0) - sigma_mu;
                    }
                    gam[j] = r->ml[j] * (w->rm[k][j] + lam[k][j] * w->rd[k][j]) / t[k][j]
                           - r->mu[j] * (w->rm[k][NG + j] + lam[k][NG + j] * w->rd[k][NG + j]) / t[k][NG + j];
                }
                for (int c = 0; c < NZ; c++) w->qt[k][c] = w->rg[k][c] + f_ct(r, gam, c);
            }
            fast_backward_vectors(f, w, N, dt);
            fast_forward(f, w, N, dt, 1);
            fast_slack_step(f, w, N, lam, t, w->dt, w->dlam, &a_p, &a_d, cs);
            if (pass == 1 || !o->cond_pred_corr) break;
            const double al = a_p < a_d ? a_p : a_d;
            const double mu_c = (cs[0] + al * (cs[1] + al * cs[2])) * inv_nc;
            if (!(mu_c > 2.0 * mu_aff)) break;
            info->ncond++;
        }
        double alpha_p = a_p, alpha_d = a_d;
        alpha_p *= (1.0 - alpha_p) * 0.99 + alpha_p * 0.9999999;
        alpha_d *= (1.0 - alpha_d) * 0.99 + alpha_d * 0.9999999;
        if (alpha_p < IPM_ALPHA_MIN && alpha_d < IPM_ALPHA_MIN) { info->status = 2; break; }
        for (int k = 0; k <= N; k++) {
            const stage_sr* r = &f->sr[k];
            for (int c = 0; c < NZ; c++) z[k][c] += alpha_p * w->dz[k][c];
            if (k < N) for (int i = 0; i < NX; i++) pi[k][i] += alpha_d * w->dpi[k][i];
            for (int id = 0; id < 2 * NG; id++) {
                const double msk = id < NG ? r->ml[id] : r->mu[id - NG];
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
