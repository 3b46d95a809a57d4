"""Dense (numpy) view of an exported OCP-QP, for algorithm-independent checks of the Riccati IPM.

Layouts are those of include/nmpc_b200.h (`nmpc_export_qp`):  z_k = [u_k; x_k],
min sum_k 1/2 z_k' H_k z_k + g_k' z_k   s.t.  x_{k+1} = A_k x_k + B_k u_k + b_k,  x_0 = dx0,
lg_k <= C_k z_k (rows with mask_l) , C_k z_k <= ug_k (rows with mask_u); u_N is absent.
"""
import numpy as np

NX, NU, NZ, NG = 42, 30, 72, 42


def unpack(qp, b):
    N = qp["BAbt"].shape[1]
    st = []
    for k in range(N + 1):
        R = qp["RSQrq"][b, k]
        H = np.tril(R[:NZ]) + np.tril(R[:NZ], -1).T
        g = R[NZ].copy()
        C = qp["DCt"][b, k].T.copy()
        d = dict(H=H, g=g, C=C, lg=qp["lg"][b, k], ug=qp["ug"][b, k],
                 ml=qp["mask"][b, k, :NG] > 0.5, mu=qp["mask"][b, k, NG:] > 0.5)
        if k < N:
            BA = qp["BAbt"][b, k]
            d["G"] = BA[:NZ].T.copy()     # [B A]  (NX x NZ)
            d["b"] = BA[NZ].copy()
        st.append(d)
    return st


def kkt_residuals(st, dx0, z, pi, lam):
    """inf-norms of stationarity, dynamics, inequality violation, complementarity of (z, pi, lam)
    computed densely; the free variables are u_0..u_{N-1}, x_1..x_N (x_0 fixed, u_N absent)."""
    N = len(st) - 1
    rs = re = ri = rc = 0.0
    re = max(re, np.abs(z[0, NU:] - dx0).max())
    for k, s in enumerate(st):
        r = s["H"] @ z[k] + s["g"]
        if k < N:
            r += s["G"].T @ pi[k]
        if k > 0:
            r[NU:] -= pi[k - 1]
        ll = np.where(s["ml"], lam[k, :NG], 0.0)
        lu = np.where(s["mu"], lam[k, NG:], 0.0)
        r -= s["C"].T @ (ll - lu)
        if k == 0:
            r = r[:NU]
        if k == N:
            r = r[NU:]
        rs = max(rs, np.abs(r).max())
        if k < N:
            re = max(re, np.abs(s["G"] @ z[k] + s["b"] - z[k + 1, NU:]).max())
        cz = s["C"] @ z[k]
        sl, su = cz - s["lg"], s["ug"] - cz
        if s["ml"].any():
            ri = max(ri, (-sl[s["ml"]]).max())
            rc = max(rc, np.abs(ll * sl)[s["ml"]].max())
        if s["mu"].any():
            ri = max(ri, (-su[s["mu"]]).max())
            rc = max(rc, np.abs(lu * su)[s["mu"]].max())
    return rs, re, max(ri, 0.0), rc


def solve_equality_qp(st, dx0):
    """Dense KKT solve of the QP with the inequality rows dropped (one Newton step of the
    barrier problem with Gamma = 0): returns z [N+1,NZ], pi [N,NX]."""
    N = len(st) - 1
    nz = (N + 1) * NZ
    ne = N * NX + NX + NU
    K = np.zeros((nz + ne, nz + ne))
    rhs = np.zeros(nz + ne)
    for k, s in enumerate(st):
        K[k * NZ:(k + 1) * NZ, k * NZ:(k + 1) * NZ] = s["H"]
        rhs[k * NZ:(k + 1) * NZ] = -s["g"]
    row = nz
    for k in range(N):                  # G z_k - x_{k+1} = -b
        s = st[k]
        K[row:row + NX, k * NZ:(k + 1) * NZ] = s["G"]
        K[row:row + NX, (k + 1) * NZ + NU:(k + 2) * NZ] = -np.eye(NX)
        rhs[row:row + NX] = -s["b"]
        row += NX
    K[row:row + NX, NU:NZ] = np.eye(NX)          # x_0 = dx0
    rhs[row:row + NX] = dx0
    row += NX
    K[row:row + NU, N * NZ:N * NZ + NU] = np.eye(NU)   # u_N = 0
    K[nz:, :nz] = K[nz:, :nz]
    K[:nz, nz:] = K[nz:, :nz].T
    sol = np.linalg.solve(K, rhs)
    z = sol[:nz].reshape(N + 1, NZ)
    pi = sol[nz:nz + N * NX].reshape(N, NX)
    return z, pi
