"""Prototype / validation of the quadrature used by csrc/stieltjes.cu (numpy, CPU).

m_tilde_0 = 0.5 (x + 2I - sqrt(x^2 + 4x)) = g(x) with g(xi) = (1/2pi) int_0^4 sqrt(u(4-u)) / (xi + u) du.
Trapezoid rule after u = 4 sigmoid(s); resolvents of x = E + U U' by Woodbury.  Run: python tools/proto_stieltjes.py
"""
import numpy as np


def nodes(h=0.5, smin=-36.0, smax=24.0):
    s = np.arange(smin, smax + 1e-9, h)
    sg = 1 / (1 + np.exp(-s)); sgm = 1 / (1 + np.exp(s))
    return 4 * sg, h * (8 / np.pi) * (sg * sgm) ** 1.5


def g_exact_stable(xi):
    return 2 / (xi + 2 + np.sqrt(xi * xi + 4 * xi))


def mt0_woodbury(E, Ut, u, w):
    """diag(sum_j w_j D_j) - G G' with G_j = sqrt(w_j) D_j U~ L_j^-T, T_j = I + U~' D_j U~ = L_j L_j'."""
    import scipy.linalg as sl
    n, K = Ut.shape
    dd = np.zeros(n); G = []
    for uj, wj in zip(u, w):
        d = 1 / (E + uj); dd += wj * d
        Lc = np.linalg.cholesky(np.eye(K) + Ut.T @ (d[:, None] * Ut))
        G.append(np.sqrt(wj) * sl.solve_triangular(Lc, (d[:, None] * Ut).T, lower=True).T)
    G = np.concatenate(G, axis=1)
    return np.diag(dd) - G @ G.T


if __name__ == "__main__":
    xi = np.logspace(-8, 3, 3000)
    for h, a, b in ((0.6, -36, 24), (0.5, -36, 24), (0.5, -32, 22), (0.4, -36, 24)):
        u, w = nodes(h, a, b)
        err = np.abs((w[None, :] / (xi[:, None] + u[None, :])).sum(1) - g_exact_stable(xi)).max()
        print(f"h={h} s in [{a},{b}] nodes={len(u)} max abs err on [1e-8,1e3] = {err:.2e}")
    rng = np.random.default_rng(0)
    n, K = 300, 25
    E = 10 ** rng.uniform(-5, -1, n); Ut = rng.standard_normal((n, K)) * 0.05
    x = np.diag(E) + Ut @ Ut.T
    ev, Q = np.linalg.eigh(x)
    u, w = nodes()
    print("woodbury + quadrature vs eigh:", np.abs(mt0_woodbury(E, Ut, u, w) - (Q * g_exact_stable(ev)) @ Q.T).max())
