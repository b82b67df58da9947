"""numpy restatement of the p-multigrid preconditioner (yaourt-fem-dg_b200/csrc/yfd_pmg.cu): test infrastructure.

Levels = the degrees k, k-1, ..., 1 as INDEX SUBSETS of the hierarchical scaled-monomial basis (core/bases.hpp:96-122),
level operators = sub-blocks of A, smoother = Chebyshev iteration on D^-1 A (D = the level's block diagonal), coarsest
level = a fixed Chebyshev polynomial from zero; one symmetric V-cycle per application."""
import numpy as np
import scipy.sparse as sp


def bs_of(k):
    return (k + 2) * (k + 1) // 2


def cheb_coeffs(lmin, lmax, count):
    theta, delta = 0.5 * (lmax + lmin), 0.5 * (lmax - lmin)
    sigma = theta / delta
    rho = 1.0 / sigma
    out = [(0.0, 1.0 / theta)]
    for _ in range(1, count):
        rho_new = 1.0 / (2.0 * sigma - rho)
        out.append((rho_new * rho, 2.0 * rho_new / delta))
        rho = rho_new
    return out


class PmgNumpy:
    def __init__(self, A, ncells, k, nu, coarse_degree, smooth_ratio, lambda_max):
        """lambda_max: the bounds per level, finest first (take them from the device's yfd_system_pmg_info so both sides
        run the same polynomials)."""
        self.k, self.ncells, self.nu, self.mc, self.ratio = k, ncells, nu, coarse_degree, smooth_ratio
        bs = bs_of(k)
        self.levels = []
        for l, deg in enumerate(range(k, 0, -1)):
            bl = bs_of(deg)
            idx = (np.arange(ncells)[:, None] * bs + np.arange(bl)[None, :]).ravel()
            Al = A[idx][:, idx].tocsr()
            D = np.stack([Al[c * bl:(c + 1) * bl, c * bl:(c + 1) * bl].toarray() for c in range(ncells)])
            self.levels.append(dict(A=Al, Dinv=np.linalg.inv(D), bs=bl, lmax=lambda_max[l]))

    def _dinv(self, L, v):
        return np.einsum("cij,cj->ci", L["Dinv"], v.reshape(-1, L["bs"])).ravel()

    def _sweeps(self, L, b, x, count, lmin, lmax):
        d = np.zeros_like(b)
        for c1, c2 in cheb_coeffs(lmin, lmax, count):
            r = b - L["A"] @ x if x is not None else b
            d = c1 * d + c2 * self._dinv(L, r)
            x = d.copy() if x is None else x + d
        return x

    def _vcycle(self, l, b):
        L = self.levels[l]
        if l == len(self.levels) - 1:
            return self._sweeps(L, b, None, self.mc, L["lmax"] / self.mc ** 2, L["lmax"])
        lmin = L["lmax"] / self.ratio
        x = self._sweeps(L, b, None, self.nu, lmin, L["lmax"])
        r = b - L["A"] @ x
        bc = self.levels[l + 1]["bs"]
        e = self._vcycle(l + 1, r.reshape(-1, L["bs"])[:, :bc].ravel().copy())
        x = x.reshape(-1, L["bs"]).copy()
        x[:, :bc] += e.reshape(-1, bc)
        return self._sweeps(L, b, x.ravel(), self.nu, lmin, L["lmax"])

    def __call__(self, r):
        return self._vcycle(0, r)


def lambda_max_estimate(A, ncells, k, its=30):
    """1.1 x power-iteration estimates of lambda_max(D^-1 A_l), finest first (what ensure_pmg computes on the device)."""
    out = []
    bs = bs_of(k)
    rng = np.random.default_rng(0)
    for deg in range(k, 0, -1):
        bl = bs_of(deg)
        idx = (np.arange(ncells)[:, None] * bs + np.arange(bl)[None, :]).ravel()
        Al = A[idx][:, idx].tocsr()
        Dinv = np.linalg.inv(np.stack([Al[c * bl:(c + 1) * bl, c * bl:(c + 1) * bl].toarray() for c in range(ncells)]))
        v = rng.standard_normal(Al.shape[0])
        lam = 1.0
        for _ in range(its):
            w = np.einsum("cij,cj->ci", Dinv, (Al @ v).reshape(-1, bl)).ravel()
            lam = np.linalg.norm(w) / np.linalg.norm(v)
            v = w / np.linalg.norm(w)
        out.append(1.1 * lam)
    return out


def pcg(A, b, apply_m, rr_tol, max_iter):
    """the reference's preconditioned recurrences (core/solvers.hpp:196-227)"""
    x = np.zeros_like(b); r = b.copy(); z = apply_m(r); d = z.copy()
    nr0 = nr = np.linalg.norm(r); it = 0; hist = [1.0]
    while nr / nr0 > rr_tol and it < max_iter:
        y = A @ d
        rho = r @ z
        alpha = rho / (d @ y)
        x += alpha * d; r -= alpha * y
        z = apply_m(r)
        d = z + (r @ z) / rho * d
        nr = np.linalg.norm(r); it += 1; hist.append(nr / nr0)
    return x, it, np.array(hist)
