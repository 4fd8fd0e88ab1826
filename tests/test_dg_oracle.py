"""Pins the DG oracle (oracle/dg_oracle.cpp, restating laplace_operator_dg_face.h) by properties, since the
reference holds no golden vectors for it: symmetry and definiteness of the SIP form, consistency
a(u,v) = -int Laplace(u) v for global polynomials on the sheared mesh, and the jump terms on piecewise
constants."""
import numpy as np
import pytest


@pytest.mark.parametrize("p,ncell", [(1, (3, 2, 2)), (2, (2, 3, 2)), (4, (2, 2, 3)), (3, (3, 3, 3))])
def test_symmetric_positive_definite(orc, p, ncell):
    A = orc.DGOperator(p, ncell)
    rng = np.random.default_rng(1)
    n = A.n_dofs()
    x, y = rng.random(n) - 0.5, rng.random(n) - 0.5
    Ax, Ay = A.vmult(x), A.vmult(y)
    assert abs(y @ Ax - x @ Ay) <= 1e-12 * abs(y @ Ax)
    assert x @ Ax > 0 and y @ Ay > 0
    assert np.allclose(A.vmult(2 * x - 3 * y), 2 * Ax - 3 * Ay, rtol=1e-12, atol=1e-12)


def cell_integrals(orc, A):
    """int phi_i over a cell (cell-local index i), by the tensor Gauss rule"""
    b = orc.basis(A.degree)
    m1 = b["S"].T @ b["wq"]
    h = [(A.upper[d] - A.lower[d]) / A.ncell[d] for d in range(3)]
    detJ = abs(np.linalg.det(A.trafo)) * h[0] * h[1] * h[2]
    return detJ * np.einsum("k,j,i->kji", m1, m1, m1).ravel()


def interior_mask(A):
    n3 = (A.degree + 1) ** 3
    nc = A.ncell
    c = np.arange(nc[0] * nc[1] * nc[2])
    cx, cy, cz = c % nc[0], (c // nc[0]) % nc[1], c // (nc[0] * nc[1])
    inner = (cx > 0) & (cx < nc[0] - 1) & (cy > 0) & (cy < nc[1] - 1) & (cz > 0) & (cz < nc[2] - 1)
    return np.repeat(inner, n3)


@pytest.mark.parametrize("p", [2, 3, 4])
def test_consistency_on_global_polynomials(orc, p):
    """continuous u: jumps vanish, a(u, phi_i) = -int Laplace(u) phi_i for test functions away from the boundary"""
    A = orc.DGOperator(p, (3, 3, 3))
    X = A.support_points()
    x, y, z = X[:, 0], X[:, 1], X[:, 2]
    inner = interior_mask(A)
    mi = np.tile(cell_integrals(orc, A), 27)
    scale = np.abs(A.vmult(x * x)).max()
    # harmonic: A u = 0 in the interior
    for u in (1 + 0 * x, x - 2 * y + 0.5 * z, x * x - y * y + 3 * x * z, x * y + y * z):
        r = A.vmult(u)
        assert np.abs(r[inner]).max() <= 1e-11 * scale
    # Laplace(u) = const: A u = -const * int phi_i
    for u, lap in ((x * x, 2.0), (x * x + 2 * y * y - z * z + x * y, 4.0)):
        r = A.vmult(u)
        assert np.abs(r[inner] + lap * mi[inner]).max() <= 1e-11 * scale


@pytest.mark.parametrize("p", [1, 3])
def test_jump_terms_on_piecewise_constants(orc, p):
    """u constant per cell: (A u)_i = sum_f jump_f (sigma_f int_f phi_i - 1/2 int_f dn phi_i), jump = 2 u at the boundary"""
    nc = (3, 2, 2)
    A = orc.DGOperator(p, nc)
    b = orc.basis(p)
    n = p + 1
    rng = np.random.default_rng(5)
    uc = rng.random(nc[::-1])  # [cz, cy, cx]
    u = np.repeat(uc.ravel(), n ** 3)
    r = A.vmult(u).reshape(nc[2], nc[1], nc[0], n, n, n)
    h = np.array([(A.upper[d] - A.lower[d]) / nc[d] for d in range(3)])
    J = A.trafo * h[None, :]
    Jinv = np.linalg.inv(J)
    detJ = abs(np.linalg.det(J))
    gn = np.linalg.norm(Jinv, axis=1)
    sigma = A.penalty()
    assert np.allclose(sigma, (p + 1) ** 2 * gn)
    m1 = b["S"].T @ b["wq"]       # int phi_i
    g1 = b["G"].T @ b["wq"]       # int phi_i'
    nodes = b["gll"]

    def lagr_d(i, x):             # phi_i'(x)
        s = 0.0
        for m in range(n):
            if m == i:
                continue
            v = 1.0 / (nodes[i] - nodes[m])
            for l in range(n):
                if l != i and l != m:
                    v *= (x - nodes[l]) / (nodes[i] - nodes[l])
            s += v
        return s
    end_v = {0: np.eye(n)[0], 1: np.eye(n)[p]}
    end_d = {0: np.array([lagr_d(i, 0.0) for i in range(n)]), 1: np.array([lagr_d(i, 1.0) for i in range(n)])}
    expect = np.zeros_like(r)
    for cz in range(nc[2]):
        for cy in range(nc[1]):
            for cx in range(nc[0]):
                c = [cx, cy, cz]
                for d in range(3):
                    for side in (0, 1):
                        cn = list(c)
                        cn[d] += 1 if side else -1
                        jump = 2 * uc[cz, cy, cx] if not (0 <= cn[d] < nc[d]) else uc[cz, cy, cx] - uc[cn[2], cn[1], cn[0]]
                        sgn = 1.0 if side else -1.0
                        # 1-D factors per direction for int_f phi_i and int_f (grad xi_e . n) d_e phi_i
                        val = [end_v[side] if e == d else m1 for e in range(3)]
                        iv = np.einsum("k,j,i->kji", val[2], val[1], val[0])
                        dn = np.zeros((n, n, n))
                        for e in range(3):
                            f = [(end_d[side] if e == d else end_v[side]) if t == d else (g1 if t == e else m1) for t in range(3)]
                            coef = sgn * (Jinv[e] @ Jinv[d]) / gn[d]
                            dn += coef * np.einsum("k,j,i->kji", f[2], f[1], f[0])
                        area = detJ * gn[d]
                        expect[cz, cy, cx] += jump * area * (sigma[d] * iv - 0.5 * dn)
    assert np.abs(r - expect).max() <= 1e-12 * np.abs(expect).max()


def test_diagonal_probe(orc):
    A = orc.DGOperator(2, (2, 2, 2))
    d = A.diagonal()
    rng = np.random.default_rng(0)
    for i in rng.integers(0, A.n_dofs(), 5):
        e = np.zeros(A.n_dofs())
        e[i] = 1
        assert abs(A.vmult(e)[i] - d[i]) <= 1e-13 * abs(d[i])
    assert d.min() > 0
