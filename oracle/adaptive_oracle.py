"""CPU oracle (test infrastructure only) for the FE_Q(p) Laplacian on adaptively refined box meshes with hanging
nodes -- SURVEY row H, the operator of poisson_l/program.cc:232-260 (deal.II AffineConstraints from
DoFTools::make_hanging_node_constraints + zero boundary values, applied by read_dof_values /
distribute_local_to_global).  deal.II is not available here and the reference holds no golden vectors for this
path: **parity unpinned** against the reference; the restatement is pinned by properties (tests/test_adaptive_oracle.py).

Written independently of the library's integer-lattice numbering: everything is geometric.  Input is the list of
active cells as boxes.  A node of a cell is constrained when it lies in the closure of an active cell twice as
large; its value is that cell's polynomial at the point (Lagrange evaluation), chains are resolved by powers of the
constraint matrix, and the global matrix is assembled as sum_K C_K^T K_K C_K with scipy.sparse."""
import numpy as np
import scipy.sparse as sp

from . import orc


def lagrange(nodes, x):
    """values of all Lagrange polynomials on `nodes` at the points x: shape (len(x), len(nodes))"""
    x = np.atleast_1d(x)
    out = np.ones((x.size, nodes.size))
    for i in range(nodes.size):
        for l in range(nodes.size):
            if l != i:
                out[:, i] *= (x - nodes[l]) / (nodes[i] - nodes[l])
    return out


class AdaptiveOracle:
    def __init__(self, degree, cell_lower, cell_size):
        self.p, self.n = degree, degree + 1
        self.lo = np.asarray(cell_lower, dtype=np.float64)
        self.sz = np.asarray(cell_size, dtype=np.float64)
        b = orc.basis(degree)
        self.gll, self.M1, self.K1 = b["gll"], b["S"].T @ np.diag(b["wq"]) @ b["S"], b["G"].T @ np.diag(b["wq"]) @ b["G"]
        self.scale = self.sz.min()
        self._build()

    def _key(self, pts):
        return np.round(pts / (self.scale * 1e-6)).astype(np.int64)

    def _build(self):
        n, nc = self.n, self.lo.shape[0]
        g = self.gll
        k, j, i = np.meshgrid(np.arange(n), np.arange(n), np.arange(n), indexing="ij")
        ref = np.stack([g[i], g[j], g[k]], axis=-1).reshape(-1, 3)                 # local node l = (k n + j) n + i
        X = self.lo[:, None, :] + self.sz[:, None, :] * ref[None, :, :]            # (cells, n^3, 3)
        smin = self.sz[:, 0].min()
        level = np.round(np.log2(self.sz[:, 0].max() / self.sz[:, 0])).astype(int)  # 0 = coarsest present
        # generalised nodes: (point, level of the owning cell)
        node_id, node_pt, node_lv = {}, [], []
        cell_nodes = np.zeros((nc, n ** 3), dtype=np.int64)
        keys = self._key(X)
        for c in range(nc):
            for l in range(n ** 3):
                kk = (int(keys[c, l, 0]), int(keys[c, l, 1]), int(keys[c, l, 2]), int(level[c]))
                if kk not in node_id:
                    node_id[kk] = len(node_pt)
                    node_pt.append(X[c, l])
                    node_lv.append(level[c])
                cell_nodes[c, l] = node_id[kk]
        P = np.array(node_pt)
        lv = np.array(node_lv)
        nn = P.shape[0]
        # constraint matrix on the generalised nodes
        rows, cols, vals = [], [], []
        slave = np.zeros(nn, dtype=bool)
        tol = 1e-9 * self.scale
        for c in range(nc):
            cand = np.where((lv == level[c] + 1) & ~slave)[0]      # nodes of cells half as large
            if cand.size == 0:
                continue
            inside = np.all((P[cand] >= self.lo[c] - tol) & (P[cand] <= self.lo[c] + self.sz[c] + tol), axis=1)
            for t in cand[inside]:
                xi = (P[t] - self.lo[c]) / self.sz[c]
                w = [lagrange(g, xi[d])[0] for d in range(3)]
                W = np.einsum("c,b,a->cba", w[2], w[1], w[0]).ravel()
                nz = np.abs(W) > 1e-13
                rows += [t] * int(nz.sum())
                cols += list(cell_nodes[c, nz])
                vals += list(W[nz])
                slave[t] = True
        reg = np.where(~slave)[0]
        C = sp.csr_matrix((vals + [1.0] * reg.size, (rows + list(reg), cols + list(reg))), shape=(nn, nn))
        Ct = C.copy()
        for _ in range(20):                                         # resolve chains: masters that are constrained themselves
            if abs(Ct[:, np.where(slave)[0]]).sum() < 1e-14:
                break
            Ct = Ct @ C
        else:
            raise RuntimeError("constraint chains did not resolve")
        self.regular = reg
        self.points = P[reg]
        col_of = -np.ones(nn, dtype=np.int64)
        col_of[reg] = np.arange(reg.size)
        self.C = Ct[:, reg].tocsr()                                  # generalised nodes -> DoFs
        self.cell_nodes = cell_nodes
        # Dirichlet: a DoF is on the boundary if one of the 8 octants around its point leaves the domain
        eps = 1e-4 * smin
        bnd = np.zeros(reg.size, dtype=bool)
        for sx in (-eps, eps):
            for sy in (-eps, eps):
                for sz in (-eps, eps):
                    Q = self.points + np.array([sx, sy, sz])
                    inside = np.zeros(reg.size, dtype=bool)
                    for c in range(nc):
                        inside |= np.all((Q > self.lo[c]) & (Q < self.lo[c] + self.sz[c]), axis=1)
                    bnd |= ~inside
        self.dirichlet = bnd
        # assembly
        K = sp.csr_matrix((reg.size, reg.size))
        M1, K1 = self.M1, self.K1
        blocks = []
        for c in range(nc):
            h = self.sz[c]
            vol = h[0] * h[1] * h[2]
            Kc = vol * (np.kron(M1, np.kron(M1, K1)) / h[0] ** 2 + np.kron(M1, np.kron(K1, M1)) / h[1] ** 2 +
                        np.kron(K1, np.kron(M1, M1)) / h[2] ** 2)
            Cc = self.C[cell_nodes[c]]
            blocks.append(Cc.T @ sp.csr_matrix(Kc) @ Cc)
        for bmat in blocks:
            K = K + bmat
        self.K_free = K.tocsr()                                       # before the boundary condition
        keep = sp.diags((~bnd).astype(float))
        self.K = (keep @ self.K_free @ keep + sp.diags(bnd.astype(float))).tocsr()

    def n_dofs(self):
        return self.regular.size

    def permutation_to(self, library_points):
        """perm with oracle_dof[perm[i]] = library DoF i, matched by coordinates"""
        lut = {tuple(k): d for d, k in enumerate(self._key(self.points))}
        perm = np.array([lut[tuple(k)] for k in self._key(np.asarray(library_points))])
        assert np.unique(perm).size == perm.size == self.n_dofs()
        return perm

    def vmult(self, x):
        return self.K @ x

    def interpolate(self, fn):
        return fn(self.points[:, 0], self.points[:, 1], self.points[:, 2])

    def cell_rows(self, c):
        """constraint rows of cell c: dense (n^3, n_dofs) with the Dirichlet columns zeroed"""
        R = self.C[self.cell_nodes[c]].toarray()
        R[:, self.dirichlet] = 0
        return R
