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


# ======================================================================================================================
# Local-smoothing multigrid on the adaptive mesh (config 5, poisson_l/program.cc:283-420)
# ======================================================================================================================
def _chebyshev_setup(A, inv_diag, constrained, smoothing_range, degree, eig_cg_n_iterations):
    """deal.II PreconditionChebyshev set-up on an assembled matrix: CG/Lanczos estimate of the
    largest eigenvalue of D^-1 A from the start vector "index mod 11, mean removed, constrained entries zero" (the
    indices are those of the level numbering, sorted by z, y, x of the support point), then theta / delta."""
    n = A.shape[0]
    pattern = (np.arange(n) % 11).astype(np.float64)
    g = -(pattern - pattern.sum() / n)
    g[constrained] = 0
    diag, off = [], []
    res = np.sqrt(g @ g)
    its = 0
    min_eig = max_eig = 1.0
    if res > 1e-10 and eig_cg_n_iterations > 0:
        h = inv_diag * g
        d = -h
        gh = g @ h
        alpha_old = beta_old = 0.0
        it = 1
        while True:
            h = A @ d
            alpha = gh / (d @ h)
            g = g + alpha * h
            res = np.sqrt(g @ g)
            if it == 1:
                diag.append(1.0 / alpha)
            else:
                diag.append(1.0 / alpha + beta_old / alpha_old)
                off.append(np.sqrt(beta_old) / alpha_old)
            its = it
            if it >= eig_cg_n_iterations or res <= 1e-10:
                break
            h = inv_diag * g
            gh_new = g @ h
            beta = gh_new / gh
            gh = gh_new
            d = beta * d - h
            alpha_old, beta_old = alpha, beta
            it += 1
        T = np.diag(diag) + np.diag(off, 1) + np.diag(off, -1)
        ev = np.linalg.eigvalsh(T)
        min_eig, max_eig = ev[0], ev[-1]
    max_est = 1.2 * max_eig
    alpha = max_est / smoothing_range if smoothing_range > 1.0 else min(0.9 * max_est, min_eig)
    if degree is None:
        actual_range = max_est / alpha
        sigma = (1.0 - np.sqrt(1.0 / actual_range)) / (1.0 + np.sqrt(1.0 / actual_range))
        eps = smoothing_range
        degree = 1 + int(np.log(1.0 / eps + np.sqrt(1.0 / eps / eps - 1.0)) / np.log(1.0 / sigma))
    return dict(theta=0.5 * (max_est + alpha), delta=0.5 * (max_est - alpha), degree=degree, min_eig=min_eig,
                max_eig=max_eig, cg_iterations=its)


def _chebyshev_apply(A, inv_diag, info, b, x=None):
    """vmult (x = None: zero initial guess, degree - 1 matrix products) or step (degree products)"""
    theta, delta, degree = info["theta"], info["delta"], info["degree"]
    if x is None:
        x = inv_diag * b / theta
        x_old = np.zeros_like(x)
    else:
        x_old = x
        x = x + inv_diag * (b - A @ x) / theta
    if degree < 2:
        return x
    rhok, sigma = delta / theta, theta / delta
    for _ in range(degree - 1):
        rhokp = 1.0 / (2.0 * sigma - rhok)
        f1, f2 = rhokp * rhok, 2.0 * rhokp / delta
        rhok = rhokp
        x, x_old = x + f1 * (x - x_old) + f2 * inv_diag * (b - A @ x), x
    return x


class LocalSmoothingOracle:
    """PreconditionMG with deal.II's local-smoothing Multigrid on the 2:1-balanced forest, assembled with scipy.

    Level l holds all cells of size root_size / 2^l that are active or refined (parallel::distributed::Triangulation
    with construct_multigrid_hierarchy).  Level operators (poisson_l/program.cc:283-305): no hanging nodes inside a
    level; the DoFs on the domain boundary AND on the refinement edge (faces towards coarser active cells) are
    constrained to zero with identity rows (mg_matrices[level].initialize(mf, mg_constrained_dofs, level)).  Edge
    matrices (:374-383, MGInterfaceOperator; semantics at gpu/poisson_l.cu:390-436): interface_down(x) = rows of the
    refinement-edge DoFs of K_level applied to x with the constrained entries zeroed; interface_up(x) = K_level applied
    to the refinement-edge part of x, constrained rows zeroed.  Transfer: MGTransferMatrixFree with
    initialize_constraints (:312-313): interpolation of the parent's polynomial, boundary DoFs zero, restriction its
    transpose.  copy_to_mg / copy_from_mg: the DoFs of the active cells of a level except those on its refinement
    edge.  Smoothers (:346-363): Chebyshev(point Jacobi), range 15, degree n_smoother, 15 CG steps for the eigenvalue
    estimate; level 0: range 1e-3 as tolerance, degree automatic, m() CG steps.  V-cycle: deal.II
    Multigrid::level_v_step with edge_out = edge_in = the interface matrices.  Outer solver (:385-392): SolverCG,
    ReductionControl(m(), 1e-13, 1e-9), zero start vector.  **Parity unpinned** against deal.II (not available)."""

    def __init__(self, degree, cell_lower, cell_size, root_size, origin, n_smoother=3):
        self.active = AdaptiveOracle(degree, cell_lower, cell_size)
        self.p, self.n = degree, degree + 1
        self.origin = np.asarray(origin, dtype=np.float64)
        self.root = np.asarray(root_size, dtype=np.float64)
        lo, sz = self.active.lo, self.active.sz
        lev = np.round(np.log2(self.root[0] / sz[:, 0])).astype(int)
        self.n_levels = int(lev.max()) + 1
        self.scale = sz.min()
        b = orc.basis(degree)
        self.gll = b["gll"]
        # ---- level cell sets: integer positions of active cells and of their ancestors
        pos = np.round((lo - self.origin) / sz).astype(np.int64)
        cells = [set() for _ in range(self.n_levels)]
        active_at = [set() for _ in range(self.n_levels)]
        for c in range(lo.shape[0]):
            l, q = int(lev[c]), tuple(int(v) for v in pos[c])
            active_at[l].add(q)
            while True:
                cells[l].add(q)
                if l == 0:
                    break
                l, q = l - 1, tuple(v // 2 for v in q)
        self.levels = []
        for l in range(self.n_levels):
            self.levels.append(self._build_level(l, sorted(cells[l], key=lambda q: (q[2], q[1], q[0])), active_at[l]))
        for l in range(1, self.n_levels):
            self._build_transfer(l)
        self._build_copy_indices()
        for l, L in enumerate(self.levels):
            n = L["A"].shape[0]
            L["inv_diag"] = 1.0 / L["A"].diagonal()
            if l > 0:
                L["cheb"] = _chebyshev_setup(L["A"], L["inv_diag"], L["constrained"], 15.0, n_smoother, 15)
            else:
                L["cheb"] = _chebyshev_setup(L["A"], L["inv_diag"], L["constrained"], 1e-3, None, n)

    # ------------------------------------------------------------------------------------------------------------------
    def _key(self, pts):
        return np.round((pts - self.origin) / (self.scale * 1e-6)).astype(np.int64)

    def _build_level(self, l, cell_pos, active_pos):
        n, g = self.n, self.gll
        h = self.root / 2 ** l
        P = np.array(cell_pos, dtype=np.int64).reshape(-1, 3)
        lo = self.origin + P * h
        k, j, i = np.meshgrid(np.arange(n), np.arange(n), np.arange(n), indexing="ij")
        ref = np.stack([g[i], g[j], g[k]], axis=-1).reshape(-1, 3)
        X = lo[:, None, :] + h * ref[None, :, :]
        keys = self._key(X).reshape(-1, 3)
        # level numbering: sorted by (z, y, x) of the support point
        uniq, inverse = np.unique(keys[:, ::-1], axis=0, return_inverse=True)
        nd = uniq.shape[0]
        cell_dofs = inverse.reshape(P.shape[0], n ** 3)
        points = np.zeros((nd, 3))
        points[cell_dofs.ravel()] = X.reshape(-1, 3)
        # boundary / refinement edge: look into the eight octants around a support point
        in_level = set(cell_pos)
        alo, asz = self.active.lo, self.active.sz
        eps = 1e-4 * self.scale
        boundary = np.zeros(nd, dtype=bool)
        edge = np.zeros(nd, dtype=bool)
        for sx in (-eps, eps):
            for sy in (-eps, eps):
                for sz_ in (-eps, eps):
                    Q = points + np.array([sx, sy, sz_])
                    inside = np.zeros(nd, dtype=bool)
                    for c in range(alo.shape[0]):
                        inside |= np.all((Q > alo[c]) & (Q < alo[c] + asz[c]), axis=1)
                    boundary |= ~inside
                    qpos = np.floor((Q - self.origin) / h).astype(np.int64)
                    covered = np.array([tuple(v) in in_level for v in qpos])
                    edge |= inside & ~covered
        edge &= ~boundary
        constrained = boundary | edge
        # assembly without constraints
        M1, K1 = self.active.M1, self.active.K1
        vol = h[0] * h[1] * h[2]
        Kc = vol * (np.kron(M1, np.kron(M1, K1)) / h[0] ** 2 + np.kron(M1, np.kron(K1, M1)) / h[1] ** 2 +
                    np.kron(K1, np.kron(M1, M1)) / h[2] ** 2)
        rows = np.repeat(cell_dofs, n ** 3, axis=1).ravel()
        cols = np.tile(cell_dofs, (1, n ** 3)).ravel()
        K = sp.csr_matrix((np.tile(Kc.ravel(), P.shape[0]), (rows, cols)), shape=(nd, nd))
        keep = sp.diags((~constrained).astype(float))
        A = (keep @ K @ keep + sp.diags(constrained.astype(float))).tocsr()
        return dict(pos=cell_pos, pos_index={q: c for c, q in enumerate(cell_pos)}, active=active_pos, h=h, lo=lo,
                    cell_dofs=cell_dofs, points=points, boundary=boundary, edge=edge, constrained=constrained, K=K, A=A,
                    n=nd)

    def _build_transfer(self, l):
        """P (n_l x n_{l-1}): value of the parent's polynomial at every support point of its children"""
        F, C = self.levels[l], self.levels[l - 1]
        n, g = self.n, self.gll
        rows, cols, vals = [], [], []
        done = np.zeros(F["n"], dtype=bool)
        for c, q in enumerate(F["pos"]):
            parent = C["pos_index"][tuple(v // 2 for v in q)]
            plo, ph = C["lo"][parent], C["h"]
            for loc, f in enumerate(F["cell_dofs"][c]):
                if done[f]:
                    continue
                done[f] = True
                xi = (F["points"][f] - plo) / ph
                w = [lagrange(g, xi[d])[0] for d in range(3)]
                W = np.einsum("c,b,a->cba", w[2], w[1], w[0]).ravel()
                nz = np.abs(W) > 1e-13
                rows += [f] * int(nz.sum())
                cols += list(C["cell_dofs"][parent][nz])
                vals += list(W[nz])
        P = sp.csr_matrix((vals, (rows, cols)), shape=(F["n"], C["n"]))
        P = sp.diags((~F["boundary"]).astype(float)) @ P @ sp.diags((~C["boundary"]).astype(float))
        F["P"] = P.tocsr()

    def _build_copy_indices(self):
        """(active DoF, level DoF) pairs per level: support points of the active cells of the level that carry an
        unconstrained active DoF and do not lie on the level's refinement edge"""
        act = self.active
        lut = {tuple(k): d for d, k in enumerate(act._key(act.points))}
        for L in self.levels:
            pairs = {}
            for q in L["active"]:
                c = L["pos_index"][q]
                for f in L["cell_dofs"][c]:
                    if L["edge"][f] or L["boundary"][f]:
                        continue
                    d = lut.get(tuple(act._key(L["points"][f][None, :])[0]))
                    if d is not None and not act.dirichlet[d]:
                        pairs[int(f)] = d
            L["copy_level"] = np.array(sorted(pairs), dtype=np.int64)
            L["copy_active"] = np.array([pairs[f] for f in sorted(pairs)], dtype=np.int64)

    # ------------------------------------------------------------------------------------------------------------------
    def interface_down(self, l, x):
        L = self.levels[l]
        y = L["K"] @ np.where(L["constrained"], 0.0, x)
        return np.where(L["edge"], y, 0.0)

    def interface_up(self, l, x):
        L = self.levels[l]
        y = L["K"] @ np.where(L["edge"], x, 0.0)
        return np.where(L["constrained"], 0.0, y)

    def level_v_step(self, l, defect, solution):
        L = self.levels[l]
        if l == 0:
            solution[0] = _chebyshev_apply(L["A"], L["inv_diag"], L["cheb"], defect[0])
            return
        solution[l] = _chebyshev_apply(L["A"], L["inv_diag"], L["cheb"], defect[l])
        t = L["A"] @ solution[l] + self.interface_down(l, solution[l])
        t = defect[l] - t
        defect[l - 1] = defect[l - 1] + L["P"].T @ t
        self.level_v_step(l - 1, defect, solution)
        solution[l] = solution[l] + L["P"] @ solution[l - 1]
        defect[l] = defect[l] - self.interface_up(l, solution[l])
        solution[l] = _chebyshev_apply(L["A"], L["inv_diag"], L["cheb"], defect[l], solution[l])

    def vmult(self, src):
        """PreconditionMG::vmult: copy_to_mg, one V-cycle, copy_from_mg"""
        defect = [np.zeros(L["n"]) for L in self.levels]
        solution = [np.zeros(L["n"]) for L in self.levels]
        for l, L in enumerate(self.levels):
            defect[l][L["copy_level"]] = src[L["copy_active"]]
        self.level_v_step(self.n_levels - 1, defect, solution)
        dst = np.zeros_like(src)
        for l, L in enumerate(self.levels):
            dst[L["copy_active"]] = solution[l][L["copy_level"]]
        return dst

    def solve_cg(self, rhs, rel_tol=1e-9, abs_tol=1e-13):
        """deal.II SolverCG on the active operator, preconditioned by vmult; returns (x, iterations, residual history)"""
        K = self.active.K
        x = np.zeros_like(rhs)
        g = -rhs.copy()
        res0 = res = np.sqrt(g @ g)
        hist = [res]
        it = 0
        if res > abs_tol:
            h = self.vmult(g)
            d = -h
            gh = g @ h
            while True:
                it += 1
                h = K @ d
                alpha = gh / (d @ h)
                x += alpha * d
                g += alpha * h
                res = np.sqrt(g @ g)
                hist.append(res)
                if res <= abs_tol or res <= rel_tol * res0 or it >= rhs.size:
                    break
                h = self.vmult(g)
                gh_new = g @ h
                beta = gh_new / gh
                gh = gh_new
                d = beta * d - h
        return x, it, np.array(hist)


# ======================================================================================================================
# The adaptive loop around the solver (poisson_l/program.cc:497-602): boundary data, right-hand side, Kelly error
# estimator, fixed-number refinement and coarsening, error norms.  Geometric like the classes above; small meshes only.
# ======================================================================================================================
def lagrange_derivative(nodes, x):
    """derivatives of all Lagrange polynomials on `nodes` at the points x: shape (len(x), len(nodes))"""
    x = np.atleast_1d(x)
    out = np.zeros((x.size, nodes.size))
    for i in range(nodes.size):
        for m in range(nodes.size):
            if m == i:
                continue
            t = np.ones(x.size) / (nodes[i] - nodes[m])
            for l in range(nodes.size):
                if l != i and l != m:
                    t *= (x - nodes[l]) / (nodes[i] - nodes[l])
            out[:, i] += t
    return out


def l_singularity(x, y, z=None):
    """deal.II Functions::LSingularityFunction (the exact solution of poisson_l/program.cc:85-135)"""
    phi = np.arctan2(y, -x) + np.pi
    r2 = x * x + y * y
    return np.where((x >= 0) & (y >= 0), 0., r2 ** (1. / 3.) * np.sin(2. / 3. * phi))


def l_singularity_gradient(x, y, z=None):
    phi = np.arctan2(y, -x) + np.pi
    r43 = (x * x + y * y) ** (2. / 3.)
    gx = 2. / 3. * (np.sin(2. / 3. * phi) * x + np.cos(2. / 3. * phi) * y) / r43
    gy = 2. / 3. * (np.sin(2. / 3. * phi) * y - np.cos(2. / 3. * phi) * x) / r43
    out = (x >= 0) & (y >= 0)
    return np.where(out, 0., gx), np.where(out, 0., gy), np.zeros_like(gx)


def gauss_01(nq):
    x, w = np.polynomial.legendre.leggauss(nq)
    return 0.5 * (x + 1), 0.5 * w


class AdaptiveLoopOracle(AdaptiveOracle):
    def cell_values(self, u):
        """nodal values of every cell, hanging nodes interpolated, Dirichlet values kept: (cells, n^3)"""
        return np.stack([self.C[self.cell_nodes[c]] @ u for c in range(self.lo.shape[0])])

    def boundary_values(self, fn):
        """interpolate_boundary_values (:321-333): fn on the Dirichlet DoFs, zero elsewhere"""
        u = np.zeros(self.n_dofs())
        P = self.points[self.dirichlet]
        u[self.dirichlet] = fn(P[:, 0], P[:, 1], P[:, 2])
        return u

    def compute_residual(self, u, f=0.):
        """(f, phi) - K u on the rows without Dirichlet condition (laplace_operator.h:806-845)"""
        xq, wq = gauss_01(self.n)
        m1 = wq @ lagrange(self.gll, xq)
        m3 = np.einsum("c,b,a->cba", m1, m1, m1).ravel()
        load = np.zeros(self.n_dofs())
        for c in range(self.lo.shape[0]):
            load += self.C[self.cell_nodes[c]].T @ (np.prod(self.sz[c]) * m3)
        r = f * load - self.K_free @ u
        r[self.dirichlet] = 0
        return r

    def evaluate(self, vals_c, c, X):
        """value and gradient of the polynomial of cell c (nodal values vals_c) at physical points X (m, 3)"""
        n = self.n
        xi = (X - self.lo[c]) / self.sz[c]
        L = [lagrange(self.gll, xi[:, d]) for d in range(3)]
        D = [lagrange_derivative(self.gll, xi[:, d]) / self.sz[c, d] for d in range(3)]
        V = vals_c.reshape(n, n, n)
        val = np.einsum("mk,mj,mi,kji->m", L[2], L[1], L[0], V)
        g = [np.einsum("mk,mj,mi,kji->m", L[2], L[1], D[0], V), np.einsum("mk,mj,mi,kji->m", L[2], D[1], L[0], V),
             np.einsum("mk,mj,mi,kji->m", D[2], L[1], L[0], V)]
        return val, g

    def _cell_at(self, point):
        hit = np.where(np.all((point > self.lo) & (point < self.lo + self.sz), axis=1))[0]
        return int(hit[0]) if hit.size else -1

    def kelly(self, u):
        """KellyErrorEstimator::estimate with QGauss(p+1) on the faces, strategy cell_diameter_over_24, no Neumann data
        (:524-531): eta_K = sqrt(diameter_K / 24 * sum over the faces of int_F [du/dn]^2); a face between cells of
        different size is integrated over the small cells' faces and counted on both sides"""
        vals = self.cell_values(u)
        xq, wq = gauss_01(self.n)
        nc = self.lo.shape[0]
        eps = 1e-3 * self.sz.min()

        def face_integral(small, d, side, other):
            t1, t2 = [k for k in range(3) if k != d]
            A, B = np.meshgrid(xq, xq, indexing="ij")
            X = np.zeros((A.size, 3))
            X[:, d] = self.lo[small, d] + (self.sz[small, d] if side else 0.)
            X[:, t1] = self.lo[small, t1] + self.sz[small, t1] * A.ravel()
            X[:, t2] = self.lo[small, t2] + self.sz[small, t2] * B.ravel()
            W = np.outer(wq, wq).ravel() * self.sz[small, t1] * self.sz[small, t2]
            _, g1 = self.evaluate(vals[small], small, X)
            _, g2 = self.evaluate(vals[other], other, X)
            return float(W @ (g1[d] - g2[d]) ** 2)

        eta = np.zeros(nc)
        for c in range(nc):
            total = 0.
            for d in range(3):
                t1, t2 = [k for k in range(3) if k != d]
                for side in (0, 1):
                    nbs = []
                    for a in (0.25, 0.75):
                        for b in (0.25, 0.75):
                            P = self.lo[c].copy()
                            P[d] += self.sz[c, d] + eps if side else -eps
                            P[t1] += a * self.sz[c, t1]
                            P[t2] += b * self.sz[c, t2]
                            nbs.append(self._cell_at(P))
                    if nbs[0] < 0:
                        continue                                   # boundary face
                    if len(set(nbs)) == 1:                         # same size or larger
                        total += face_integral(c, d, side, nbs[0])
                    else:                                          # four smaller neighbours: their faces
                        for nb in nbs:
                            total += face_integral(nb, d, 1 - side, c)
            eta[c] = np.sqrt(np.linalg.norm(self.sz[c]) / 24. * total)
        return eta

    def integrate_difference(self, u, fn, grad, norm, nq):
        """VectorTools::integrate_difference (:556-575): L2 (norm 0) or H1 seminorm (norm 1) -> (norm, cell squares)"""
        vals = self.cell_values(u)
        xq, wq = gauss_01(nq)
        K, J, I = np.meshgrid(xq, xq, xq, indexing="ij")
        W = np.einsum("c,b,a->cba", wq, wq, wq).ravel()
        sq = np.zeros(self.lo.shape[0])
        for c in range(self.lo.shape[0]):
            X = self.lo[c] + self.sz[c] * np.stack([I.ravel(), J.ravel(), K.ravel()], axis=1)
            v, g = self.evaluate(vals[c], c, X)
            if norm == 0:
                d2 = (v - fn(X[:, 0], X[:, 1], X[:, 2])) ** 2
            else:
                e = grad(X[:, 0], X[:, 1], X[:, 2])
                d2 = sum((g[k] - e[k]) ** 2 for k in range(3))
            sq[c] = np.prod(self.sz[c]) * (W @ d2)
        return np.sqrt(sq.sum()), sq


def fixed_number_flags(indicators, top_fraction, bottom_fraction):
    """GridRefinement::refine_and_coarsen_fixed_number (serial): thresholds at the int(fraction N)-th largest / smallest
    indicator, ties flagged, refinement wins"""
    e = np.asarray(indicators, dtype=np.float32)
    n_ref, n_coa = int(top_fraction * e.size), int(bottom_fraction * e.size)
    ref = np.zeros(e.size, dtype=bool)
    coa = np.zeros(e.size, dtype=bool)
    if n_ref > 0:
        ref = e >= np.sort(e)[::-1][n_ref - 1]
    if n_coa > 0:
        coa = (e <= np.sort(e)[n_coa - 1]) & ~ref
    return ref, coa


def refine_and_coarsen(cells, refine, coarsen, n_roots, root_mask):
    """execute_coarsening_and_refinement on a forest given as a list of (level, x, y, z): refine the flagged cells keeping
    the 2:1 balance over faces, edges and corners; then replace complete families of coarsen-flagged siblings by their
    parent unless it would touch a cell more than one level finer.  Returns the new set of cells."""
    def inside(l, q):
        for d in range(3):
            if q[d] < 0 or q[d] >= (n_roots[d] << l):
                return False
        r = [q[d] >> l for d in range(3)]
        return root_mask is None or bool(root_mask[r[2]][r[1]][r[0]])

    active = set(cells)
    wish = {c for c, r, k in zip(cells, refine, coarsen) if k and not r and c[0] > 0}
    todo = [c for c, r in zip(cells, refine) if r]
    marked = set(todo)
    while todo:
        l, x, y, z = todo.pop()
        if l == 0:
            continue
        for dz in (-1, 0, 1):
            for dy in (-1, 0, 1):
                for dx in (-1, 0, 1):
                    q = (x + dx, y + dy, z + dz)
                    if (dx, dy, dz) == (0, 0, 0) or not inside(l, q):
                        continue
                    par = (l - 1, q[0] >> 1, q[1] >> 1, q[2] >> 1)
                    if par in active and par not in marked:
                        marked.add(par)
                        todo.append(par)
    new = set()
    for c in active:
        if c in marked:
            l, x, y, z = c
            new |= {(l + 1, 2 * x + a, 2 * y + b, 2 * z + e) for a in (0, 1) for b in (0, 1) for e in (0, 1)}
        else:
            new.add(c)
    parents = []
    for c in sorted(new):
        l, x, y, z = c
        if l == 0 or (x | y | z) & 1:
            continue
        family = [(l, x + a, y + b, z + e) for a in (0, 1) for b in (0, 1) for e in (0, 1)]
        if not all(k in new and k in wish for k in family):
            continue
        ok = True
        for dz in range(-1, 3):
            for dy in range(-1, 3):
                for dx in range(-1, 3):
                    if 0 <= dx <= 1 and 0 <= dy <= 1 and 0 <= dz <= 1:
                        continue
                    q = (x + dx, y + dy, z + dz)
                    if not inside(l, q):
                        continue
                    covered = any((l - s, q[0] >> s, q[1] >> s, q[2] >> s) in new for s in range(l + 1))
                    ok &= covered
        if ok:
            parents.append((family, (l - 1, x >> 1, y >> 1, z >> 1)))
    for family, par in parents:
        new -= set(family)
        new.add(par)
    return new
