"""Meshes shared by the adaptive-operator tests: hyper_L / box forests refined towards the re-entrant edge or a
corner (the pattern the Kelly estimator produces in poisson_l/program.cc:508-541)."""
import numpy as np

HYPER_L = np.ones((2, 2, 2), dtype=np.uint8)
HYPER_L[1, 1, 1] = 0          # GridGenerator::hyper_L: [-1,1]^3 without the octant (0,1)^3   [z][y][x]
L_EXTRUDED = np.ones((1, 2, 2), dtype=np.uint8)
L_EXTRUDED[0, 1, 1] = 0       # 2-D L extruded in z (three root cells)

CASES = {   # name: (degree, n_roots, mask, n_global, refinement rounds: (kind, parameter))
    "hyper_l_p2": (2, (2, 2, 2), HYPER_L, 1, [("edge", 0.55), ("edge", 0.3)]),
    "hyper_l_p3": (3, (2, 2, 2), HYPER_L, 0, [("corner", 1.2), ("corner", 0.6)]),
    "extruded_l_p1": (1, (2, 2, 1), L_EXTRUDED, 1, [("edge", 0.6), ("edge", 0.35), ("edge", 0.2)]),
    "box_p4": (4, (2, 1, 1), None, 0, [("corner", 1.1)]),
    "box_point_p2": (2, (1, 1, 1), None, 2, [("point", 0.3), ("point", 0.2)]),
}


def make(mgb, name):
    p, nr, mask, ng, rounds = CASES[name]
    lower = (-1., -1., -1.)
    upper = (1., 1., 1.) if nr[2] == 2 or nr == (1, 1, 1) else ((1., 1., 0.) if nr == (2, 2, 1) else (3., 1., 1.))
    op = mgb.LaplaceOperatorAdaptive(p, nr, mask, lower, upper, ng)
    for kind, r in rounds:
        lo, sz = op.cell_boxes()
        c = lo + 0.5 * sz
        if kind == "edge":        # distance to the re-entrant edge x = y = 0
            d = np.hypot(c[:, 0], c[:, 1])
        elif kind == "corner":    # distance to the origin
            d = np.linalg.norm(c, axis=1)
        else:                     # an interior point off the lattice: three levels next to each other
            d = np.linalg.norm(c - np.array([0.37, -0.21, 0.13]), axis=1)
        op.refine(d < r)
    return op
