"""Structure fixtures for the tests: the data of the reference's example structures
(examples-structures/*.yaml, expressions evaluated by hand) restated as plain Python, so
that nothing under /root/reference is read at test time.  YAML line references are to
/root/reference/examples-structures/<name>.yaml.
"""
import numpy as np

from openbeam_b200.structure import build_structure

G = 9.81
DOWN = (0.0, -1.0, 0.0)
IPE200 = {"IPE200": (2.1e11, 28.5e-4, 1940e-8)}       # demo3_frame.yaml:21-24
MY_BAR = {"MY_BAR": (2.0e11, 2e-4, 1940e-8)}          # demo1.yaml:19-22


def demo1(tilted=False):
    """demo1.yaml:27-47 (P = 1000, demo1.yaml parameters) / demo1_tilded_plane.yaml (rot_z 30 on node 1,
    load 5*P on node 1 DX)."""
    P = 1000.0
    nodes = [(0, 0), (0.5, 0, 30.0) if tilted else (0.5, 0), (0, 1)]
    elements = [("BEAM2D_AA", 0, 1, "MY_BAR"), ("BEAM2D_AA", 1, 2, "MY_BAR"), ("BEAM2D_AA", 0, 2, "MY_BAR")]
    cons = [(0, "DXDY"), (1, "DY")]
    loads = [(1, "DX", 5 * P)] if tilted else [(2, "DX", +P)]
    return build_structure(nodes, elements, MY_BAR, cons, loads)


def demo2(bis=False):
    """demo2.yaml:33-57 (RA+AR+RR) / demo2_bis.yaml (RR,RR,RR; q = 1000G, 2000G)."""
    L = 1.0
    P = G * 1000
    nodes = [(0, 0), (L, 0), (2 * L, 0), (2 * L, L)]
    if bis:
        elements = [("BEAM2D_RR", 0, 1, "IPE200"), ("BEAM2D_RR", 1, 2, "IPE200"), ("BEAM2D_RR", 2, 3, "IPE200")]
        eloads = [(0, 1000 * G, DOWN), (1, 2000 * G, DOWN)]
    else:
        elements = [("BEAM2D_RA", 0, 1, "IPE200"), ("BEAM2D_AR", 1, 2, "IPE200"), ("BEAM2D_RR", 2, 3, "IPE200")]
        eloads = [(0, 2000 * G, DOWN), (1, 2000 * G, DOWN)]
    cons = [(0, "DXDYRZ"), (3, "DXDYRZ")]
    return build_structure(nodes, elements, IPE200, cons, [(2, "DX", -P)], eloads)


def demo3_frame(variant="fixed"):
    """demo3_frame.yaml:30-51; variants: 'slider' (demo3_frame_slider.yaml:42-43, node 3 only DY),
    'tilted_slider' (demo3_frame_tilted_slider.yaml:13-14,33: H = 3, node 3 rot_z 30)."""
    L, H = 4.0, (3.0 if variant == "tilted_slider" else 2.0)
    n3 = (L, 0, 30.0) if variant == "tilted_slider" else (L, 0)
    nodes = [(0, 0), (0, H), (L, H), n3]
    elements = [("BEAM2D_RR", 0, 1, "IPE200"), ("BEAM2D_RR", 1, 2, "IPE200"), ("BEAM2D_RR", 2, 3, "IPE200")]
    cons = [(0, "DXDYRZ"), (3, "DXDYRZ" if variant == "fixed" else "DY")]
    return build_structure(nodes, elements, IPE200, cons, [], [(1, 1000 * G, DOWN)])


def ex2021_01():
    """ex2021_01.yaml:11-50 — G=10, P=6*1000*G, MY_BAR = (2.0e11, 6e-4, 1940e-8), 5 nodes, 4 RR."""
    P = 6 * 1000 * 10.0
    sec = {"MY_BAR": (2.0e11, 6e-4, 1940e-8)}
    nodes = [(0, 0), (2, 0), (7, 0), (2, 2), (2.5, 2)]
    elements = [("BEAM2D_RR", 0, 1, "MY_BAR"), ("BEAM2D_RR", 1, 2, "MY_BAR"),
                ("BEAM2D_RR", 1, 3, "MY_BAR"), ("BEAM2D_RR", 3, 4, "MY_BAR")]
    cons = [(0, "DXDYRZ"), (2, "DY")]
    return build_structure(nodes, elements, sec, cons, [(4, "DY", -P)])


def cantilever(L=2.0, P=1000.0, n_el=1, etype="BEAM2D_RR"):
    """Closed-form check: tip deflection PL^3/3EI, rotation PL^2/2EI."""
    nodes = [(L * i / n_el, 0.0) for i in range(n_el + 1)]
    elements = [(etype, i, i + 1, "IPE200") for i in range(n_el)]
    return build_structure(nodes, elements, IPE200, [(0, "DXDYRZ")], [(n_el, "DY", -P)])


def fixed_fixed_udl(L=3.0, q=5000.0):
    """Two RR elements, both ends clamped, UDL: end moments qL^2/12, midspan deflection qL^4/384EI."""
    nodes = [(0, 0), (L / 2, 0), (L, 0)]
    elements = [("BEAM2D_RR", 0, 1, "IPE200"), ("BEAM2D_RR", 1, 2, "IPE200")]
    return build_structure(nodes, elements, IPE200, [(0, "DXDYRZ"), (2, "DXDYRZ")], [],
                           [(0, q, DOWN), (1, q, DOWN)])


def propped_cantilever_udl(L=3.0, q=5000.0):
    """One RA element (clamped at node 0, pinned at node 1): reactions 5/8 qL, 3/8 qL, M = qL^2/8."""
    nodes = [(0, 0), (L, 0)]
    elements = [("BEAM2D_RA", 0, 1, "IPE200")]
    return build_structure(nodes, elements, IPE200, [(0, "DXDYRZ"), (1, "DXDY")], [], [(0, q, DOWN)])


def simply_supported_udl(L=4.0, q=3000.0):
    """AR + RA pair = simply supported beam with a rigid joint at midspan: 5qL^4/384EI."""
    nodes = [(0, 0), (L / 2, 0), (L, 0)]
    elements = [("BEAM2D_AR", 0, 1, "IPE200"), ("BEAM2D_RA", 1, 2, "IPE200")]
    return build_structure(nodes, elements, IPE200, [(0, "DXDY"), (2, "DY")], [], [(0, q, DOWN), (1, q, DOWN)])


def inclined_mixed():
    """Inclined members, mixed end conditions, an inclined UDL and a prescribed displacement-free
    mechanism-free layout: exercises the orientation threshold and every element type at once."""
    nodes = [(0, 0), (1.5, 2.0), (4.0, 2.5), (5.0, 0.0), (2.5, 0.0)]
    sec = dict(IPE200)
    sec["HEB"] = (2.1e11, 53.8e-4, 2492e-8)
    elements = [("BEAM2D_RR", 0, 1, "IPE200"), ("BEAM2D_RA", 1, 2, "HEB"), ("BEAM2D_AR", 2, 3, "IPE200"),
                ("BEAM2D_AA", 1, 4, "HEB"), ("BEAM2D_AA", 4, 2, "IPE200"), ("BEAM2D_RR", 4, 3, "IPE200")]
    cons = [(0, "DXDYRZ"), (3, "DXDY")]
    d = np.array([0.6, -0.8, 0.0])
    return build_structure(nodes, elements, sec, cons, [(1, "DX", 4000.0), (2, "RZ", -700.0), (4, "DY", -2500.0)],
                           [(1, 8000.0, DOWN), (0, 3000.0, tuple(d)), (5, 1200.0, DOWN)])


def load_kinds():
    """Every element-load class of the reference on one frame: DISTRIB_UNIFORM, TRIANGULAR and
    CONCENTRATED (RR beams only, the latter perpendicular to the member) and TEMPERATURE (any beam type, here also
    an inclined RA member and an AA bar)."""
    nodes = [(0, 0), (0, 3.0), (4.0, 3.0), (4.0, 0), (7.0, 3.0), (7.0, 0.5)]
    elements = [("BEAM2D_RR", 0, 1, "IPE200"), ("BEAM2D_RR", 1, 2, "IPE200"), ("BEAM2D_RR", 2, 3, "IPE200"),
                ("BEAM2D_RA", 2, 4, "IPE200"), ("BEAM2D_AA", 4, 5, "IPE200"), ("BEAM2D_RR", 3, 5, "IPE200")]
    cons = [(0, "DXDYRZ"), (3, "DXDY"), (5, "DY")]
    d = (3.0 - 0.5, 3.0, 0.0)   # perpendicular to member 5 (nodes 3 -> 5: axis (3, 0.5))... see below
    ax = np.array([7.0 - 4.0, 0.5 - 0.0])
    perp = np.array([-ax[1], ax[0]]) / np.hypot(*ax)
    eloads = [(1, "TRIANGULAR", 2000.0, 9000.0, DOWN), (1, 1500.0, DOWN), (0, "CONCENTRATED", 4000.0, 1.2, (1.0, 0.0, 0.0)),
              (5, "CONCENTRATED", -2500.0, 0.9, (float(perp[0]), float(perp[1]), 0.0)),
              (3, "TEMPERATURE", 35.0), (4, "TEMPERATURE", -20.0), (2, "TEMPERATURE", 15.0),
              (2, "TRIANGULAR", 500.0, 0.0, (1.0, 0.0, 0.0))]
    del d
    return build_structure(nodes, elements, IPE200, cons, [(1, "DX", 1000.0)], eloads)


def springs_and_slider():
    """The remaining element classes of the reference: BEAM2D_RD (rigid end + slider end), SPRING_1D
    (along its own axis, here vertical), SPRING_XY (vertical member: local x = global -y), SPRING_DXDYRZ between two
    coincident nodes (identity orientation; carries the reference's K11(5,5) = -KRz slip) and SPRING_TORSION.
    Spring constants ride in the section triple (K|Kx, Ky, KRz)."""
    sec = dict(IPE200)
    sec.update(S1=(2e6, 0.0, 0.0), SXY=(1e6, 3e6, 0.0), SXYR=(5e5, 0.0, 2e5), ST=(1e5, 0.0, 0.0))
    nodes = [(0, 0), (3.0, 0), (6.0, 0), (3.0, -1.0), (6.0, -1.5), (6.0, 0), (3.0, 1.0)]
    elements = [("BEAM2D_RR", 0, 1, "IPE200"), ("BEAM2D_RD", 1, 2, "IPE200"), ("SPRING_1D", 1, 3, "S1"),
                ("SPRING_XY", 2, 4, "SXY"), ("SPRING_DXDYRZ", 2, 5, "SXYR"), ("SPRING_TORSION", 1, 6, "ST")]
    cons = [(0, "DXDYRZ"), (3, "DXDY"), (4, "DXDY"), (5, "DXDYRZ"), (6, "RZ")]
    return build_structure(nodes, elements, sec, cons, [(1, "DY", -5e3), (2, "DX", 2e3), (2, "DY", -1.5e3)],
                           [(0, 4000.0, DOWN), (1, "TEMPERATURE", 25.0)])


def old_demo3():
    """old_demo3.yaml:13-58 — textbook frame with a 135-degree rotated support frame and two 1e20 penalty springs tying
    the coincident nodes 2 and 6 (cond(K_ff) ~ 1e13: the reference's own result carries only a few digits)."""
    Gv, D = 9.81, 2.0
    P, Q = Gv * 1000, Gv * 100
    sec = dict(IPE200)
    sec.update({"SXY": (0.0, 1e20, 0.0), "STOR": (1e20, 0.0, 0.0)})
    nodes = [(0, 0, 135.0), (D, D), (2 * D, D), (3 * D, D), (0, D), (2 * D, 0), (2 * D, D)]
    elements = [("BEAM2D_RR", 0, 1, "IPE200"), ("BEAM2D_RR", 1, 2, "IPE200"), ("BEAM2D_RR", 3, 6, "IPE200"),
                ("BEAM2D_AA", 5, 6, "IPE200"), ("BEAM2D_RR", 4, 1, "IPE200"), ("SPRING_XY", 2, 6, "SXY"),
                ("SPRING_TORSION", 2, 6, "STOR")]
    cons = [(0, "DYRZ"), (5, "DXDY"), (3, "DX")]
    return build_structure(nodes, elements, sec, cons, [(6, "DX", -P)], [(4, Q, DOWN)])


def mechanism():
    """demo1 without the DY support at node 1 -> singular K_ff (the reference throws
    'Ill-conditioned matrix. Too few restrictions?', solver_static.cpp:98-105)."""
    st = demo1()
    keep = ~((st.con_node == 1) & (st.con_dof == 1))
    st.con_node, st.con_dof, st.con_val = st.con_node[keep], st.con_dof[keep], st.con_val[keep]
    return st


ALL = {
    "demo1": demo1,
    "demo1_tilded_plane": lambda: demo1(True),
    "demo2": demo2,
    "demo2_bis": lambda: demo2(True),
    "demo3_frame": demo3_frame,
    "demo3_frame_slider": lambda: demo3_frame("slider"),
    "demo3_frame_tilted_slider": lambda: demo3_frame("tilted_slider"),
    "ex2021_01": ex2021_01,
    "cantilever": cantilever,
    "cantilever4": lambda: cantilever(n_el=4),
    "fixed_fixed_udl": fixed_fixed_udl,
    "propped_cantilever_udl": propped_cantilever_udl,
    "simply_supported_udl": simply_supported_udl,
    "inclined_mixed": inclined_mixed,
    "load_kinds": load_kinds,
    "springs_and_slider": springs_and_slider,
}
# ill-conditioned on purpose (1e20 penalty springs): integer artefacts and equilibrium are checked, not 1e-9 value parity
ILL_CONDITIONED = {"old_demo3": old_demo3}
# fixtures without nodal rotated frames (U_b = 0, identity node frames): the round-1 GPU scope
PLAIN = [k for k in ALL if "tilted" not in k and "tilded" not in k]
