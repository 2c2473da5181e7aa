"""CPU restatement of ``CStructureProblem::mesh()`` — TEST INFRASTRUCTURE ONLY (oracle/; see oracle/README or
DESIGN.md §2: only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this).  Parity unpinned: the
reference ships no test or fixture for mesh() and cannot be built here.

Plain Python loops over one structure, written after
  libopenbeam/src/CStructureProblem_mesh.cpp:37-132      (driver: nodes, do_mesh per element, constraints, loads)
  libopenbeam/src/CBaseElementBeam.cpp:317-422           (beam subdivision and the type of each piece)
  libopenbeam/src/CElementSpring.cpp:177-202 (and the three other spring classes: "mesh to myself")
  libopenbeam/src/CLoadDistributedUniform.cpp:174-185, CLoadConstTemperature.cpp:104-115,
  CLoadDistributedTriangular.cpp:186-202, CLoadConcentratedForce.cpp:170-193   (meshLoad)
"""
import math

import numpy as np

RR, RA, AR, AA, RD = 0, 1, 2, 3, 4
UNIFORM, TRIANGULAR, CONCENTRATED, TEMPERATURE = 0, 1, 2, 3


def mesh(st, max_element_length):
    """st: any object with the Structure arrays.  Returns (dict of meshed arrays, element2elements, element2nodes)."""
    assert max_element_length > 0
    nodes = [tuple(map(float, p)) for p in st.node_xy]                      # (1) original nodes (mesh.cpp:44-48)
    rot = [0.0] * len(nodes) if st.node_rotz is None else [float(r) for r in st.node_rotz]
    elems, e2e, e2n = [], [], []
    for e in range(len(st.elem_type)):                                      # (2) do_mesh per element (mesh.cpp:53-57)
        n0, n1 = int(st.elem_nodes[e][0]), int(st.elem_nodes[e][1])
        t = int(st.elem_type[e])
        eai = tuple(float(v) for v in st.elem_EAI[e])
        if t > RD:                                                          # springs: CElementSpring.cpp:177-202
            e2n.append([n0, n1])
            e2e.append([len(elems)])
            elems.append((n0, n1, t, eai))
            continue
        v = (nodes[n1][0] - nodes[n0][0], nodes[n1][1] - nodes[n0][1])
        L = math.sqrt(v[0] * v[0] + v[1] * v[1] + 0.0)                      # CBaseElementBeam.cpp:332-334
        assert L > 0
        n_el = int(math.ceil(L / max_element_length))                       # :340-341
        mine = [n0]
        for k in range(1, n_el):                                            # :352-363
            f = k / float(n_el)
            nodes.append((nodes[n0][0] + v[0] * f, nodes[n0][1] + v[1] * f))
            rot.append(rot[n0])
            mine.append(len(nodes) - 1)
        mine.append(n1)
        mine_e = []
        for k in range(n_el):                                               # :374-420
            if k == 0:
                nt = AR if t in (AR, AA) else RR
            elif k == n_el - 1:
                nt = RA if t in (RA, AA) else (RD if t == RD else RR)
            else:
                nt = RR
            mine_e.append(len(elems))
            elems.append((mine[k], mine[k + 1], nt, eai))
        e2n.append(mine)
        e2e.append(mine_e)
    loads = []                                                              # (4) element loads (mesh.cpp:122-131)
    kinds = [UNIFORM] * len(st.eload_elem) if st.eload_type is None else [int(k) for k in st.eload_type]
    q2s = [0.0] * len(st.eload_elem) if st.eload_q2 is None else [float(v) for v in st.eload_q2]
    order = sorted(range(len(st.eload_elem)), key=lambda i: int(st.eload_elem[i]))   # multimap order (stable)
    for i in order:
        e, kind, q, q2 = int(st.eload_elem[i]), kinds[i], float(st.eload_q[i]), q2s[i]
        d = tuple(float(v) for v in st.eload_dir[i])
        pieces = e2e[e]
        nE = len(pieces)
        if kind in (UNIFORM, TEMPERATURE):
            for pe in pieces:
                loads.append((pe, kind, q, q2, d))
        elif kind == TRIANGULAR:
            Aq = (q2 - q) / nE
            for k, pe in enumerate(pieces):
                qi = q + k * Aq
                loads.append((pe, kind, qi, qi + Aq, d))
        elif kind == CONCENTRATED:
            a, b = elems[pieces[0]][0], elems[pieces[-1]][1]
            Lb = math.sqrt((nodes[b][0] - nodes[a][0]) ** 2 + (nodes[b][1] - nodes[a][1]) ** 2)
            Al = Lb / nE
            idx = min(math.floor(q2 / Al), float(nE - 1))
            loads.append((int(idx), kind, q, q2 - idx * Al, d))              # addLoadAtBeam(idx_affected_element, ...) :192
        else:
            raise ValueError("unknown load kind")
    out = dict(
        node_xy=np.array(nodes, np.float64).reshape(-1, 2), node_rotz=np.array(rot, np.float64),
        elem_nodes=np.array([[a, b] for a, b, _, _ in elems], np.uint32).reshape(-1, 2),
        elem_type=np.array([t for _, _, t, _ in elems], np.uint8),
        elem_EAI=np.array([x for _, _, _, x in elems], np.float64).reshape(-1, 3),
        con_node=np.array(st.con_node, np.uint32), con_dof=np.array(st.con_dof, np.uint8),
        con_val=np.array(st.con_val, np.float64), nload_node=np.array(st.nload_node, np.uint32),
        nload_dof=np.array(st.nload_dof, np.uint8), nload_val=np.array(st.nload_val, np.float64),
        eload_elem=np.array([l[0] for l in loads], np.uint32), eload_type=np.array([l[1] for l in loads], np.uint8),
        eload_q=np.array([l[2] for l in loads], np.float64), eload_q2=np.array([l[3] for l in loads], np.float64),
        eload_dir=np.array([l[4] for l in loads], np.float64).reshape(-1, 3))
    return out, e2e, e2n
