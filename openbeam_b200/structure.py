"""SoA description of one 2D beam/truss structure — the data the reference keeps inside
``CStructureProblem`` (nodes, elements, constraints, loads; libopenbeam/include/openbeam/
CFiniteElementProblem.h:475-560) flattened into the arrays ``include/obgpu.h`` takes.

Host-side container only: no arithmetic of the hot path happens here.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

# element types (CElementBeam_2D_{RR,RA,AR,AA}); values match include/obgpu.h
RR, RA, AR, AA = 0, 1, 2, 3
RD, SPRING_1D, SPRING_XY, SPRING_DXDYRZ, SPRING_TORSION = 4, 5, 6, 7, 8   # constants ride in the (E, A, Iz) triple
ELEM_TYPE_NAMES = {"BEAM2D_RR": RR, "BEAM2D_RA": RA, "BEAM2D_AR": AR, "BEAM2D_AA": AA, "BEAM2D_RD": RD,
                   "SPRING_1D": SPRING_1D, "SPRING_XY": SPRING_XY, "SPRING_DXDYRZ": SPRING_DXDYRZ,
                   "SPRING_TORSION": SPRING_TORSION}
# element loads (include/obgpu.h OBGPU_LOAD_*; reference classes in libopenbeam/include/openbeam/loads.h)
LOAD_UNIFORM, LOAD_TRIANGULAR, LOAD_CONCENTRATED, LOAD_TEMPERATURE = 0, 1, 2, 3
LOAD_TYPE_NAMES = {"DISTRIB_UNIFORM": LOAD_UNIFORM, "UNIFORM": LOAD_UNIFORM, "TRIANGULAR": LOAD_TRIANGULAR,
                   "CONCENTRATED": LOAD_CONCENTRATED, "TEMPERATURE": LOAD_TEMPERATURE}
# DoF_index (CFiniteElementProblem.h:53-67)
DX, DY, DZ, RX, RY, RZ = range(6)
# constraint / load DOF names as accepted by the parser (parse_file.cpp listDofNames)
DOF_NAMES = {
    "DX": [DX], "DY": [DY], "DZ": [DZ], "RX": [RX], "RY": [RY], "RZ": [RZ],
    "DXDY": [DX, DY], "DXDYRZ": [DX, DY, RZ], "DYRZ": [DY, RZ], "DXRZ": [DX, RZ],
    "DXDYDZ": [DX, DY, DZ], "RXRYRZ": [RX, RY, RZ], "DXDYDZRXRYRZ": [DX, DY, DZ, RX, RY, RZ],
}


def _a(x, dt, shape=None):
    a = np.ascontiguousarray(np.asarray(x, dtype=dt))
    if shape is not None:
        a = a.reshape(shape)
    return a


@dataclass
class Structure:
    """One structure.  All arrays are contiguous numpy arrays of the listed dtype."""

    node_xy: np.ndarray                      # [N,2] f64
    elem_nodes: np.ndarray                   # [E,2] u32
    elem_type: np.ndarray                    # [E]   u8
    elem_EAI: np.ndarray                     # [E,3] f64  (E, A, Iz)
    con_node: np.ndarray = field(default_factory=lambda: np.zeros(0, np.uint32))
    con_dof: np.ndarray = field(default_factory=lambda: np.zeros(0, np.uint8))
    con_val: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float64))
    nload_node: np.ndarray = field(default_factory=lambda: np.zeros(0, np.uint32))
    nload_dof: np.ndarray = field(default_factory=lambda: np.zeros(0, np.uint8))
    nload_val: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float64))
    eload_elem: np.ndarray = field(default_factory=lambda: np.zeros(0, np.uint32))
    eload_q: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float64))
    eload_dir: np.ndarray = field(default_factory=lambda: np.zeros((0, 3), np.float64))
    node_rotz: np.ndarray | None = None      # [N] f64 radians (nodal rotated frames), None = all zero
    eload_type: np.ndarray | None = None     # [EL] u8 LOAD_* (None = all DISTRIB_UNIFORM)
    eload_q2: np.ndarray | None = None       # [EL] f64 second load parameter (q_end / dist)

    def __post_init__(self):
        self.node_xy = _a(self.node_xy, np.float64, (-1, 2))
        self.elem_nodes = _a(self.elem_nodes, np.uint32, (-1, 2))
        self.elem_type = _a(self.elem_type, np.uint8)
        self.elem_EAI = _a(self.elem_EAI, np.float64, (-1, 3))
        self.con_node = _a(self.con_node, np.uint32)
        self.con_dof = _a(self.con_dof, np.uint8)
        self.con_val = _a(self.con_val, np.float64)
        self.nload_node = _a(self.nload_node, np.uint32)
        self.nload_dof = _a(self.nload_dof, np.uint8)
        self.nload_val = _a(self.nload_val, np.float64)
        self.eload_elem = _a(self.eload_elem, np.uint32)
        self.eload_q = _a(self.eload_q, np.float64)
        self.eload_dir = _a(self.eload_dir, np.float64, (-1, 3))
        if self.eload_type is not None:
            self.eload_type = _a(self.eload_type, np.uint8)
            self.eload_q2 = _a(self.eload_q2 if self.eload_q2 is not None else np.zeros(len(self.eload_type)), np.float64)
            if not np.any(self.eload_type):
                self.eload_type = self.eload_q2 = None
        if self.node_rotz is not None:
            self.node_rotz = _a(self.node_rotz, np.float64)
            if not np.any(self.node_rotz):
                self.node_rotz = None

    @property
    def n_nodes(self):
        return self.node_xy.shape[0]

    @property
    def n_elems(self):
        return self.elem_nodes.shape[0]


def build_structure(nodes, elements, sections, constraints, node_loads=(), element_loads=()):
    """Convenience constructor mirroring the YAML structure-definition sections
    (docs/structure-definition-format.rst): ``nodes`` = [(x, y[, rot_z_deg])],
    ``elements`` = [(type_name, n0, n1, section_name)], ``sections`` = {name: (E, A, Iz)},
    ``constraints`` = [(node, dof_name[, value])], ``node_loads`` = [(node, dof_name, value)],
    ``element_loads`` = [(element, q, (dx, dy, dz))] for DISTRIB_UNIFORM, or
    (element, "TRIANGULAR", q_ini, q_end, dir) / (element, "CONCENTRATED", P, dist, dir) /
    (element, "TEMPERATURE", deltaT)  (parse_file.cpp element-load section)."""
    xy = [(n[0], n[1]) for n in nodes]
    # parse_file.cpp:463-468: rot_z given in degrees -> DEG2RAD
    rotz = [np.deg2rad(n[2]) if len(n) > 2 else 0.0 for n in nodes]
    en = [(e[1], e[2]) for e in elements]
    et = [ELEM_TYPE_NAMES[e[0]] for e in elements]
    eai = [sections[e[3]] for e in elements]
    cn, cd, cv = [], [], []
    for c in constraints:
        for d in DOF_NAMES[c[1]]:
            cn.append(c[0]); cd.append(d); cv.append(c[2] if len(c) > 2 else 0.0)
    ln, ld, lv = [], [], []
    for l in node_loads:
        for d in DOF_NAMES[l[1]]:
            ln.append(l[0]); ld.append(d); lv.append(l[2])
    ee, eq, ed, ety, eq2 = [], [], [], [], []
    for l in element_loads:
        ee.append(l[0])
        if isinstance(l[1], str):
            kind = LOAD_TYPE_NAMES[l[1]]
            ety.append(kind); eq.append(l[2])
            eq2.append(l[3] if kind in (LOAD_TRIANGULAR, LOAD_CONCENTRATED) else 0.0)
            ed.append((1.0, 0.0, 0.0) if kind == LOAD_TEMPERATURE else l[4])
        else:
            ety.append(LOAD_UNIFORM); eq.append(l[1]); eq2.append(0.0); ed.append(l[2])
    return Structure(xy, en, et, eai, cn, cd, cv, ln, ld, lv, ee, eq,
                     np.asarray(ed, np.float64).reshape(-1, 3), rotz, ety if any(ety) else None, eq2 if any(ety) else None)
