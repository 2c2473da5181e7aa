"""Synthetic workloads of BASELINE.json's configs: counter-based RNG
``splitmix64(seed ^ index)`` so every rank / the CPU baseline / the tests see identical inputs
without exchanging them.  Host-side input generation only (numpy) — no hot-path arithmetic.
"""
from __future__ import annotations

import numpy as np

from .structure import AA, AR, RA, RR, DX, DY, RZ, Structure, build_structure

SEED = 0x0B5EED
_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    """Vectorised splitmix64 finaliser over uint64 counters."""
    with np.errstate(over="ignore"):
        z = (np.asarray(x, np.uint64) + np.uint64(0x9E3779B97F4A7C15)) & _M
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M
        return z ^ (z >> np.uint64(31))


def uniform01(stream, index):
    """U[0,1) doubles for (stream, index) pairs; 53 random mantissa bits."""
    with np.errstate(over="ignore"):
        key = splitmix64(np.uint64(SEED) ^ (np.asarray(stream, np.uint64) * np.uint64(0x100000001B3)))
        bits = splitmix64(key ^ np.asarray(index, np.uint64))
    return (bits >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def uniform(stream, index, lo, hi):
    return lo + (hi - lo) * uniform01(stream, index)


# data/beam_profiles.txt IPE rows (A cm^2 -> m^2, Iy cm^4 -> m^4): name, A, I
IPE = np.array([
    (7.64e-4, 80.1e-8), (10.3e-4, 171e-8), (13.2e-4, 318e-8), (16.4e-4, 541e-8), (20.1e-4, 869e-8),
    (23.9e-4, 1320e-8), (28.5e-4, 1940e-8), (33.4e-4, 2770e-8), (39.1e-4, 3890e-8), (45.9e-4, 5790e-8),
    (53.8e-4, 8360e-8), (62.6e-4, 11770e-8), (72.7e-4, 16270e-8), (84.5e-4, 23130e-8), (98.8e-4, 33740e-8),
    (116e-4, 48200e-8), (134e-4, 67120e-8), (156e-4, 92080e-8),
])
E_STEEL = 2.1e11


def demo3_frame() -> Structure:
    """examples-structures/demo3_frame.yaml:30-51 with the three extra nodal-load slots of config C2
    (Fx,Fy,Mz on nodes 1 and 2; base values 0 so the base structure equals the YAML file's)."""
    L, H = 4.0, 2.0
    sec = {"IPE200": (E_STEEL, 28.5e-4, 1940e-8)}
    nodes = [(0, 0), (0, H), (L, H), (L, 0)]
    elements = [("BEAM2D_RR", 0, 1, "IPE200"), ("BEAM2D_RR", 1, 2, "IPE200"), ("BEAM2D_RR", 2, 3, "IPE200")]
    cons = [(0, "DXDYRZ"), (3, "DXDYRZ")]
    nl = [(1, "DX", 0.0), (1, "DY", 0.0), (1, "RZ", 0.0), (2, "DX", 0.0), (2, "DY", 0.0), (2, "RZ", 0.0)]
    return build_structure(nodes, elements, sec, cons, nl, [(1, 1000 * 9.81, (0.0, -1.0, 0.0))])


def c2_cases(first, count, per_case_section=False, out_nl=None, out_q=None, out_eai=None):
    """Load cases [first, first+count) of config C2: returns (case_nload_val [count,6], case_eload_q [count,1],
    case_elem_EAI [count,9] or None).  q = 9810*U(0.5,1.5); Fx,Fy ~ U(-1e4,1e4); Mz ~ U(-1e3,1e3)."""
    idx = np.arange(first, first + count, dtype=np.uint64)
    nl = out_nl if out_nl is not None else np.empty((count, 6))
    q = out_q if out_q is not None else np.empty((count, 1))
    for k in range(6):
        a = 1e3 if k % 3 == 2 else 1e4
        nl[:, k] = uniform(10 + k, idx, -a, a)
    q[:, 0] = 9810.0 * uniform(1, idx, 0.5, 1.5)
    eai = None
    if per_case_section:
        eai = out_eai if out_eai is not None else np.empty((count, 9))
        row = (uniform01(20, idx) * len(IPE)).astype(np.int64)
        for e in range(3):
            eai[:, 3 * e] = E_STEEL
            eai[:, 3 * e + 1] = IPE[row, 0]
            eai[:, 3 * e + 2] = IPE[row, 1]
    return nl, q, eai


def c3_frame(index) -> Structure:
    """Structure ``index`` of config C3: random small 2D frame, b bays x s storeys with
    E = s(2b+1) in [10,60]; columns RR, beams {RR .6, RA .15, AR .15, AA .1}; bases fixed; one section
    for columns and one for beams; UDL on every beam; wind Fx on the left-column nodes."""
    i64 = np.uint64(index)
    t = 0
    while True:
        b = 1 + int(uniform01(30 + t, i64) * 5)
        s = 1 + int(uniform01(60 + t, i64) * 6)
        if 10 <= s * (2 * b + 1) <= 60:
            break
        t += 1
    W = uniform(100, i64 * np.uint64(8) + np.arange(b, dtype=np.uint64), 3.0, 6.0)
    Hh = uniform(101, i64 * np.uint64(8) + np.arange(s, dtype=np.uint64), 2.5, 4.0)
    xs = np.concatenate([[0.0], np.cumsum(W)])
    ys = np.concatenate([[0.0], np.cumsum(Hh)])
    nid = lambda i, j: j * (b + 1) + i
    xy = [(xs[i], ys[j]) for j in range(s + 1) for i in range(b + 1)]
    csec = IPE[int(uniform01(102, i64) * len(IPE))]
    bsec = IPE[int(uniform01(103, i64) * len(IPE))]
    en, et, eai, ee, eq = [], [], [], [], []
    for j in range(1, s + 1):
        for i in range(b + 1):  # columns of storey j
            en.append((nid(i, j - 1), nid(i, j))); et.append(RR); eai.append((E_STEEL, csec[0], csec[1]))
        for i in range(b):      # beams of storey j
            u = uniform01(104, i64 * np.uint64(64) + np.uint64(j * 8 + i))
            ty = RR if u < 0.6 else RA if u < 0.75 else AR if u < 0.9 else AA
            ee.append(len(en))
            eq.append(uniform(105, i64 * np.uint64(64) + np.uint64(j * 8 + i), 5e3, 30e3))
            en.append((nid(i, j), nid(i + 1, j))); et.append(ty); eai.append((E_STEEL, bsec[0], bsec[1]))
    cn = np.repeat(np.arange(b + 1), 3)
    cd = np.tile([DX, DY, RZ], b + 1)
    ln = [nid(0, j) for j in range(1, s + 1)]
    lv = uniform(106, i64 * np.uint64(8) + np.arange(s, dtype=np.uint64), 0.0, 20e3)
    return Structure(xy, en, et, eai, cn, cd, np.zeros(len(cn)), ln, [DX] * s, lv, ee, eq,
                     np.tile([0.0, -1.0, 0.0], (len(ee), 1)))


def c3_batch(first, count, pinned=None):
    """Structures [first, first + count) of config C3 as ONE packed batch (the dict obgpu.pack_batch returns), generated
    with array operations — the same structures, bit for bit, as [c3_frame(i) for i in range(first, first + count)],
    in that order.  ``pinned``: an allocator ``(shape, dtype) -> ndarray`` (obgpu.pinned_empty) for the large arrays."""
    alloc = pinned or (lambda shape, dt: np.empty(shape, dt))
    idx = np.arange(first, first + count, dtype=np.uint64)
    b = np.zeros(count, np.int64)
    s = np.zeros(count, np.int64)
    todo = np.ones(count, bool)
    t = 0
    while todo.any():
        bb = 1 + (uniform01(30 + t, idx[todo]) * 5).astype(np.int64)
        ss = 1 + (uniform01(60 + t, idx[todo]) * 6).astype(np.int64)
        ok = (ss * (2 * bb + 1) >= 10) & (ss * (2 * bb + 1) <= 60)
        w = np.nonzero(todo)[0]
        b[w[ok]], s[w[ok]] = bb[ok], ss[ok]
        todo[w[ok]] = False
        t += 1
    nN, nE = (b + 1) * (s + 1), s * (2 * b + 1)
    nC, nNL, nEL = 3 * (b + 1), s, s * b

    def off(c):
        o = np.zeros(count + 1, np.int64)
        np.cumsum(c, out=o[1:])
        return o

    node_off, elem_off, con_off, nload_off, eload_off = off(nN), off(nE), off(nC), off(nNL), off(nEL)
    Nt, Et, Ct, NLt, ELt = (int(o[-1]) for o in (node_off, elem_off, con_off, nload_off, eload_off))
    node_xy = alloc((Nt, 2), np.float64)
    elem_nodes = alloc((Et, 2), np.uint32)
    elem_type = alloc((Et,), np.uint8)
    elem_EAI = alloc((Et, 3), np.float64)
    con_node, con_dof, con_val = alloc((Ct,), np.uint32), alloc((Ct,), np.uint8), alloc((Ct,), np.float64)
    nload_node, nload_dof, nload_val = alloc((NLt,), np.uint32), alloc((NLt,), np.uint8), alloc((NLt,), np.float64)
    eload_elem, eload_q, eload_dir = alloc((ELt,), np.uint32), alloc((ELt,), np.float64), alloc((ELt, 3), np.float64)
    con_val[:] = 0.0
    nload_dof[:] = DX
    eload_dir[:] = (0.0, -1.0, 0.0)
    csec = IPE[(uniform01(102, idx) * len(IPE)).astype(np.int64)]
    bsec = IPE[(uniform01(103, idx) * len(IPE)).astype(np.int64)]
    for bv in range(1, 6):
        for sv in range(1, 7):
            m = np.nonzero((b == bv) & (s == sv))[0]
            if len(m) == 0:
                continue
            i64 = idx[m]
            k = len(m)
            W = uniform(100, (i64 * np.uint64(8))[:, None] + np.arange(bv, dtype=np.uint64)[None, :], 3.0, 6.0)
            H = uniform(101, (i64 * np.uint64(8))[:, None] + np.arange(sv, dtype=np.uint64)[None, :], 2.5, 4.0)
            xs = np.concatenate([np.zeros((k, 1)), np.cumsum(W, 1)], 1)            # [k, b+1]
            ys = np.concatenate([np.zeros((k, 1)), np.cumsum(H, 1)], 1)            # [k, s+1]
            nn = (bv + 1) * (sv + 1)
            rows = node_off[m][:, None] + np.arange(nn)[None, :]
            jj, ii = np.divmod(np.arange(nn), bv + 1)
            node_xy[rows, 0] = xs[:, ii]
            node_xy[rows, 1] = ys[:, jj]
            # elements: per storey j = 1..s the b+1 columns then the b beams
            en = np.zeros((sv * (2 * bv + 1), 2), np.uint32)
            is_beam = np.zeros(sv * (2 * bv + 1), bool)
            key = np.zeros(sv * (2 * bv + 1), np.uint64)       # j*8 + i of a beam
            e = 0
            for j in range(1, sv + 1):
                for i in range(bv + 1):
                    en[e] = ((j - 1) * (bv + 1) + i, j * (bv + 1) + i); e += 1
                for i in range(bv):
                    en[e] = (j * (bv + 1) + i, j * (bv + 1) + i + 1); is_beam[e] = True; key[e] = j * 8 + i; e += 1
            erows = elem_off[m][:, None] + np.arange(len(en))[None, :]
            elem_nodes[erows] = en[None, :, :]
            beams = np.nonzero(is_beam)[0]
            u = uniform01(104, (i64 * np.uint64(64))[:, None] + key[beams][None, :])
            ty = np.where(u < 0.6, RR, np.where(u < 0.75, RA, np.where(u < 0.9, AR, AA))).astype(np.uint8)
            et = np.full((k, len(en)), RR, np.uint8)
            et[:, beams] = ty
            elem_type[erows] = et
            eai = np.empty((k, len(en), 3))
            eai[:, :, 0] = E_STEEL
            eai[:, :, 1:] = csec[m][:, None, :]
            eai[:, beams, 1:] = bsec[m][:, None, :]
            elem_EAI[erows] = eai
            crows = con_off[m][:, None] + np.arange(3 * (bv + 1))[None, :]
            con_node[crows] = np.repeat(np.arange(bv + 1), 3)[None, :]
            con_dof[crows] = np.tile([DX, DY, RZ], bv + 1)[None, :]
            lrows = nload_off[m][:, None] + np.arange(sv)[None, :]
            nload_node[lrows] = (np.arange(1, sv + 1) * (bv + 1))[None, :]
            nload_val[lrows] = uniform(106, (i64 * np.uint64(8))[:, None] + np.arange(sv, dtype=np.uint64)[None, :], 0.0, 20e3)
            qrows = eload_off[m][:, None] + np.arange(len(beams))[None, :]
            eload_elem[qrows] = beams[None, :]
            eload_q[qrows] = uniform(105, (i64 * np.uint64(64))[:, None] + key[beams][None, :], 5e3, 30e3)
    return dict(S=count, node_off=node_off, elem_off=elem_off, con_off=con_off, nload_off=nload_off, eload_off=eload_off,
                node_xy=node_xy.reshape(-1), elem_nodes=elem_nodes.reshape(-1), elem_type=elem_type,
                elem_EAI=elem_EAI.reshape(-1), con_node=con_node, con_dof=con_dof, con_val=con_val, nload_node=nload_node,
                nload_dof=nload_dof, nload_val=nload_val, eload_elem=eload_elem, eload_q=eload_q,
                eload_dir=eload_dir.reshape(-1), node_rotz=None, eload_type=None, eload_q2=None)


def grid_frame(nx, ny, etype=RR, seed_index=0, section=(E_STEEL, 28.5e-4, 1940e-8), spacing=1.0) -> Structure:
    """Configs C4/C5: nx x ny node grid, horizontal + vertical members of ``etype``, bottom row fixed,
    UDL 9810 N/m down on horizontal members, lateral nodal loads N(0, 5e3) (Box-Muller on the counter RNG)."""
    jj, ii = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    nid = (jj * nx + ii).astype(np.uint32)
    xy = np.stack([ii.ravel() * spacing, jj.ravel() * spacing], 1).astype(np.float64)
    hor = np.stack([nid[:, :-1].ravel(), nid[:, 1:].ravel()], 1)
    ver = np.stack([nid[:-1, :].ravel(), nid[1:, :].ravel()], 1)
    # element order: row by row, horizontals of row j then verticals from row j to j+1
    rows = []
    nh, nv = nx - 1, nx
    for j in range(ny):
        rows.append(hor[j * nh:(j + 1) * nh])
        if j < ny - 1:
            rows.append(ver[j * nv:(j + 1) * nv])
    en = np.concatenate(rows).astype(np.uint32)
    is_h = (en[:, 1] - en[:, 0]) == 1
    E = len(en)
    eai = np.tile(np.asarray(section, np.float64), (E, 1))
    ee = np.nonzero(is_h & (en[:, 0] >= nx))[0].astype(np.uint32)  # bottom-row members are fully fixed: skip
    base = np.arange(nx, dtype=np.uint32)
    cn = np.repeat(base, 3)
    cd = np.tile(np.array([DX, DY, RZ], np.uint8), nx)
    ln = np.arange(nx, nx * ny, dtype=np.uint32)
    k = np.arange(len(ln), dtype=np.uint64)
    u1 = uniform01(200 + 2 * seed_index, k)
    u2 = uniform01(201 + 2 * seed_index, k)
    lv = 5e3 * np.sqrt(-2.0 * np.log(1.0 - u1)) * np.cos(2 * np.pi * u2)
    return Structure(xy, en, np.full(E, etype, np.uint8), eai, cn, cd, np.zeros(len(cn)), ln,
                     np.full(len(ln), DX, np.uint8), lv, ee, np.full(len(ee), 9810.0),
                     np.tile([0.0, -1.0, 0.0], (len(ee), 1)))
