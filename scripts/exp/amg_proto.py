"""Prototype (scipy, CPU): rigid-body-mode aggregation multigrid as a PCG preconditioner for the C4 grid-frame family.
Decides the design of the GPU preconditioner (iterations vs Jacobi).  Not part of the product."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, scipy.sparse as sp, scipy.sparse.linalg as spla
from openbeam_b200 import workloads
from oracle.pyoracle import OracleProblem

def build(n):
    st = workloads.grid_frame(n, n)
    o = OracleProblem(st).assembleProblem(skip_subK=True)
    outer, inner, val = o.K_csc_upper()
    N = len(outer) - 1
    Ku = sp.csc_matrix((val, inner, outer), shape=(N, N))
    K = (Ku + sp.triu(Ku, 1).T).tocsr()
    f = np.asarray(o.free_dof_indices)
    A = K[f][:, f].tocsr()
    dn = np.asarray(o.geti(0))[f]; dk = np.asarray(o.geti(1))[f]
    kind = np.where(dk == 5, 2, dk)
    xy = st.node_xy[dn]
    return A, np.asarray(o.F_f), dn, kind, xy

def aggregate_greedy(adj_indptr, adj_idx, nn):
    agg = -np.ones(nn, np.int64); na = 0
    for i in range(nn):                     # phase 1
        if agg[i] >= 0: continue
        nb = adj_idx[adj_indptr[i]:adj_indptr[i+1]]
        if np.all(agg[nb] < 0):
            agg[i] = na; agg[nb] = na; na += 1
    for i in range(nn):                     # phase 2: attach leftovers to a neighbouring aggregate
        if agg[i] >= 0: continue
        nb = adj_idx[adj_indptr[i]:adj_indptr[i+1]]
        c = agg[nb][agg[nb] >= 0]
        if len(c): agg[i] = -2 - c[0]
    left = agg <= -2
    agg[left] = -2 - agg[left]
    for i in range(nn):
        if agg[i] < 0: agg[i] = na; na += 1
    return agg, na

def aggregate_box(xy, h):
    key = np.floor(xy / h + 1e-9).astype(np.int64)
    key -= key.min(0)
    k = key[:, 1] * (key[:, 0].max() + 1) + key[:, 0]
    u, inv = np.unique(k, return_inverse=True)
    return inv, len(u)

def level_setup(A, node, kind, xy_node, mode, lvl):
    """A scalar CSR over dofs; node[i] node id of dof i (0..nn-1), kind in {0,1,2}; xy_node [nn,2]."""
    nn = xy_node.shape[0]
    G = sp.csr_matrix((np.ones(A.nnz), (node[A.nonzero()[0]], node[A.nonzero()[1]])), shape=(nn, nn))
    G.setdiag(0); G.eliminate_zeros()
    if mode[0] == 'greedy':
        agg, na = aggregate_greedy(G.indptr, G.indices, nn)
    else:
        agg, na = aggregate_box(xy_node, mode[1] * (mode[2] ** lvl))
    cnt = np.bincount(agg, minlength=na)
    cen = np.stack([np.bincount(agg, xy_node[:, 0], na) / cnt, np.bincount(agg, xy_node[:, 1], na) / cnt], 1)
    a = agg[node]
    dx = xy_node[node, 0] - cen[a, 0]; dy = xy_node[node, 1] - cen[a, 1]
    n = A.shape[0]
    rows, cols, vals = [], [], []
    i = np.arange(n)
    m0 = kind == 0; m1 = kind == 1; m2 = kind == 2
    rows += [i[m0], i[m0]]; cols += [3*a[m0], 3*a[m0]+2]; vals += [np.ones(m0.sum()), -dy[m0]]
    rows += [i[m1], i[m1]]; cols += [3*a[m1]+1, 3*a[m1]+2]; vals += [np.ones(m1.sum()), dx[m1]]
    rows += [i[m2]]; cols += [3*a[m2]+2]; vals += [np.ones(m2.sum())]
    P = sp.csr_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(n, 3*na))
    return P, cen, na

def gersh(A, d):
    return float((abs(A).sum(1).A1 / d).max()) if hasattr(abs(A).sum(1), 'A1') else float((np.asarray(abs(A).sum(1)).ravel() / d).max())

class MG:
    def __init__(self, A, node, kind, xy_node, mode, smooth_P=False, min_coarse=300, deg=2, alpha=8.0, eig='power'):
        self.lv = []
        lvl = 0
        while True:
            d = A.diagonal().copy(); d[d == 0] = 1.0
            # lambda_max(D^-1 A)
            if eig == 'gersh':
                lam = float((np.asarray(abs(A).sum(1)).ravel() / d).max())
            else:
                v = np.random.default_rng(0).standard_normal(A.shape[0])
                for _ in range(15):
                    v = (A @ v) / d; lam = np.linalg.norm(v); v /= lam
                lam *= 1.1
            L = dict(A=A, dinv=1.0/d, lam=lam)
            self.lv.append(L)
            if A.shape[0] <= min_coarse: break
            P, cen, na = level_setup(A, node, kind, xy_node, mode, lvl)
            if smooth_P:
                P = P - (4.0/(3.0*lam)) * sp.diags(1.0/d) @ (A @ P)
            L['P'] = P.tocsr(); L['R'] = P.T.tocsr()
            A = (L['R'] @ A @ P).tocsr()
            node = np.repeat(np.arange(na), 3); kind = np.tile([0, 1, 2], na); xy_node = cen
            lvl += 1
        Ac = self.lv[-1]['A'].toarray()
        self.coarse = np.linalg.pinv(Ac, hermitian=True)
        self.deg = deg; self.alpha = alpha
        print("levels:", [l['A'].shape[0] for l in self.lv], "nnz:", [l['A'].nnz for l in self.lv],
              "opcx=%.2f" % (sum(l['A'].nnz for l in self.lv) / self.lv[0]['A'].nnz))
    def cheb(self, L, b, x):
        # Chebyshev smoother of degree deg on [lam/alpha, lam] for D^-1 A
        lmax, lmin = L['lam'], L['lam'] / self.alpha
        theta, delta = 0.5*(lmax+lmin), 0.5*(lmax-lmin)
        sigma = theta/delta; rho = 1.0/sigma
        A, dinv = L['A'], L['dinv']
        r = b - A @ x if x is not None else b.copy()
        if x is None: x = np.zeros_like(b)
        dvec = dinv * r / theta
        x = x + dvec
        for _ in range(self.deg - 1):
            r = r - A @ dvec
            rho_n = 1.0/(2*sigma - rho)
            dvec = rho_n*rho*dvec + (2*rho_n/delta) * (dinv * r)
            rho = rho_n
            x = x + dvec
        return x
    def vcycle(self, l, b):
        L = self.lv[l]
        if l == len(self.lv) - 1:
            return self.coarse @ b
        x = self.cheb(L, b, None)
        r = b - L['A'] @ x
        xc = self.vcycle(l+1, L['R'] @ r)
        x = x + L['P'] @ xc
        x = self.cheb(L, b, x)
        return x
    def __call__(self, b):
        return self.vcycle(0, b)

def pcg(A, b, M, rtol, maxit=20000):
    x = np.zeros_like(b); r = b.copy(); z = M(r); p = z.copy(); rz = r @ z; bb = b @ b
    for it in range(1, maxit+1):
        Ap = A @ p; a = rz / (p @ Ap); x += a*p; r -= a*Ap
        if r @ r <= rtol*rtol*bb: return x, it
        z = M(r); rzn = r @ z; p = z + (rzn/rz)*p; rz = rzn
    return x, maxit

if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 120
    A, b, dn, kind, xyd = build(n)
    print("n=%d dofs=%d nnz=%d" % (n, A.shape[0], A.nnz))
    un, node = np.unique(dn, return_inverse=True)
    xy_node = np.zeros((len(un), 2)); xy_node[node] = xyd
    t0 = time.time(); xref = spla.splu(A.tocsc()).solve(b); print("splu %.1fs" % (time.time()-t0))
    d = A.diagonal()
    if '--jacobi' in sys.argv:
        x, it = pcg(A, b, lambda r: r/d, 1e-10); print("jacobi it", it, "err", abs(x-xref).max()/abs(xref).max())
    for name, kw in [
        ("box2 deg2", dict(mode=('box', 2.0, 2.0), deg=2)),
        ("box3 deg2", dict(mode=('box', 3.0, 3.0), deg=2)),
        ("box3 deg3", dict(mode=('box', 3.0, 3.0), deg=3)),
        ("box2 deg3", dict(mode=('box', 2.0, 2.0), deg=3)),
        ("greedy deg2", dict(mode=('greedy',), deg=2)),
        ("box2 deg2 SA", dict(mode=('box', 2.0, 2.0), deg=2, smooth_P=True)),
        ("box3 deg2 SA", dict(mode=('box', 3.0, 3.0), deg=2, smooth_P=True)),
    ]:
        t0 = time.time(); M = MG(A, node, kind, xy_node, **kw); ts = time.time()-t0
        for rtol in (1e-10,):
            t0 = time.time(); x, it = pcg(A, b, M, rtol)
            print("%-14s rtol %.0e: it=%d err=%.2e (setup %.1fs solve %.1fs)" % (name, rtol, it, abs(x-xref).max()/abs(xref).max(), ts, time.time()-t0))
