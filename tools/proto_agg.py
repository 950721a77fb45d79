"""CPU prototype (scipy) of strength-of-connection + MIS-2 aggregation with a filtered smoothed prolongator, used to choose the
parameters of the device implementation (csrc/mg_aggregate.cu) and its twin (oracle/mg_ref.c).  Design exploration only."""
import sys, time
import numpy as np
import scipy.sparse as sp
sys.path.insert(0, "/root/repo")
from conjugategradient_b200 import generators as g


def hash32(i):
    x = (i.astype(np.uint64) + 1) * np.uint64(0x9E3779B97F4A7C15) & np.uint64(0xFFFFFFFFFFFFFFFF)
    x ^= x >> np.uint64(32)
    x = x * np.uint64(0xD6E8FEB86659FD93) & np.uint64(0xFFFFFFFFFFFFFFFF)
    x ^= x >> np.uint64(32)
    return (x & np.uint64(0xFFFFFFFF)).astype(np.uint64)


def strength(A, theta):
    A = A.tocsr()
    n = A.shape[0]
    off = A - sp.diags(A.diagonal())
    off = off.tocsr(); off.eliminate_zeros()
    absoff = abs(off)
    rowmax = absoff.max(axis=1).toarray().ravel()
    coo = absoff.tocoo()
    thr = theta * np.minimum(rowmax[coo.row], rowmax[coo.col])
    keep = coo.data >= thr
    S = sp.csr_matrix((coo.data[keep], (coo.row[keep], coo.col[keep])), shape=(n, n))
    return S


def mis2(S):
    n = S.shape[0]
    idx = np.arange(n, dtype=np.uint64)
    key = (hash32(idx) << np.uint64(32)) | idx      # unique
    state = np.ones(n, dtype=np.uint64)              # 1 undecided, 2 in, 0 out
    Sp = S.indptr; Sj = S.indices
    rows = np.repeat(np.arange(n), np.diff(Sp))
    rounds = 0
    while (state == 1).any():
        rounds += 1
        # tuple -> single comparable: state in top bits.  key has 64 bits; use two arrays (state, key) lexicographic
        t_s, t_k = state.copy(), key.copy()
        for hop in range(2):
            ns, nk = t_s[Sj], t_k[Sj]
            # per-row max of (s,k) including self
            best_s, best_k = t_s.copy(), t_k.copy()
            # combine: sort-free reduce using np.maximum.at on a combined ordering (state major)
            comb = (ns.astype(np.float64) * 0)  # placeholder
            # lexicographic max: first max state, then max key among those
            ms = best_s.copy()
            np.maximum.at(ms, rows, ns)
            cand_self = np.where(best_s == ms, best_k, np.uint64(0))
            mk = cand_self.copy()
            nk_masked = np.where(ns == ms[rows], nk, np.uint64(0))
            np.maximum.at(mk, rows, nk_masked)
            t_s, t_k = ms, mk
        und = state == 1
        win = und & (t_k == key) & (t_s == 1)
        lose = und & (t_s == 2)
        state[win] = 2
        state[lose] = 0
    return state == 2, rounds


def aggregate(A, S, roots):
    n = S.shape[0]
    agg = -np.ones(n, dtype=np.int64)
    rid = np.cumsum(roots) - 1
    agg[roots] = rid[roots]
    Sp, Sj, Sv = S.indptr, S.indices, S.data
    rows = np.repeat(np.arange(n), np.diff(Sp))
    for p in range(2):
        src = agg.copy()
        # each unaggregated node: strongest strong neighbour that is aggregated in src
        ok = (src[Sj] >= 0) & (src[rows] < 0)
        best = np.zeros(n)
        np.maximum.at(best, rows[ok], Sv[ok])
        hit = ok & (Sv == best[rows])
        # first in storage order
        first = np.full(n, np.iinfo(np.int64).max)
        pos = np.arange(len(Sj))
        np.minimum.at(first, rows[hit], pos[hit])
        m = first < np.iinfo(np.int64).max
        agg[m] = src[Sj[first[m]]]
    # leftovers -> own aggregates
    left = agg < 0
    nc = int(rid[-1]) + 1 if n else 0
    agg[left] = nc + np.arange(left.sum())
    nc += int(left.sum())
    return agg, nc


def build_P(A, S, agg, nc, omega, filtered):
    n = A.shape[0]
    T = sp.csr_matrix((np.ones(n), (np.arange(n), agg)), shape=(n, nc))
    if omega == 0:
        return T
    if filtered:
        # A_F: strong off-diagonals kept; weak ones added to the diagonal
        Soff = S.copy(); Soff.data[:] = 1
        off = A - sp.diags(A.diagonal())
        strong = off.multiply(Soff)
        weak = off - strong
        dF = A.diagonal() + np.asarray(weak.sum(axis=1)).ravel()
        AF = strong + sp.diags(dF)
        D = dF
    else:
        AF = A; D = A.diagonal()
    return (T - omega * sp.diags(1.0 / D) @ AF @ T).tocsr()


class MG:
    def __init__(self, A, theta=0.25, omega_p=2 / 3, filtered=True, coarse=500, max_levels=10, mode="mis2", verbose=True):
        self.levels = []
        t0 = time.time()
        while True:
            n = A.shape[0]
            L = {"A": A.tocsr(), "dinv": 1.0 / A.diagonal()}
            self.levels.append(L)
            if n <= coarse or len(self.levels) >= max_levels:
                break
            if mode == "mis2":
                S = strength(A, theta)
                roots, rounds = mis2(S)
                agg, nc = aggregate(A, S, roots)
            else:
                S = strength(A, 0.0)
                agg = np.arange(n) // 8; nc = (n + 7) // 8; rounds = 0
            P = build_P(A, S, agg, nc, omega_p, filtered and mode == "mis2")
            L["P"] = P; L["R"] = P.T.tocsr()
            Ac = (L["R"] @ (A @ P)).tocsr()
            if verbose:
                print(f"  level {len(self.levels)-1}: n={n} nnz/row={A.nnz/n:.1f} strong/row={S.nnz/n:.1f} rounds={rounds} nc={nc} ratio={n/nc:.1f} "
                      f"P/row={P.nnz/n:.2f} -> coarse nnz/row={Ac.nnz/nc:.1f}", flush=True)
            A = Ac
        self.Ainv = np.linalg.inv(self.levels[-1]["A"].toarray())
        self.opc = sum(l["A"].nnz for l in self.levels) / self.levels[0]["A"].nnz
        if verbose:
            print(f"  levels={len(self.levels)} op complexity={self.opc:.3f} setup {time.time()-t0:.1f}s", flush=True)

    def apply(self, r, omega=0.85, l=0):
        L = self.levels[l]
        if l == len(self.levels) - 1:
            return self.Ainv @ r
        A = L["A"]
        x = omega * L["dinv"] * r
        res = r - A @ x
        xc = self.apply(L["R"] @ res, omega, l + 1)
        x = x + L["P"] @ xc
        x = x + omega * L["dinv"] * (r - A @ x)
        return x


def pcg(A, b, M, tol=1e-8, maxit=500):
    x = np.zeros_like(b); r = b.copy(); rr0 = r @ r
    z = M(r); p = z.copy(); rz = r @ z
    for it in range(maxit):
        Ap = A @ p; a = rz / (p @ Ap)
        x += a * p; r -= a * Ap
        if (r @ r) / rr0 < tol * tol:
            return x, it
        z = M(r); rzn = r @ z; p = z + (rzn / rz) * p; rz = rzn
    return x, maxit


if __name__ == "__main__":
    ns = int(sys.argv[1]); re = float(sys.argv[2])
    v, rp, c = g.mps_particles(ns, re_over_l0=re)
    n = len(rp) - 1
    A = sp.csr_matrix((v, c, rp), shape=(n, n))
    b = g.rhs_uniform(n, 2)
    print(f"MPS {ns}^3 re={re}: n={n} nnz/row={A.nnz/n:.1f}")
    for cfg in sys.argv[3:]:
        mode, theta, om, filt = cfg.split(",")
        print(f"config mode={mode} theta={theta} omega_p={om} filtered={filt}")
        H = MG(A, theta=float(theta), omega_p=float(om), filtered=bool(int(filt)), mode=mode)
        x, it = pcg(A, b, lambda r: H.apply(r))
        print(f"  -> PCG iterations {it}  true rel res {np.linalg.norm(b - A@x)/np.linalg.norm(b):.2e}", flush=True)
