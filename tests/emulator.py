"""numpy emulation of the device numeric phase, driven by the integer plan (test infrastructure).

Mirrors the CUDA algorithm step by step (uno_b200/csrc/kernels.cu): assembly through the
scatter map, extend-add through the front row lists, blocked Bunch-Kaufman with the pivot search
confined to NBI x NBI diagonal blocks, null-pivot dropping, forward/diagonal/backward solves.
It validates the plan and the pivoting strategy on the CPU; it is NOT bit-exact with the GPU
(different summation order in the Schur updates) and it is far too slow for real sizes.
"""
from __future__ import annotations

import numpy as np

NBI = 32
ALPHA = (1.0 + np.sqrt(17.0)) / 8.0


def equilibrate(kkt, values, sweeps=3):
    """Power-of-two Ruiz sweeps, same rule as k_scale_sweep."""
    n = kkt.dim
    r = kkt.row.astype(np.int64)
    c = kkt.col.astype(np.int64)
    a = np.abs(values)
    sc = np.ones(n)
    for _ in range(sweeps):
        m = np.zeros(n)
        np.maximum.at(m, r, a * sc[c])
        np.maximum.at(m, c, a * sc[r])
        m *= sc
        ok = (m > 0) & (m < 1e300)
        e = np.zeros(n, dtype=np.int64)
        _, ex = np.frexp(m[ok])
        e[ok] = ex - 1
        sh = -((e + 1) >> 1)
        sc = np.where(ok, np.ldexp(sc, sh), sc)
    return sc


def bk_block(B, n, tol):
    """Bunch-Kaufman on the leading n x n of the full symmetric block B (modified in place).
    Returns perm, d, e, ptype exactly like bk_factor_block."""
    p = np.arange(B.shape[0])
    d = np.ones(B.shape[0])
    e = np.zeros(B.shape[0])
    ty = np.zeros(B.shape[0], dtype=np.int32)
    j = 0
    while j < n:
        absajj = abs(B[j, j])
        if j + 1 < n:
            colv = np.abs(B[j + 1 : n, j])
            r = j + 1 + int(np.argmax(colv))
            lam = colv[r - j - 1]
        else:
            r, lam = j, 0.0
        kp, step, isnull = j, 1, False
        if not (max(absajj, lam) > tol):
            isnull = True
        elif absajj >= ALPHA * lam:
            pass
        else:
            mask = np.ones(n, dtype=bool)
            mask[:j] = False
            mask[r] = False
            sigma = np.max(np.abs(B[:n, r][mask]))
            if absajj * sigma >= ALPHA * lam * lam:
                pass
            elif abs(B[r, r]) >= ALPHA * sigma:
                kp = r
            else:
                kp, step = r, 2
        kk = j + step - 1
        if kp != kk:
            B[[kk, kp], :] = B[[kp, kk], :]
            B[:, [kk, kp]] = B[:, [kp, kk]]
            p[[kk, kp]] = p[[kp, kk]]
        if step == 1:
            dj = B[j, j]
            if isnull:
                B[j + 1 :, j] = 0.0
                d[j], e[j], ty[j] = 1.0, 0.0, 3
            else:
                l = B[j + 1 : n, j] / dj
                B[j + 1 : n, j + 1 : n] -= np.outer(l, B[j + 1 : n, j])
                B[j + 1 : n, j] = l
                d[j], e[j], ty[j] = dj, 0.0, 0
        else:
            a, b, c = B[j, j], B[j + 1, j], B[j + 1, j + 1]
            det = a * c - b * b
            w1 = B[j + 2 : n, j].copy()
            w2 = B[j + 2 : n, j + 1].copy()
            l1 = (c * w1 - b * w2) / det
            l2 = (a * w2 - b * w1) / det
            B[j + 2 : n, j + 2 : n] -= np.outer(l1, w1) + np.outer(l2, w2)
            B[j + 2 : n, j] = l1
            B[j + 2 : n, j + 1] = l2
            B[j + 1, j] = 0.0
            d[j], d[j + 1], e[j], e[j + 1] = a, c, b, 0.0
            ty[j], ty[j + 1] = 1, 2
        j += step
    return p, d, e, ty


class EmulatedFactorization:
    def __init__(self, plan, perm, scatter, kkt, values, sweeps=3, null_tol=1e-14, defer=False, delay_tol=1e-12, growth_limit=0.0):
        self.plan = plan
        self.perm = np.asarray(perm)
        self.n = kkt.dim
        self.null_tol = null_tol
        self.delay_tol = delay_tol
        self.growth_limit = growth_limit
        self.delayed_size = []       # largest Schur entry outside the block of each null-flagged column
        self.delayed = []            # plan columns flagged for a delay (kernels.cu: flag_delay)
        self.scale = equilibrate(kkt, values, sweeps)
        nsn = plan["num_fronts"]
        pool = np.zeros(plan["pool_doubles"])
        r = kkt.row.astype(np.int64)
        c = kkt.col.astype(np.int64)
        np.add.at(pool, scatter, values * self.scale[r] * self.scale[c])
        self.assembled = pool.copy()
        self.dd = np.ones(self.n)
        self.de = np.zeros(self.n)
        self.ptype = np.zeros(self.n, dtype=np.int32)
        self.lperm = np.zeros(self.n, dtype=np.int64)
        self.fronts = [None] * nsn
        children = [[] for _ in range(nsn)]
        for s in range(nsn):
            if plan["parent"][s] >= 0:
                children[plan["parent"][s]].append(s)
        self.children = children
        self.relmap = [None] * nsn
        if not defer:
            for s in range(nsn):
                self.factor_front(s)

    def child_map(self, ch):
        """Local index in the parent front of every boundary row of front ch (the extend-add map)."""
        plan = self.plan
        k, rows_off, rows, first = plan["k"], plan["rows_off"], plan["rows"], plan["first_col"]
        s = int(plan["parent"][ch])
        ks, kc = int(k[s]), int(k[ch])
        crow = rows[rows_off[ch] + kc : rows_off[ch + 1]]
        prow = rows[rows_off[s] : rows_off[s + 1]]
        loc = np.searchsorted(prow[ks:], crow) + ks
        inpiv = crow < first[s] + ks
        loc = np.where(inpiv, crow - first[s], loc)
        assert np.all(prow[loc] == crow)
        return loc

    def factor_front(self, s):
        """Assemble front s (its children's contribution blocks must be in self.fronts) and factor its pivot columns."""
        plan, pool, null_tol = self.plan, self.assembled, self.null_tol
        k, f, ld, off, first = plan["k"], plan["f"], plan["ld"], plan["pool_offset"], plan["first_col"]
        if True:
            fs, ks = int(f[s]), int(k[s])
            F = np.zeros((fs, fs))
            F[:, :ks] = pool[off[s] : off[s] + ld[s] * ks].reshape(ks, ld[s]).T[:fs, :]   # ld x k column-major factor panel
            F = np.tril(F) + np.tril(F, -1).T
            for ch in self.children[s]:
                Fc = self.fronts[ch]
                loc = self.child_map(ch)
                self.relmap[ch] = loc
                F[np.ix_(loc, loc)] += Fc["cb"]
            # blocked restricted Bunch-Kaufman on the first ks columns
            lp = np.arange(ks)
            for j0 in range(0, ks, NBI):
                nbw = min(NBI, ks - j0)
                j1 = j0 + nbw
                B = F[j0:j1, j0:j1].copy()
                p, d, e, ty = bk_block(B, nbw, null_tol)
                L11 = np.tril(B, -1) + np.eye(nbw)
                q = first[s] + j0
                self.dd[q : q + nbw], self.de[q : q + nbw], self.ptype[q : q + nbw] = d, e, ty
                lp[j0:j1] = lp[j0:j1][p]
                F[j0:j1, :j0] = F[j0:j1, :j0][p, :]
                X = F[j1:, j0:j1][:, p]
                Wm = np.linalg.solve(L11, X.T).T      # X L11^{-T}
                D = np.diag(d) + np.diag(e[:-1], -1) + np.diag(e[:-1], 1)
                Lm = np.linalg.solve(D, Wm.T).T
                null = ty == 3
                # a null pivot whose column still has entries outside the block is not null: the column has to leave the front
                if Wm.shape[0] > 0:
                    for c in np.flatnonzero(null):
                        if np.max(np.abs(Wm[:, c])) > self.delay_tol:
                            self.delayed.append(int(first[s] + lp[j0 + c]))
                            self.delayed_size.append(float(np.max(np.abs(Wm[:, c]))))
                    if self.growth_limit > 0:
                        for c in np.flatnonzero(~null):
                            if np.max(np.abs(Lm[:, c])) > self.growth_limit:
                                self.delayed.append(int(first[s] + lp[j0 + c]))
                Lm[:, null] = 0.0
                Wm[:, null] = 0.0
                F[j0:j1, j0:j1] = np.tril(B, -1) + np.diag(d)
                F[j1:, j0:j1] = Lm
                F[j1:, j1:] -= Wm @ Lm.T
            self.lperm[first[s] : first[s] + ks] = lp
            self.fronts[s] = {"L": F[:, :ks].copy(), "cb": F[ks:, ks:].copy()}

    def inertia(self):
        pos = neg = zero = 0
        q = 0
        while q < self.n:
            t = self.ptype[q]
            if t == 0:
                if self.dd[q] > 0: pos += 1
                else: neg += 1
                q += 1
            elif t == 1:
                a, b, c = self.dd[q], self.de[q], self.dd[q + 1]
                det = a * c - b * b
                if det < 0: pos += 1; neg += 1
                elif a > 0: pos += 2
                else: neg += 2
                q += 2
            else:
                zero += 1
                q += 1
        return pos, neg, zero

    def forward_front(self, s, zg, z2, cvec):
        plan = self.plan
        k, first = plan["k"], plan["first_col"]
        ks = int(k[s])
        L = self.fronts[s]["L"]
        v = np.zeros(L.shape[0])
        v[:ks] = zg[first[s] : first[s] + ks]
        for ch in self.children[s]:
            v[self.child_map(ch)] += cvec[ch]
        y = v[:ks][self.lperm[first[s] : first[s] + ks]]
        L11 = np.tril(L[:ks, :], -1) + np.eye(ks)
        y = np.linalg.solve(L11, y)
        cvec[s] = v[ks:] - L[ks:, :] @ y
        z2[first[s] : first[s] + ks] = y

    def backward_front(self, s, z2, xg):
        plan = self.plan
        k, first = plan["k"], plan["first_col"]
        rows_off, rows = plan["rows_off"], plan["rows"]
        ks = int(k[s])
        L = self.fronts[s]["L"]
        y = z2[first[s] : first[s] + ks].copy()
        q0 = first[s]
        j = 0
        while j < ks:
            t = self.ptype[q0 + j]
            if t == 0:
                y[j] /= self.dd[q0 + j]; j += 1
            elif t == 1:
                a, bb, c = self.dd[q0 + j], self.de[q0 + j], self.dd[q0 + j + 1]
                det = a * c - bb * bb
                y1, y2 = y[j], y[j + 1]
                y[j], y[j + 1] = (c * y1 - bb * y2) / det, (a * y2 - bb * y1) / det
                j += 2
            else:
                y[j] = 0.0; j += 1
        xb = xg[rows[rows_off[s] + ks : rows_off[s + 1]]]
        y -= L[ks:, :].T @ xb
        L11 = np.tril(L[:ks, :], -1) + np.eye(ks)
        y = np.linalg.solve(L11.T, y)
        xg[first[s] + self.lperm[first[s] : first[s] + ks]] = y

    def solve(self, b):
        nsn = self.plan["num_fronts"]
        zg = (self.scale * b)[self.perm]
        z2 = np.zeros(self.n)
        cvec = [None] * nsn
        for s in range(nsn):
            self.forward_front(s, zg, z2, cvec)
        xg = np.zeros(self.n)
        for s in range(nsn - 1, -1, -1):
            self.backward_front(s, z2, xg)
        x = np.zeros(self.n)
        x[self.perm] = xg
        return x * self.scale


def factor_with_delays(solver, kkt, values, max_replans=8, **kw):
    """The library's run_factor loop (solver.cpp) on the emulation: factorize, move the flagged columns into ancestor fronts
    through the library's own re-plan (b200kkt_debug_delay_columns), factorize again.  Returns (emulation, number of re-plans)."""
    replans = 0
    while True:
        perm, scat = solver.permutation()
        emu = EmulatedFactorization(solver.plan(), perm, scat, kkt, values, **kw)
        plan = emu.plan
        sn_of = np.searchsorted(plan["first_col"], np.asarray(emu.delayed, dtype=np.int64), side="right") - 1
        cols = [q for q, s in zip(emu.delayed, sn_of) if plan["parent"][s] >= 0]
        if not cols or replans >= max_replans:
            return emu, replans
        solver.debug_delay_columns(sorted(set(cols)))
        replans += 1


class PartitionedEmulation(EmulatedFactorization):
    """The same numeric phase run the way the library runs it on several GPUs (solver.cpp: enqueue_factor / run_solve with
    a subtree partition): every rank walks the tree levels in ascending order over ITS fronts, the cut fronts of a level hand
    their contribution block (factorization) or boundary partial sums (forward solve) to the parent's rank, the backward sweep
    ships the boundary solution the other way, the solution is put together by a sum over ranks of the owned entries.
    `transport.exchange(sends, recvs)` moves arrays between ranks (sends: [(peer, array)], recvs: [(peer, shape)] -> arrays),
    `transport.allreduce_sum(array)` sums over ranks."""

    def __init__(self, plan, part, rank, transport, perm, scatter, kkt, values, **kw):
        super().__init__(plan, perm, scatter, kkt, values, defer=True, **kw)
        self.rank, self.transport = rank, transport
        self.owner = np.asarray(part["owner"])
        nlev = plan["num_levels"]
        self.level_fronts = [np.flatnonzero(plan["level"] == l) for l in range(nlev)]
        self.cut_at = [[] for _ in range(nlev)]      # cut fronts touching this rank, per level, ascending front index
        for s in part["cut_fronts"]:
            s = int(s)
            if rank in (int(self.owner[s]), int(self.owner[plan["parent"][s]])):
                self.cut_at[int(plan["level"][s])].append(s)
        self.shared = np.asarray(part.get("shared", np.zeros(len(self.owner), dtype=bool)), dtype=bool)
        self.world = int(len(part["part_flops"]))
        self.dd[:] = 0.0
        self.ptype[:] = -1                           # other ranks' pivots: skipped by the inertia count
        parent = plan["parent"]
        for l in range(nlev):
            for s in self.level_fronts[l]:
                s = int(s)
                if self.shared[s]:
                    self.factor_shared_front(s)      # every rank, same order
                elif self.owner[s] == rank:
                    self.factor_front(s)
            # factorization transfers of the level (solver.cpp: build_cut_schedule): below a shared parent the block goes to every
            # rank, below a parent on another rank to that rank alone; shared fronts spread their own block themselves
            p2p_fronts = [s for s in self.cut_at[l] if not self.shared[s] and not self.shared[parent[s]]]
            self._transfer(l, lambda s: self.fronts[s]["cb"], self._store_cb, up=True, fronts=p2p_fronts)
            for s in self.level_fronts[l]:
                s = int(s)
                if self.shared[s] or parent[s] < 0 or not self.shared[parent[s]] or self._r(s) == 0:
                    continue
                r = self._r(s)
                cb = self.fronts[s]["cb"] if self.owner[s] == rank else np.zeros((r, r))
                self._store_cb_keep_L(s, self.transport.bcast(cb, int(self.owner[s])))
        # pivots of shared fronts are counted by their owner only
        self.count_mask = np.repeat(self.owner == rank, plan["k"])

    # ---- a front shared by all ranks (solver.cpp: enqueue_shared_batch) ---------------------------------------------------------
    NBO_E = 64       # outer panel of the emulation (the library: 256); a multiple of NBI
    CHUNK = 2        # column chunk in outer panels (option share_chunk_panels)

    def _store_cb_keep_L(self, s, a):
        old = self.fronts[s]
        self.fronts[s] = {"L": None if old is None else old["L"], "cb": a}

    def factor_shared_front(self, s):
        plan, pool, tol = self.plan, self.assembled, self.null_tol
        k, f, ld, off, first = plan["k"], plan["f"], plan["ld"], plan["pool_offset"], plan["first_col"]
        G, me, NBO, CH = self.world, self.rank, self.NBO_E, self.CHUNK
        shift = s % G                                      # the library rotates by the front's place in its batch
        fs, ks = int(f[s]), int(k[s])
        F = np.zeros((fs, fs))
        F[:, :ks] = pool[off[s] : off[s] + ld[s] * ks].reshape(ks, ld[s]).T[:fs, :]
        F = np.tril(F) + np.tril(F, -1).T
        for ch in self.children[s]:                        # every rank assembles the whole front (K5 everywhere)
            loc = self.child_map(ch)
            self.relmap[ch] = loc
            F[np.ix_(loc, loc)] += self.fronts[ch]["cb"]
        chunk_of = (np.arange(fs) // (CH * NBO))
        mine_col = ((chunk_of + shift) % G) == me          # columns of the trailing matrix this rank updates
        lower = np.tril(np.ones((fs, fs), dtype=bool))
        lp = np.arange(ks)
        q0 = first[s]
        npan = (ks + NBO - 1) // NBO
        for p in range(npan):
            J0, J1 = p * NBO, min((p + 1) * NBO, ks)
            root = ((p // CH) + shift) % G
            perms = []
            Wp = np.zeros((fs - J1, J1 - J0))
            if root == me:
                for j0 in range(J0, J1, NBI):
                    j1 = min(j0 + NBI, J1)
                    nbw = j1 - j0
                    B = F[j0:j1, j0:j1]
                    B = np.tril(B) + np.tril(B, -1).T      # only the lower triangle of a rank's own columns is kept current
                    pb, d, e, ty = bk_block(B, nbw, tol)
                    L11 = np.tril(B, -1) + np.eye(nbw)
                    self.dd[q0 + j0 : q0 + j1], self.de[q0 + j0 : q0 + j1], self.ptype[q0 + j0 : q0 + j1] = d, e, ty
                    perms.append(pb.copy())
                    lp[j0:j1] = lp[j0:j1][pb]
                    F[j0:j1, J0:j0] = F[j0:j1, J0:j0][pb, :]          # the panel's own earlier columns follow at once
                    X = F[j1:, j0:j1][:, pb]
                    Wm = np.linalg.solve(L11, X.T).T
                    D = np.diag(d) + np.diag(e[:-1], -1) + np.diag(e[:-1], 1)
                    Lm = np.linalg.solve(D, Wm.T).T
                    null = ty == 3
                    Lm[:, null] = 0.0
                    Wm[:, null] = 0.0
                    F[j0:j1, j0:j1] = np.tril(B, -1) + np.diag(d)
                    F[j1:, j0:j1] = Lm
                    if j1 < J1:                                       # inside the panel: left-looking on the panel's own columns
                        U = Wm[: , :] @ Lm[: J1 - j1, :].T
                        blk = F[j1:, j1:J1]
                        m = np.tril(np.ones_like(blk, dtype=bool))
                        blk[m] -= U[m]
                        F[j1:, j1:J1] = blk
                    Wp[:, j0 - J0 : j1 - J0] = Wm[J1 - j1 :, :]
            # the factored panel to every rank: L columns (with the diagonal block), W panel, D, pivot types and orders
            panel = self.transport.bcast(F[J0:, J0:J1].copy(), root)
            Wp = self.transport.bcast(Wp, root)
            dinfo = np.stack([self.dd[q0 + J0 : q0 + J1], self.de[q0 + J0 : q0 + J1], self.ptype[q0 + J0 : q0 + J1].astype(float), lp[J0:J1].astype(float)])
            dinfo = self.transport.bcast(dinfo, root)
            nblk = (J1 - J0 + NBI - 1) // NBI
            pm = np.zeros((nblk, NBI))
            for bi, pb in enumerate(perms):
                pm[bi, : len(pb)] = pb
            pm = self.transport.bcast(pm, root).astype(int)
            if root != me:
                F[J0:, J0:J1] = panel
                self.dd[q0 + J0 : q0 + J1], self.de[q0 + J0 : q0 + J1] = dinfo[0], dinfo[1]
                self.ptype[q0 + J0 : q0 + J1] = dinfo[2].astype(np.int32)
                lp[J0:J1] = dinfo[3].astype(int)
            # rows of the L columns left of the panel follow the panel's interchanges (k_permute_rows, on every rank)
            for bi, j0 in enumerate(range(J0, J1, NBI)):
                j1 = min(j0 + NBI, J1)
                F[j0:j1, :J0] = F[j0:j1, :J0][pm[bi, : j1 - j0], :]
            # trailing update of this rank's column chunks, lower triangle only
            if J1 < fs:
                Lp = F[J1:, J0:J1]
                U = Wp @ Lp.T
                m = lower[J1:, J1:] & mine_col[None, J1:]
                T = F[J1:, J1:]
                T[m] -= U[m]
                F[J1:, J1:] = T
        self.lperm[q0 : q0 + ks] = lp
        # contribution block: every chunk from its owner (sum of the owned columns over the ranks = gather)
        r = fs - ks
        cb = None
        if r > 0:
            M = np.where(lower[ks:, ks:] & mine_col[None, ks:], F[ks:, ks:], 0.0)
            M = self.transport.allreduce_sum(M)
            cb = np.tril(M) + np.tril(M, -1).T
        self.fronts[s] = {"L": F[:, :ks].copy(), "cb": cb}

    def _r(self, s):
        return int(self.plan["f"][s] - self.plan["k"][s])

    def _store_cb(self, s, a):
        self.fronts[s] = {"L": None, "cb": a}

    def _transfer(self, l, payload, store, up, fronts=None):
        sends, recvs, targets = [], [], []
        for s in (self.cut_at[l] if fronts is None else fronts):
            mine = self.owner[s] == self.rank
            peer = int(self.owner[self.plan["parent"][s]] if mine else self.owner[s])
            if mine == up:
                sends.append((peer, np.ascontiguousarray(payload(s))))
            else:
                recvs.append((peer, s))
                targets.append(s)
        got = self.transport.exchange(sends, [(peer, self._shape(s)) for peer, s in recvs])
        for s, a in zip(targets, got):
            store(s, a)

    _mode = "cb"

    def _shape(self, s):
        r = self._r(s)
        return (r, r) if self._mode == "cb" else (r,)

    def inertia(self):
        pos = neg = zero = 0
        for q in range(self.n):
            if not self.count_mask[q]:       # k_inertia with the owner mask
                continue
            t = self.ptype[q]
            if t == 0:
                if self.dd[q] > 0: pos += 1
                else: neg += 1
            elif t == 1:
                a, b, c = self.dd[q], self.de[q], self.dd[q + 1]
                det = a * c - b * b
                if det < 0: pos += 1; neg += 1
                elif a > 0: pos += 2
                else: neg += 2
            elif t == 3:
                zero += 1
        tot = self.transport.allreduce_sum(np.array([pos, neg, zero], dtype=np.float64))
        return tuple(int(v) for v in tot)

    def solve(self, b):
        plan = self.plan
        nsn, nlev = plan["num_fronts"], plan["num_levels"]
        rows_off, rows, k = plan["rows_off"], plan["rows"], plan["k"]
        zg = (self.scale * b)[self.perm]
        z2 = np.zeros(self.n)
        cvec = [None] * nsn
        self._mode = "vec"
        try:
            for l in range(nlev):
                for s in self.level_fronts[l]:
                    if self.owner[s] == self.rank:
                        self.forward_front(int(s), zg, z2, cvec)
                self._transfer(l, lambda s: cvec[s], lambda s, a: cvec.__setitem__(s, a), up=True)
            xg = np.zeros(self.n)
            boundary = lambda s: rows[rows_off[s] + int(k[s]) : rows_off[s + 1]]   # noqa: E731
            for l in range(nlev - 1, -1, -1):
                self._transfer(l, lambda s: xg[boundary(s)], lambda s, a: xg.__setitem__(boundary(s), a), up=False)
                for s in self.level_fronts[l]:
                    if self.owner[s] == self.rank:
                        self.backward_front(int(s), z2, xg)
        finally:
            self._mode = "cb"
        col_owner = np.repeat(self.owner, np.diff(np.append(plan["first_col"], self.n)))
        xg[col_owner != self.rank] = 0.0
        xg = self.transport.allreduce_sum(xg)
        x = np.zeros(self.n)
        x[self.perm] = xg
        return x * self.scale
