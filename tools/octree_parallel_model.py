"""Executable model of the prefix-sum formulation of DistributeOctTree used by k_octree (csrc/orbx_extract.cu).

Mirrors the kernel pass by pass in numpy (no list, no sequential splitting) so the reformulation itself can be
checked against the sequential oracle on the CPU:  python tools/octree_parallel_model.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def distribute_parallel(xyr, min_x, max_x, min_y, max_y, N):
    xyr = np.asarray(xyr, np.int64)
    M = len(xyr)
    if M == 0:
        return np.zeros(0, np.int64)
    x, y, resp = xyr[:, 0], xyr[:, 1], xyr[:, 2]
    n_ini = int(np.round(np.float32(max_x - min_x) / np.float32(max_y - min_y)))  # half away from zero below
    f = np.float32(max_x - min_x) / np.float32(max_y - min_y)
    n_ini = int(np.floor(float(f) + 0.5))
    hX = np.float32(np.float32(max_x - min_x) / np.float32(n_ini))
    geo = np.zeros((n_ini, 4), np.int64)
    for i in range(n_ini):
        geo[i] = [int(np.float32(hX * np.float32(i))), 0, int(np.float32(hX * np.float32(i + 1))), max_y - min_y]
    node = (x.astype(np.float32) / hX).astype(np.int64)
    cnt = np.bincount(node, minlength=n_ini)
    keepm = cnt > 0
    pos = np.cumsum(keepm) - keepm
    geo, cnt = geo[keepm], cnt[keepm]
    node = pos[node]
    n = len(cnt)

    def children(geo_i, ks):
        ulx, uly, brx, bry = geo_i
        mx = ulx + (brx - ulx + 1) // 2
        my = uly + (bry - uly + 1) // 2
        c = (x[ks] >= mx).astype(np.int64) + 2 * (y[ks] >= my).astype(np.int64)
        rects = [[ulx, uly, mx, my], [mx, uly, brx, my], [ulx, my, mx, bry], [mx, my, brx, bry]]
        return c, rects

    V = []
    finish = False
    while not finish:
        prev = n
        exp = cnt > 1
        child_cnt = np.zeros((n, 4), np.int64)
        kc = np.zeros(M, np.int64)
        rects_of = {}
        for i in np.nonzero(exp)[0]:
            ks = np.nonzero(node == i)[0]
            c, rects = children(geo[i], ks)
            kc[ks] = c
            rects_of[i] = rects
            child_cnt[i] = np.bincount(c, minlength=4)
        nch = (child_cnt > 0).sum(1) * exp
        P = np.cumsum(nch) - nch
        Q = np.cumsum(~exp) - (~exp)
        C = int(nch.sum())
        n_new = C + int((~exp).sum())
        geo_n = np.zeros((n_new, 4), np.int64)
        cnt_n = np.zeros(n_new, np.int64)
        child_pos = -np.ones((n, 4), np.int64)
        node_pos = -np.ones(n, np.int64)
        V = []
        for i in range(n):
            if exp[i]:
                ci = P[i]
                for c in range(4):
                    if child_cnt[i, c] > 0:
                        p = C - 1 - ci
                        geo_n[p] = rects_of[i][c]
                        cnt_n[p] = child_cnt[i, c]
                        child_pos[i, c] = p
                        if child_cnt[i, c] > 1:
                            V.append(p)
                        ci += 1
            else:
                p = C + Q[i]
                geo_n[p] = geo[i]
                cnt_n[p] = cnt[i]
                node_pos[i] = p
        node = np.where(exp[node], child_pos[node, kc], node_pos[node])
        geo, cnt, n = geo_n, cnt_n, n_new
        n_to_expand = len(V)
        if n >= N or n == prev:
            finish = True
        elif n + 3 * n_to_expand > N:
            while not finish:
                if not V:
                    finish = True
                    break
                prev2 = n
                nV = len(V)
                sizes = cnt[V]
                # processing order: (size desc, slot desc)
                order = sorted(range(nV), key=lambda s: (-sizes[s], -s))
                slot_of = -np.ones(n, np.int64)
                slot_of[V] = np.arange(nV)
                proc_of_slot = np.zeros(nV, np.int64)
                proc_of_slot[order] = np.arange(nV)
                child_cnt = np.zeros((nV, 4), np.int64)
                kc = np.zeros(M, np.int64)
                rects_of = {}
                for s in range(nV):
                    i = V[s]
                    ks = np.nonzero(node == i)[0]
                    c, rects = children(geo[i], ks)
                    kc[ks] = c
                    rects_of[s] = rects
                    child_cnt[s] = np.bincount(c, minlength=4)
                nch_p = np.array([(child_cnt[order[p]] > 0).sum() for p in range(nV)])
                excl = np.cumsum(nch_p) - nch_p
                hit = np.nonzero(n + excl + nch_p - (np.arange(nV) + 1) >= N)[0]
                cut = int(hit[0]) if len(hit) else nV - 1
                Ctot = int(excl[cut] + nch_p[cut])
                processed_node = np.zeros(n, bool)
                for p in range(cut + 1):
                    processed_node[V[order[p]]] = True
                keep_old = ~processed_node
                Qo = np.cumsum(keep_old) - keep_old
                n_new = Ctot + int(keep_old.sum())
                geo_n = np.zeros((n_new, 4), np.int64)
                cnt_n = np.zeros(n_new, np.int64)
                child_pos = -np.ones((nV, 4), np.int64)
                node_pos = -np.ones(n, np.int64)
                Vn = []
                for p in range(cut + 1):
                    s = order[p]
                    ci = excl[p]
                    for c in range(4):
                        if child_cnt[s, c] > 0:
                            q = Ctot - 1 - ci
                            geo_n[q] = rects_of[s][c]
                            cnt_n[q] = child_cnt[s, c]
                            child_pos[s, c] = q
                            if child_cnt[s, c] > 1:
                                Vn.append(q)
                            ci += 1
                for i in range(n):
                    if keep_old[i]:
                        q = Ctot + Qo[i]
                        geo_n[q] = geo[i]
                        cnt_n[q] = cnt[i]
                        node_pos[i] = q
                s_of_key = slot_of[node]
                isproc = processed_node[node]
                node = np.where(isproc, child_pos[np.maximum(s_of_key, 0), kc], node_pos[node])
                geo, cnt, n = geo_n, cnt_n, n_new
                V = Vn
                if n >= N or n == prev2:
                    finish = True
    # selection: max response, first in candidate order
    out = np.zeros(n, np.int64)
    for i in range(n):
        ks = np.nonzero(node == i)[0]
        out[i] = ks[np.argmax(resp[ks])]  # argmax returns the first maximum
    return out


if __name__ == "__main__":
    import oracle

    rng = np.random.default_rng(0)
    bad = 0
    trials = 0
    for t in range(400):
        W = int(rng.integers(40, 700))
        H = int(rng.integers(40, 500))
        if round(W / H) < 1:
            continue
        M = int(rng.integers(0, 2500))
        N = int(rng.integers(1, 500))
        # distinct integer positions like FAST output
        allpos = rng.permutation(W * H)[: min(M, W * H)]
        if t % 5 == 0 and len(allpos):  # clustered
            cx, cy = rng.integers(0, W), rng.integers(0, H)
            xs = np.clip((rng.normal(cx, W / 10, len(allpos))).astype(int), 0, W - 1)
            ys = np.clip((rng.normal(cy, H / 10, len(allpos))).astype(int), 0, H - 1)
            uniq = np.unique(ys * W + xs)
            rng.shuffle(uniq)
            allpos = uniq
        xs, ys = allpos % W, allpos // W
        # candidate order = cell-major like the reference; any fixed order is fine for the equivalence check
        resp = rng.integers(7, 60 if t % 2 else 255, len(allpos))
        xyr = np.stack([xs, ys, resp], 1).astype(np.int32)
        ref = oracle.distribute(xyr, 16, 16 + W, 16, 16 + H, N)
        got = distribute_parallel(xyr, 16, 16 + W, 16, 16 + H, N)
        trials += 1
        if len(ref) != len(got) or not np.array_equal(ref, got):
            bad += 1
            print("MISMATCH trial", t, W, H, len(xyr), N, len(ref), len(got))
    print("trials", trials, "mismatches", bad)
    sys.exit(1 if bad else 0)
