"""Step-level numpy model of the data-parallel quadtree of my_orb_slam2_b200/csrc/k_quadtree.cu.

The kernel replaces the reference's std::list surgery (src/ORBextractor.cc:539-765) by rank / scan steps; this
model performs exactly those steps (same formulas, sequentially) so that the ALGORITHM can be checked against the
CPU oracle without a GPU (tests/test_quadtree_model.py).  It is test infrastructure, not a product path."""
import numpy as np


def distribute(xs, ys, scores, width, height, n_target, w_cell, h_cell):
    """xs, ys relative to minBorder; width/height = maxX-minX, maxY-minY. Returns kept candidate indices in list order."""
    n = len(xs)
    if n == 0:
        return []
    f32 = np.float32
    n_ini = int(np.floor(f32(width) / f32(height) + f32(0.5)))  # roundf for positive values
    h_x = f32(width) / f32(n_ini)
    box = [(int(h_x * f32(i)), int(h_x * f32(i + 1)), 0, height) for i in range(n_ini)]
    node_of = np.minimum((xs.astype(np.float32) / h_x).astype(np.int64), n_ini - 1)
    cnt = np.bincount(node_of, minlength=n_ini)
    keep = np.nonzero(cnt > 0)[0]
    remap = -np.ones(n_ini, np.int64)
    remap[keep] = np.arange(len(keep))
    box = [box[i] for i in keep]
    cnt = cnt[keep]
    node_of = remap[node_of]
    S, phase = len(box), 1

    def quadrant(bx, x, y):
        mx = bx[0] + ((bx[1] - bx[0] + 1) >> 1)
        my = bx[2] + ((bx[3] - bx[2] + 1) >> 1)
        return (0 if x < mx else 1) + (0 if y < my else 2)

    while True:
        cc = np.zeros((S, 4), np.int64)
        for k in range(n):
            nd = node_of[k]
            if cnt[nd] > 1:
                cc[nd, quadrant(box[nd], xs[k], ys[k])] += 1
        elist = [i for i in range(S) if cnt[i] > 1]
        E = len(elist)
        if E == 0:
            break
        nrank = -np.ones(S, np.int64)
        vis = [0] * E
        for a, i in enumerate(elist):
            r = a
            if phase == 2:
                r = sum(1 for a2, j in enumerate(elist) if cnt[j] > cnt[i] or (cnt[j] == cnt[i] and a2 < a))
            nrank[i] = r
            vis[r] = i
        nne = np.array([(cc[vis[r]] > 0).sum() for r in range(E)])
        push = np.concatenate([[0], np.cumsum(nne)])  # exclusive scan, push[E] = total
        P = E
        if phase == 2:
            for r in range(E):
                if S + push[r] + nne[r] - (r + 1) >= n_target:
                    P = r + 1
                    break
        TP = push[P]
        surv_flag = np.array([0 if (0 <= nrank[i] < P) else 1 for i in range(S)])
        surv = np.concatenate([[0], np.cumsum(surv_flag)])
        S2 = TP + surv[S]
        nbox = [None] * S2
        ncnt = np.zeros(S2, np.int64)
        slot = np.zeros((S, 4), np.int64)
        newpos = np.zeros(S, np.int64)
        n_expand = 0
        for i in range(S):
            r = nrank[i]
            if 0 <= r < P:
                bx = box[i]
                mx = bx[0] + ((bx[1] - bx[0] + 1) >> 1)
                my = bx[2] + ((bx[3] - bx[2] + 1) >> 1)
                pi = push[r]
                for q in range(4):
                    c = cc[i, q]
                    if c > 0:
                        pos = TP - 1 - pi
                        pi += 1
                        nbox[pos] = (mx if q & 1 else bx[0], bx[1] if q & 1 else mx, my if q & 2 else bx[2],
                                     bx[3] if q & 2 else my)
                        ncnt[pos] = c
                        slot[i, q] = pos
                        n_expand += c > 1
            else:
                pos = TP + surv[i]
                nbox[pos] = box[i]
                ncnt[pos] = cnt[i]
                newpos[i] = pos
        for k in range(n):
            nd = node_of[k]
            r = nrank[nd]
            if 0 <= r < P:
                node_of[k] = slot[nd, quadrant(box[nd], xs[k], ys[k])]
            else:
                node_of[k] = newpos[nd]
        prev, S, box, cnt = S, S2, nbox, ncnt
        if S >= n_target or S == prev:
            break
        if phase == 1 and S + 3 * n_expand > n_target:
            phase = 2
    best = [None] * S
    for k in range(n):
        cy, cx = (ys[k] - 3) // h_cell, (xs[k] - 3) // w_cell
        key = (int(scores[k]), -int(cy), -int(cx), -int(ys[k]), -int(xs[k]))
        nd = node_of[k]
        if best[nd] is None or key > best[nd][0]:
            best[nd] = (key, k)
    return [b[1] for b in best]
