"""Offline tree-quality study (CPU only): box / triangle tests per ray segment of the ORDERED policy on candidate
internal trees, counted by the oracle.  Candidates are numpy restatements of csrc/rt_lbvh.cuh (Morton order + radix
tree, leaves of <= L triangles) with optional changes, written out in the reference BvhNode format.
  python scripts/tree_study.py [nx nz w h]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpu_raytracer_b200 as g
from gpu_raytracer_b200 import abi
import oracle_binding as ob

NONE = 0xFFFFFFFF


def expand21(v):
    v = v.astype(np.uint64) & np.uint64(0x1FFFFF)
    v = (v | v << np.uint64(32)) & np.uint64(0x1F00000000FFFF)
    v = (v | v << np.uint64(16)) & np.uint64(0x1F0000FF0000FF)
    v = (v | v << np.uint64(8)) & np.uint64(0x100F00F00F00F00F)
    v = (v | v << np.uint64(4)) & np.uint64(0x10C30C30C30C30C3)
    v = (v | v << np.uint64(2)) & np.uint64(0x1249249249249249)
    return v


def tri_boxes(sc):
    v = sc.vertices["position"]
    t = sc.triangles
    p = np.stack([v[t["v0_index"]], v[t["v1_index"]], v[t["v2_index"]]], 1)
    return p.min(1), p.max(1)


def morton(lo, hi, axes=(0, 1, 2)):
    c = 0.5 * (lo + hi)
    mn, mx = c.min(0), c.max(0)
    ext = float((mx - mn).max())
    q = np.minimum(2097151.0, np.clip((c - mn) / ext, 0, 1) * 2097152.0).astype(np.uint64)
    if len(axes) == 3:
        return expand21(q[:, 0]) << np.uint64(2) | expand21(q[:, 1]) << np.uint64(1) | expand21(q[:, 2])
    a, b = axes  # 2D: interleave two axes only (expand to every 2nd bit)
    def e2(v):
        v = v & np.uint64(0x1FFFFF)
        v = (v | v << np.uint64(16)) & np.uint64(0x0000FFFF0000FFFF)
        v = (v | v << np.uint64(8)) & np.uint64(0x00FF00FF00FF00FF)
        v = (v | v << np.uint64(4)) & np.uint64(0x0F0F0F0F0F0F0F0F)
        v = (v | v << np.uint64(2)) & np.uint64(0x3333333333333333)
        v = (v | v << np.uint64(1)) & np.uint64(0x5555555555555555)
        return v
    return e2(q[:, a]) << np.uint64(1) | e2(q[:, b])


def morton_hybrid(lo, hi, yb):
    """x/z code whose levels b <= yb (cells of side extent / 2^(20 - b) and finer) also carry the y bit."""
    c = 0.5 * (lo + hi)
    mn, mx = c.min(0), c.max(0)
    ext = float((mx - mn).max())
    q = np.minimum(2097151.0, np.clip((c - mn) / ext, 0, 1) * 2097152.0).astype(np.uint64)
    key = np.zeros(len(q), np.uint64)
    one = np.uint64(1)
    for b in range(20, -1, -1):
        for k in (0, 1, 2):
            if k == 1 and b > yb: continue
            key = key << one | (q[:, k] >> np.uint64(b) & one)
    return key


def radix_tree(keys):
    """children[i] = (left, right), leaves encoded as ~slot (negative).  Top-down highest-differing-bit splits:
    the same tree Karras' construction yields."""
    n = len(keys)
    children = np.zeros((n - 1, 2), np.int64)
    rng = np.zeros((n - 1, 2), np.int64)
    nxt = 0
    root = 0
    stack = [(0, n - 1, -1, 0)]
    ks = keys
    while stack:
        lo, hi, par, side = stack.pop()
        if lo == hi:
            children[par, side] = ~lo
            continue
        me = nxt; nxt += 1
        if par >= 0: children[par, side] = me
        rng[me] = (lo, hi)
        a, b = int(ks[lo]), int(ks[hi])
        if a == b:
            x = lo ^ hi
            bit = x.bit_length() - 1
            split = (hi >> bit) << bit  # first index with that bit set
            split -= 1
        else:
            bit = (a ^ b).bit_length() - 1
            pref = (b >> bit) << bit
            split = int(np.searchsorted(ks[lo:hi + 1], np.uint64(pref), side="left")) + lo - 1
        stack.append((split + 1, hi, me, 1))
        stack.append((lo, split, me, 0))
    return children[:nxt], rng[:nxt]


def area(lo, hi):
    d = np.maximum(hi - lo, 0)
    return 2 * (d[..., 0] * d[..., 1] + d[..., 1] * d[..., 2] + d[..., 2] * d[..., 0])


class Tree:
    """Binary tree over sorted slots; node k: children (int, leaf = ~slot), box."""
    def __init__(self, children, rng, plo, phi):
        self.ch, self.rng = children.copy(), rng.copy()
        self.plo, self.phi = plo, phi  # per sorted slot
        n = len(children)
        self.lo = np.zeros((n, 3), np.float32); self.hi = np.zeros((n, 3), np.float32)
        self.order = self.postorder()
        for k in self.order:
            self.refit(k)

    def postorder(self):
        out, st = [], [0]
        while st:
            k = st.pop(); out.append(k)
            for c in self.ch[k]:
                if c >= 0: st.append(int(c))
        return out[::-1]

    def box(self, c):
        return (self.plo[~c], self.phi[~c]) if c < 0 else (self.lo[c], self.hi[c])

    def refit(self, k):
        (al, ah), (bl, bh) = self.box(self.ch[k, 0]), self.box(self.ch[k, 1])
        self.lo[k] = np.minimum(al, bl); self.hi[k] = np.maximum(ah, bh)

    def count(self, c):
        return 1 if c < 0 else int(self.rng[c, 1] - self.rng[c, 0] + 1)

    def rotate(self, leaf_max):
        """One bottom-up pass of tree rotations (Kensler 2008): at node X swap a child with a grandchild under the
        other child when that shrinks the area of the rebuilt child.  Subtrees that will be collapsed into leaves
        (<= leaf_max triangles) are treated as leaves.  NOTE: rotations break the contiguous-slot property of
        subtrees, so ranges are recomputed by flatten()."""
        nrot = 0
        self.order = self.postorder()
        cnt = {}
        def cnt_of(c):
            if c < 0: return 1
            return cnt[c]
        for k in self.order:
            a, b = int(self.ch[k, 0]), int(self.ch[k, 1])
            cnt[k] = cnt_of(a) + cnt_of(b)
            best, bestv = None, 0.0
            for side in (0, 1):
                keep, other = (a, b) if side == 0 else (b, a)   # swap `keep` with a grandchild under `other`
                if other < 0 or cnt_of(other) <= leaf_max: continue
                cur = float(area(self.lo[other], self.hi[other]))
                kl, kh = self.box(keep)
                for gs in (0, 1):
                    stay = int(self.ch[other, 1 - gs])
                    sl, sh = self.box(stay)
                    new = float(area(np.minimum(kl, sl), np.maximum(kh, sh)))
                    if new - cur < bestv:
                        bestv, best = new - cur, (side, gs)
            if best:
                side, gs = best
                keep, other = (a, b) if side == 0 else (b, a)
                gc = int(self.ch[other, gs])
                self.ch[other, gs] = keep
                self.ch[k, side] = gc
                cnt[other] = cnt_of(int(self.ch[other, 0])) + cnt_of(int(self.ch[other, 1]))
                self.refit(other)
                nrot += 1
        return nrot

    def flatten(self, sorted_tri, leaf_max):
        """-> reference-format nodes + triangle_indices (DFS order, leaves of <= leaf_max triangles)."""
        # subtree triangle lists after rotations: gather by DFS
        nodes, idx = [], []
        def leaf_slots(c, out):
            st = [c]
            while st:
                x = st.pop()
                if x < 0: out.append(~x)
                else: st.append(int(self.ch[x, 1])); st.append(int(self.ch[x, 0]))
        cnt = {}
        self.order = self.postorder()
        for k in self.order:
            cnt[k] = sum(1 if c < 0 else cnt[int(c)] for c in self.ch[k])
        out_nodes = []
        st = [(0, -1, 0)]
        while st:
            c, par, side = st.pop()
            me = len(out_nodes)
            if par >= 0: out_nodes[par][2 + side] = me
            n_c = 1 if c < 0 else cnt[c]
            if c < 0 or n_c <= leaf_max:
                sl = []
                leaf_slots(c, sl)
                lo = self.plo[sl].min(0); hi = self.phi[sl].max(0)
                out_nodes.append([lo, hi, NONE, NONE, len(idx), len(sl)])
                idx.extend(int(sorted_tri[s]) for s in sl)
            else:
                out_nodes.append([self.lo[c].copy(), self.hi[c].copy(), NONE, NONE, 0, 0])
                st.append((int(self.ch[c, 1]), me, 1))
                st.append((int(self.ch[c, 0]), me, 0))
        arr = np.zeros(len(out_nodes), abi.BvhNode)
        for i, (lo, hi, l, r, s, cn) in enumerate(out_nodes):
            arr[i]["bounds"]["min"] = lo; arr[i]["bounds"]["max"] = hi
            arr[i]["left_child"] = l; arr[i]["right_child"] = r
            arr[i]["triangle_start"] = s; arr[i]["triangle_count"] = cn
        return arr, np.asarray(idx, np.uint32)

    def sah(self, leaf_max):
        cnt = {}
        tot = 0.0
        self.order = self.postorder()
        for k in self.order:
            cnt[k] = sum(1 if c < 0 else cnt[int(c)] for c in self.ch[k])
        for k in self.order:
            if cnt[k] > leaf_max: tot += float(area(self.lo[k], self.hi[k]))
        return tot / float(area(self.lo[0], self.hi[0]))


def evaluate(sc, nodes, idx, w, h, label):
    sc.bvh_nodes, sc.triangle_indices = nodes, idx
    t = time.time()
    r0 = ob.render_path(sc, w, h, spp=1, max_depth=0, policy=1, want_hits=False)["counters"]
    r4 = ob.render_path(sc, w, h, spp=1, max_depth=4, policy=1, want_hits=False)["counters"]
    leaves = int((nodes["left_child"] == NONE).sum())
    print("%-34s nodes %7d leaves %7d | primary V %.1f T %.2f | depth4 V %.1f T %.2f  (%.1fs)" % (
        label, len(nodes), leaves, r0["aabb_tests"] / r0["rays"], r0["tri_tests"] / r0["rays"],
        r4["aabb_tests"] / r4["rays"], r4["tri_tests"] / r4["rays"], time.time() - t), flush=True)


def main():
    a = [int(x) for x in sys.argv[1:5]] if len(sys.argv) >= 5 else [500, 250, 960, 540]
    nx, nz, w, h = a
    if os.environ.get("STUDY_SCENE", "").startswith("field"):  # STUDY_SCENE=field:499 -> 499 tori + ground, not a height field
        n_tori = int(os.environ["STUDY_SCENE"].partition(":")[2] or 499)
        sc = g.field_scene(n_tori=n_tori, gnx=20, gnz=(250000 - 500 * n_tori) // 40 if n_tori == 499 else 25, strategy=abi.BVH_SMALL_LEAF)
        print("field scene: %d tori, %d triangles" % (n_tori, len(sc.triangles)), flush=True)
    else:
        sc = g.terrain_scene(nx, nz, seed=1, strategy=abi.BVH_SMALL_LEAF)
    evaluate(sc, sc.bvh_nodes, sc.triangle_indices, w, h, "binned SAH <=4 (rt_bvh_build)")
    if len(sc.triangles) <= 100000:
        n, i = g.build_bvh(sc.triangles, sc.vertices, abi.BVH_STANDARD)
        evaluate(sc, n, i, w, h, "standard format, 1 tri/leaf")
    lo, hi = tri_boxes(sc)
    cands = [("morton3", (0, 1, 2)), ("morton2-xz", (0, 2))]
    cands += [("hybrid-y<=%s" % b, int(b)) for b in os.environ.get("STUDY_HYBRID", "").split(",") if b]
    if os.environ.get("STUDY_ONLY"): cands = [c for c in cands if c[0].startswith(tuple(os.environ["STUDY_ONLY"].split(",")))]
    for name, axes in cands:
        keys = morton_hybrid(lo, hi, axes) if isinstance(axes, int) else morton(lo, hi, axes)
        order = np.argsort(keys, kind="stable")
        ks = keys[order]
        ch, rg = radix_tree(ks)
        for L in [int(x) for x in os.environ.get("STUDY_LEAF", "2,4").split(",")]:
            tr = Tree(ch, rg, lo[order], hi[order])
            n, i = tr.flatten(order, L)
            evaluate(sc, n, i, w, h, "%s leaf<=%d  sah %.1f" % (name, L, tr.sah(L)))
            for it in range(int(os.environ.get("STUDY_ROT", "2"))):
                nr = tr.rotate(L)
                for k in tr.order: tr.refit(k)
                n, i = tr.flatten(order, L)
                evaluate(sc, n, i, w, h, "%s leaf<=%d rot%d (%d) sah %.1f" % (name, L, it + 1, nr, tr.sah(L)))


if __name__ == "__main__" and not (len(sys.argv) > 1 and sys.argv[1] == "ploc"):
    main()


def ploc(lo, hi, order, radius=16):
    """PLOC (Meister & Bittner 2018) over the Morton order `order`: repeated merging of mutually nearest clusters within
    `radius` positions.  Returns children (root = node 0, leaves = ~sorted slot) like radix_tree()."""
    n = len(order)
    clo, chi = lo[order].astype(np.float32).copy(), hi[order].astype(np.float32).copy()
    cid = -(np.arange(n, dtype=np.int64) + 1)  # leaf k encoded as ~k = -(k+1)
    nodes = []  # (left, right) in creation order
    while len(cid) > 1:
        m = len(cid)
        best = np.full(m, np.inf, np.float32)
        nn = np.full(m, -1, np.int64)
        for d in range(1, min(radius, m - 1) + 1):
            a = area(np.minimum(clo[:-d], clo[d:]), np.maximum(chi[:-d], chi[d:]))
            i = np.arange(m - d)
            upd = a < best[:-d]
            best[:-d][upd] = a[upd]; nn[:-d][upd] = (i + d)[upd]
            upd = a < best[d:]
            best[d:][upd] = a[upd]; nn[d:][upd] = i[upd]
        idx = np.arange(m)
        mutual = (nn[nn] == idx) & (idx < nn)
        left = idx[mutual]; right = nn[left]
        base = len(nodes)
        for l, r in zip(cid[left], cid[right]):
            nodes.append((int(l), int(r)))
        new_ids = base + np.arange(len(left))
        clo[left] = np.minimum(clo[left], clo[right]); chi[left] = np.maximum(chi[left], chi[right])
        cid[left] = new_ids
        keep = np.ones(m, bool); keep[right] = False
        clo, chi, cid = clo[keep], chi[keep], cid[keep]
    t = len(nodes)
    ch = np.zeros((t, 2), np.int64)
    for k, (l, r) in enumerate(nodes):  # renumber: creation index k -> t - 1 - k (root = 0)
        ch[t - 1 - k] = (l if l < 0 else t - 1 - l, r if r < 0 else t - 1 - r)
    return ch


def main_ploc():
    a = [int(x) for x in sys.argv[2:6]] if len(sys.argv) >= 6 else [500, 250, 960, 540]
    nx, nz, w, h = a
    if os.environ.get("STUDY_SCENE", "").startswith("field"):
        n_tori = int(os.environ["STUDY_SCENE"].partition(":")[2] or 499)
        sc = g.field_scene(n_tori=n_tori, gnx=20, gnz=(250000 - 500 * n_tori) // 40 if n_tori == 499 else 25, strategy=abi.BVH_SMALL_LEAF)
    else:
        sc = g.terrain_scene(nx, nz, seed=1, strategy=abi.BVH_SMALL_LEAF)
    evaluate(sc, sc.bvh_nodes, sc.triangle_indices, w, h, "binned SAH <=4 (rt_bvh_build)")
    lo, hi = tri_boxes(sc)
    ext = (0.5 * (lo + hi)).max(0) - (0.5 * (lo + hi)).min(0)
    axes = (0, 1, 2) if ext.min() >= 0.125 * ext.max() else tuple(int(k) for k in np.argsort(-ext)[:2])
    axes = tuple(sorted(axes))
    keys = morton(lo, hi, axes)
    order = np.argsort(keys, kind="stable")
    ch, rg = radix_tree(keys[order])
    for L in (2,):
        tr = Tree(ch, rg, lo[order], hi[order])
        n, i = tr.flatten(order, L)
        evaluate(sc, n, i, w, h, "morton%s radix tree leaf<=%d sah %.1f" % (axes, L, tr.sah(L)))
        for r in [int(x) for x in os.environ.get("STUDY_RADIUS", "8,16,32").split(",")]:
            pc = ploc(lo, hi, order, r)
            tp = Tree(pc, np.zeros((len(pc), 2), np.int64), lo[order], hi[order])
            n, i = tp.flatten(order, L)
            evaluate(sc, n, i, w, h, "PLOC r=%d leaf<=%d sah %.1f" % (r, L, tp.sah(L)))


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "ploc":
    main_ploc()
