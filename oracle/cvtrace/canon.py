"""Canonicalise the shadow DAG of a straight-line region into ordered term lists using only exact identities
(commutativity of + and *, exactness of negation):  out = fold(+/-) of products, product = left-associated chain
over leaves {inputs, inputs^2, inputs^3, constants}.  TEST-INFRASTRUCTURE tool, see tracer.c."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from shadow import *
sys.setrecursionlimit(100000)


class Canon:
    def __init__(self, S, inbase, nin):
        self.S = S
        self.leafid = {}
        for i, n in enumerate(S.nodes):
            if n[0] == "in" and n[1] is not None and inbase <= n[1] < inbase + 8 * nin:
                self.leafid[i] = ("v", (n[1] - inbase) // 8)
            if n[0] == "const":
                self.leafid[i] = ("c", u2f(n[3]))
        for i, n in enumerate(S.nodes):   # squares and cubes of inputs (OpenCV's ep2 / ep3 tables)
            if n[0] == "*" and n[1] in self.leafid and n[2] in self.leafid:
                a, b = self.leafid[n[1]], self.leafid[n[2]]
                if a[0] == "v" and b == a: self.leafid[i] = ("v2", a[1])
                elif a[0] == "v2" and b == ("v", a[1]): self.leafid[i] = ("v3", a[1])
                elif b[0] == "v2" and a == ("v", b[1]): self.leafid[i] = ("v3", b[1])

    def inner(self, nd):
        return nd[1] if self.S.nodes[nd[2]][0] == "const" else nd[2]

    def is_sum(self, n): return self.S.nodes[n][0] in ("+", "-") and n not in self.leafid

    def isneg_sum(self, n):
        nd = self.S.nodes[n]
        if nd[0] != "xor": return False
        i = self.inner(nd)
        return self.is_sum(i) or self.isneg_sum(i)

    def schain(self, n):
        nd = self.S.nodes[n]
        if n in self.leafid: return 1, [self.leafid[n]]
        if nd[0] == "xor":
            s, c = self.schain(self.inner(nd)); return -s, c
        if nd[0] != "*": raise ValueError(f"not a product {nd}")
        x, y = nd[1], nd[2]
        def neg_leaf(z):
            zd = self.S.nodes[z]
            if zd[0] == "xor" and self.inner(zd) in self.leafid: return self.inner(zd)
            return None
        if y in self.leafid:
            s, c = self.schain(x); return s, c + [self.leafid[y]]
        if x in self.leafid:
            s, c = self.schain(y); return s, c + [self.leafid[x]]
        if neg_leaf(y) is not None:
            s, c = self.schain(x); return -s, c + [self.leafid[neg_leaf(y)]]
        if neg_leaf(x) is not None:
            s, c = self.schain(y); return -s, c + [self.leafid[neg_leaf(x)]]
        raise ValueError("product of two non-leaves")

    def flat(self, n):
        nd = self.S.nodes[n]
        if nd[0] == "xor" and self.isneg_sum(n):
            return [(-s, t) for s, t in self.flat(self.inner(nd))]
        if not self.is_sum(n): return [self.schain(n)]
        x, y = nd[1], nd[2]
        xs = self.is_sum(x) or self.isneg_sum(x)
        ys = self.is_sum(y) or self.isneg_sum(y)
        neg = lambda f: [(-s, t) for s, t in f]
        if nd[0] == "+":
            if xs and not ys: return self.flat(x) + [self.schain(y)]
            if ys and not xs: return self.flat(y) + [self.schain(x)]
            if not xs and not ys: return [self.schain(x), self.schain(y)]
            fx, fy = self.flat(x), self.flat(y)
            if len(fy) == 1: return fx + fy
            if len(fx) == 1: return fy + fx
            raise ValueError("sum of two chains")
        if not ys: return (self.flat(x) if xs else [self.schain(x)]) + neg([self.schain(y)])
        if not xs: return neg(self.flat(y)) + [self.schain(x)]
        fx, fy = self.flat(x), self.flat(y)
        if len(fy) == 1: return fx + neg(fy)
        if len(fx) == 1: return neg(fy) + fx
        raise ValueError("difference of two chains")

    @staticmethod
    def evaluate(terms, vin):
        def lv(l):
            k, i = l
            if k == "c": return np.float64(i)
            v = np.float64(vin[i])
            return v if k == "v" else v * v if k == "v2" else (v * v) * v
        acc = None
        for s, ch in terms:
            p = lv(ch[0])
            for l in ch[1:]: p = p * lv(l)
            acc = (p if s > 0 else -p) if acc is None else (acc + p if s > 0 else acc - p)
        return acc
