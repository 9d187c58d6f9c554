"""Shadow (symbolic) execution of a recorded cv2 instruction trace: every double that flows through an SSE2 register
or through memory carries the expression that produced it (concrete addresses come from the recorded registers, so
there is no alias analysis).  TEST-INFRASTRUCTURE tool, see tracer.c."""
import numpy as np, pickle, re, struct, sys, collections, os, cv2
SO = os.path.join(os.path.dirname(cv2.__file__), "cv2.abi3.so")
GPR = ["rax", "rcx", "rdx", "rbx", "rsp", "rbp", "rsi", "rdi", "r8", "r9", "r10", "r11", "r12", "r13", "r14", "r15"]
G32 = ["eax", "ecx", "edx", "ebx", "esp", "ebp", "esi", "edi"] + [f"r{i}d" for i in range(8, 16)]
GIDX = {n: i for i, n in enumerate(GPR)}
GIDX.update({n: i for i, n in enumerate(G32)})
DT = np.dtype([("rip", "<u8"), ("gpr", "<u8", 16), ("xmm", "<u8", (16, 2))])


def f2u(x): return struct.unpack("<Q", struct.pack("<d", x))[0]
def u2f(u): return struct.unpack("<d", struct.pack("<Q", int(u)))[0]


TRIVIAL = {0, f2u(1.0), f2u(-1.0), f2u(2.0), f2u(0.5)}
AR = {"addsd": "+", "subsd": "-", "mulsd": "*", "divsd": "/", "maxsd": "max", "minsd": "min",
      "addpd": "+", "subpd": "-", "mulpd": "*", "divpd": "/", "maxpd": "max", "minpd": "min",
      "andpd": "and", "andps": "and", "xorpd": "xor", "xorps": "xor", "orpd": "or", "orps": "or",
      "andnpd": "andn", "andnps": "andn", "pand": "and", "pxor": "xor", "por": "or", "pandn": "andn"}
NOWRITE = ("ucomisd", "comisd", "ucomiss", "comiss", "cvttsd2si", "cvtsd2si", "movmskpd", "pextrq")


class Shadow:
    def __init__(self, tr):
        self.T = np.fromfile(tr, dtype=DT)
        self.va = np.load(tr + ".va.npy")
        self.ins = pickle.load(open(tr + ".ins.pkl", "rb"))
        self.so = open(SO, "rb").read()
        self.nodes = []          # (op, a, b, valuebits or -1, step)
        self.x = [[None, None] for _ in range(16)]
        self.mem = {}
        self.g = {}
        self.byval = {}
        self.parsed = {}
        self.store_log = []      # (step, addr, node)
        self.unknown = collections.Counter()
        self.delta = int(self.T["rip"][0]) - int(self.va[0])

    def node(self, op, a, b, val, step):
        self.nodes.append((op, a, b, int(val), step))
        i = len(self.nodes) - 1
        if int(val) not in TRIVIAL and int(val) != -1:
            self.byval[int(val)] = i
        return i

    def leaf(self, kind, addr, val, step):
        v = int(val)
        if v not in TRIVIAL and v in self.byval:
            return self.byval[v]
        i = self.node(kind, addr, None, v, step)
        if kind == "in" and addr is not None:
            self.mem[addr] = i
        return i

    def parse(self, text):
        text = text.split("#")[0].strip()
        m = re.match(r"(\S+)\s*(.*)$", text)
        mn, rest = m.group(1), m.group(2)
        ops, depth, cur = [], 0, ""
        for ch in rest:
            if ch == "[": depth += 1
            if ch == "]": depth -= 1
            if ch == "," and depth == 0:
                ops.append(cur.strip()); cur = ""
            else:
                cur += ch
        if cur.strip(): ops.append(cur.strip())
        return mn, ops

    def addr(self, op, k):
        e = re.search(r"\[(.*)\]", op).group(1)
        g = self.T["gpr"][k]
        total = 0
        for sign, term in re.findall(r"([+-]?)([^+-]+)", e):
            term = term.strip()
            s = -1 if sign == "-" else 1
            if "*" in term:
                r, sc = term.split("*")
                total += s * int(g[GIDX[r]]) * int(sc, 0)
            elif term in GIDX:
                total += s * int(g[GIDX[term]])
            else:
                total += s * int(term, 0)
        return total & 0xFFFFFFFFFFFFFFFF

    def memaddr(self, op, k, text):
        if "rip" in op:
            m = re.search(r"#\s*([0-9a-f]+)", text)
            return int(m.group(1), 16) + self.delta, True
        return self.addr(op, k), False

    def constval(self, a):
        return struct.unpack_from("<Q", self.so, a - self.delta)[0]

    def load(self, a, val, k, const=False):
        n = self.mem.get(a)
        if n is not None and self.nodes[n][3] == -1:     # address-keyed leaf whose value was not visible yet
            self.nodes[n] = self.nodes[n][:3] + (int(val),) + self.nodes[n][4:]
            return n
        if n is not None and self.nodes[n][3] == int(val):
            return n
        return self.leaf("const" if const else "in", a, val, k)

    def isx(self, o): return o.startswith("xmm")
    def xi(self, o): return int(o[3:])

    def regnode(self, r, l, k):
        n = self.x[r][l]
        if n is None:
            n = self.leaf("reg", None, self.T["xmm"][k][r][l], k)
        return n

    def memnode(self, ad, c, k):
        if c:
            return self.leaf("const", ad, self.constval(ad), k)
        n = self.mem.get(ad)
        if n is None:
            n = self.node("in", ad, None, -1, k)   # value not visible in a register yet
            self.mem[ad] = n
        return n

    def run(self, k0=0, k1=None):
        T = self.T
        n = len(T) if k1 is None else k1
        for k in range(k0, min(n, len(T) - 1)):
            a = int(self.va[k])
            p = self.parsed.get(a)
            if p is None:
                p = self.parse(self.ins[a]); self.parsed[a] = p
            self.step(k, p[0], p[1], self.ins[a], T["xmm"][k + 1])

    def store(self, a, nd, k):
        self.mem[a] = nd
        self.store_log.append((k, a, nd))

    def step(self, k, mn, ops, text, post):
        x = self.x
        if mn in AR:
            d = self.xi(ops[0])
            packed = not mn.endswith("sd")
            if mn in ("xorpd", "xorps", "pxor") and ops[0] == ops[1]:
                z = self.node("zero", None, None, 0, k)
                x[d] = [z, z]
                return
            for l in range(2 if packed else 1):
                a = self.regnode(d, l, k)
                if self.isx(ops[1]):
                    b = self.regnode(self.xi(ops[1]), l, k)
                else:
                    ad, c = self.memaddr(ops[1], k, text)
                    b = self.memnode(ad + 8 * l, c, k)
                x[d][l] = self.node(AR[mn], a, b, post[d][l], k)
            return
        if mn in ("sqrtsd", "sqrtpd"):
            d = self.xi(ops[0])
            for l in range(2 if mn == "sqrtpd" else 1):
                if self.isx(ops[1]):
                    b = self.regnode(self.xi(ops[1]), l, k)
                else:
                    ad, c = self.memaddr(ops[1], k, text)
                    b = self.memnode(ad + 8 * l, c, k)
                x[d][l] = self.node("sqrt", b, None, post[d][l], k)
            return
        if mn in ("movsd", "movq", "movlpd", "movlps"):
            if self.isx(ops[0]) and self.isx(ops[1]):
                x[self.xi(ops[0])][0] = x[self.xi(ops[1])][0]
                if mn == "movq": x[self.xi(ops[0])][1] = self.node("zero", None, None, 0, k)
                return
            if self.isx(ops[0]) and "[" in ops[1]:
                d = self.xi(ops[0]); a, c = self.memaddr(ops[1], k, text)
                x[d][0] = self.load(a, post[d][0], k, c)
                if mn in ("movsd", "movq"): x[d][1] = self.node("zero", None, None, 0, k)
                return
            if "[" in ops[0] and self.isx(ops[1]):
                a, _ = self.memaddr(ops[0], k, text)
                self.store(a, self.regnode(self.xi(ops[1]), 0, k), k)
                return
            if self.isx(ops[0]) and ops[1] in GIDX:   # movq xmm, r64
                d = self.xi(ops[0]); gsh = self.g.get(GIDX[ops[1]])
                val = post[d][0]
                x[d][0] = gsh if gsh is not None and self.nodes[gsh][3] == int(val) else self.leaf("gpr", None, val, k)
                x[d][1] = self.node("zero", None, None, 0, k)
                return
            if ops[0] in GIDX and self.isx(ops[1]):   # movq r64, xmm
                self.g[GIDX[ops[0]]] = x[self.xi(ops[1])][0]
                return
        if mn in ("movhpd", "movhps"):
            if self.isx(ops[0]):
                d = self.xi(ops[0]); a, c = self.memaddr(ops[1], k, text)
                x[d][1] = self.load(a, post[d][1], k, c)
                return
            a, _ = self.memaddr(ops[0], k, text)
            self.store(a, self.regnode(self.xi(ops[1]), 1, k), k)
            return
        if mn in ("movapd", "movaps", "movupd", "movups", "movdqa", "movdqu", "lddqu"):
            if self.isx(ops[0]) and self.isx(ops[1]):
                x[self.xi(ops[0])] = list(x[self.xi(ops[1])])
                return
            if self.isx(ops[0]):
                d = self.xi(ops[0]); a, c = self.memaddr(ops[1], k, text)
                x[d] = [self.load(a, post[d][0], k, c), self.load(a + 8, post[d][1], k, c)]
                return
            a, _ = self.memaddr(ops[0], k, text); s = self.xi(ops[1])
            for l in range(2):
                self.store(a + 8 * l, self.regnode(s, l, k), k)
            return
        if mn in ("unpcklpd", "unpckhpd", "punpcklqdq", "punpckhqdq", "movlhps", "movhlps"):
            d = self.xi(ops[0])
            if self.isx(ops[1]):
                s = list(x[self.xi(ops[1])])
            else:
                a, c = self.memaddr(ops[1], k, text)
                s = [self.mem.get(a), self.mem.get(a + 8)]
            if mn in ("unpcklpd", "punpcklqdq", "movlhps"):
                x[d] = [x[d][0], s[0]]
            elif mn == "movhlps":
                x[d] = [s[1], x[d][1]]
            else:
                x[d] = [x[d][1], s[1]]
            for l in range(2):
                if x[d][l] is None or self.nodes[x[d][l]][3] != int(post[d][l]):
                    x[d][l] = self.leaf("in", None, post[d][l], k)
            return
        if mn == "movddup":
            d = self.xi(ops[0])
            if self.isx(ops[1]):
                n = self.regnode(self.xi(ops[1]), 0, k)
            else:
                a, c = self.memaddr(ops[1], k, text)
                n = self.load(a, post[d][0], k, c)
            x[d] = [n, n]
            return
        if mn in ("shufpd", "pshufd"):
            d = self.xi(ops[0])
            x[d] = [self.leaf("shuf", None, post[d][0], k), self.leaf("shuf", None, post[d][1], k)]
            return
        if mn == "mov":   # 64-bit GPR <-> memory moves may carry doubles
            if ops[0] in GPR and "[" in ops[1] and "QWORD" in ops[1]:
                a, c = self.memaddr(ops[1], k, text)
                self.g[GIDX[ops[0]]] = self.mem.get(a)
                return
            if "[" in ops[0] and "QWORD" in ops[0] and ops[1] in GPR:
                a, _ = self.memaddr(ops[0], k, text)
                nd = self.g.get(GIDX[ops[1]])
                if nd is not None and self.nodes[nd][3] == int(self.T["gpr"][k][GIDX[ops[1]]]):
                    self.store(a, nd, k)
                else:
                    self.mem.pop(a, None)
                return
            if ops[0] in GPR and ops[1] in GPR:
                self.g[GIDX[ops[0]]] = self.g.get(GIDX[ops[1]])
                return
            if ops[0] in GIDX:
                self.g.pop(GIDX[ops[0]], None)
            elif "[" in ops[0]:
                a, _ = self.memaddr(ops[0], k, text)
                self.mem.pop(a, None); self.mem.pop(a - 4, None)
            return
        if ops and self.isx(ops[0]) and mn not in NOWRITE:   # anything else: fresh leaves for the destination
            d = self.xi(ops[0])
            self.unknown[mn] += 1
            x[d] = [self.leaf(mn, None, post[d][0], k), self.leaf(mn, None, post[d][1], k)]
            return
        if ops and ops[0] in GIDX and mn not in ("cmp", "test", "push"):
            self.g.pop(GIDX[ops[0]], None)

    def expr(self, n, names, maxdepth=40):
        op, a, b, v, st = self.nodes[n]
        if n in names: return names[n]
        if maxdepth == 0: return "..."
        if op == "zero": return "0"
        if op == "const": return repr(u2f(v))
        if a is None or op in ("in", "reg", "gpr", "shuf"):
            return f"<{op}:{a if a is None else hex(a)}:{u2f(v) if v >= 0 else '?'}>"
        if op == "sqrt": return f"sqrt({self.expr(a, names, maxdepth - 1)})"
        return f"({self.expr(a, names, maxdepth - 1)} {op} {self.expr(b, names, maxdepth - 1)})"
