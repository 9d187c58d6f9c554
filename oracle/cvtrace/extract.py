"""Extract the two generated straight-line blocks of OpenCV's EMEstimatorCallback::runKernel from a trace:
   getCoeffMat (36 null-space entries e[] -> the 10x20 constraint matrix A[]) and the determinant polynomial
   (39 entries b[] of B(z) -> c[0..10]), as ordered term tables, verified bit for bit against the values the binary
   produced.  Writes TRACE.tables.pkl.  TEST-INFRASTRUCTURE tool, see tracer.c."""
from canon import *
import pickle

A_LOOP = 0x15b8c80      # ep2/ep3 loop head of getCoeffMat (cv2 4.13.0.92 manylinux x86_64 wheel)
A_BODY = 0x15b8ca6      # first instruction after the loop
A_PERM = 0x15de542      # getCoeffMat's closing column permutation (table of 20 ints in .rodata, copied to the stack)
A_PERM_TABLE = 0x3a5dd10
P_BODY = 0x15e8acf      # first instruction of the c[] block
POLY_CALL = 0x6fa8c0    # cv::solvePoly


def first(va, addr, start=0):
    idx = np.nonzero(va[start:] == addr)[0]
    return int(idx[0]) + start


def main(tr):
    S = Shadow(tr)
    va = S.va
    kA0 = first(va, A_LOOP); kA = first(va, A_BODY); kP = first(va, P_BODY); kE = first(va, POLY_CALL, kP)
    # end of getCoeffMat: first ret after the body
    kAe = kA
    while not S.ins[int(va[kAe])].startswith("ret"): kAe += 1
    print("steps: coeff", kA0, kA, kAe, "poly", kP, kE)
    S.run(kA0, kAe)
    g = S.T["gpr"][kA]
    eb, ab = int(g[GIDX["rax"]]), int(g[GIDX["rdx"]])
    C = Canon(S, eb, 36)
    kPerm = first(va, A_PERM)
    outs = {(a - ab) // 8: nd for (k, a, nd) in S.store_log if ab <= a < ab + 1600 and k < kPerm}
    perm = list(struct.unpack_from("<20i", S.so, A_PERM_TABLE))
    assert sorted(perm) == list(range(20))
    print("column permutation (dest column j <- computed column perm[j]):", perm)
    ev = np.zeros(36)
    for n in S.nodes:
        if n[0] == "in" and n[1] is not None and eb <= n[1] < eb + 288 and n[3] != -1:
            ev[(n[1] - eb) // 8] = u2f(n[3])
    assert len(outs) == 200 and np.count_nonzero(ev) == 36
    coef = {}
    for k in range(200):
        t = C.flat(outs[k])
        assert C.evaluate(t, ev) == u2f(S.nodes[outs[k]][3]), k
        coef[k] = t
    print("getCoeffMat: 200 outputs,", sum(len(t) for t in coef.values()), "terms, bit-exact re-evaluation")

    # b[] values from the stage before, then the polynomial block
    S2 = Shadow(tr); S2.run(kAe, kP)
    S3 = Shadow(tr); S3.run(kP, kE)
    rbp = int(S3.T["gpr"][kP][GIDX["rbp"]])
    first_ins = S3.ins[int(va[kP])]
    b0 = int(re.search(r"\[rbp-(0x[0-9a-f]+)\]", first_ins).group(1), 16)   # first load is b[0]
    bb = rbp - b0
    C3 = Canon(S3, bb, 39)
    st = [(k, a, nd) for (k, a, nd) in S3.store_log]
    # c[k]: eleven 8-byte slots at consecutive addresses, stored in the order c[10] .. c[0]
    cand = {}
    for (k, a, nd) in st:
        cand[a] = (k, nd)
    addrs = sorted(cand)
    cb = None
    for a in addrs:
        if all(a + 8 * j in cand for j in range(11)) and all(cand[a + 8 * j][0] > cand[a + 8 * (j + 1)][0] for j in range(10)):
            cb = a
    assert cb is not None
    bv = np.zeros(39)
    for i in range(39):
        n = S2.mem.get(bb + 8 * i)
        if n is not None and S2.nodes[n][3] != -1: bv[i] = u2f(S2.nodes[n][3])
    for n in S3.nodes:
        if n[0] == "in" and n[1] is not None and bb <= n[1] < bb + 312 and n[3] != -1:
            bv[(n[1] - bb) // 8] = u2f(n[3])
    # an input only ever used as a memory operand never shows in a register: recover it from its products
    for i, n in enumerate(S3.nodes):
        if n[0] == "in" and n[1] is not None and bb <= n[1] < bb + 312 and n[3] == -1:
            uses = [(u2f(S3.nodes[m[1] if m[2] == i else m[2]][3]), u2f(m[3])) for m in S3.nodes
                    if m[0] == "*" and i in (m[1], m[2]) and m[3] != -1]
            x0 = np.float64(uses[0][1]) / np.float64(uses[0][0])
            cands = [x0]
            for _ in range(8):
                cands = cands + [np.nextafter(cands[-1], np.inf)]
                cands = [np.nextafter(cands[0], -np.inf)] + cands
            good = [c for c in cands if all(np.float64(o) * c == np.float64(r) for o, r in uses)]
            assert len(good) >= 1, (i, uses)
            bv[(n[1] - bb) // 8] = good[len(good) // 2]
            print("recovered b[%d] from %d products, %d candidates fit" % ((n[1] - bb) // 8, len(uses), len(good)))
    print("b known", np.count_nonzero(bv))
    poly = {}
    for k in range(11):
        t = C3.flat(cand[cb + 8 * k][1])
        ok = C3.evaluate(t, bv) == u2f(S3.nodes[cand[cb + 8 * k][1]][3])
        print("c", k, len(t), ok)
        poly[k] = t
    pickle.dump({"coef": coef, "poly": poly, "perm": perm, "e": ev, "b": bv,
                 "A": [u2f(S.nodes[outs[k]][3]) for k in range(200)],
                 "c": [u2f(S3.nodes[cand[cb + 8 * k][1]][3]) for k in range(11)]}, open(tr + ".tables.pkl", "wb"))


if __name__ == "__main__":
    main(sys.argv[1])
