"""show.py/timeline helpers: python tl.py TRACE show A B | uniq A B | timeline A B SHIFT MINLEN"""
import numpy as np, pickle, sys, collections
tr, cmd = sys.argv[1], sys.argv[2]
va = np.load(tr + ".va.npy"); ins = pickle.load(open(tr + ".ins.pkl", "rb"))
a, b = int(sys.argv[3]), int(sys.argv[4])
FP = {"mulsd", "addsd", "subsd", "divsd", "sqrtsd", "mulpd", "addpd", "subpd", "divpd"}
if cmd == "show":
    for k in range(a, b): print(k, hex(int(va[k])), ins[int(va[k])].split("<")[0])
elif cmd == "uniq":
    c = collections.Counter(va[a:b].tolist())
    for v in sorted(c): print(hex(v), c[v], ins[v].split("<")[0])
elif cmd == "timeline":
    sh, minlen = int(sys.argv[5]), int(sys.argv[6])
    pg = va >> sh
    k = a
    while k < b:
        j = k
        while j < b and pg[j] == pg[k]: j += 1
        if j - k >= minlen:
            fp = sum(1 for q in range(k, j) if ins[int(va[q])].split()[0] in FP)
            print(k, j - k, hex(int(pg[k]) << sh), "fp", fp)
        k = j
