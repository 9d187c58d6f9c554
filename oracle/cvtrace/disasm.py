"""Map a recorded trace to disassembly: writes TRACE.va.npy (virtual address per step) and TRACE.ins.pkl
(address -> instruction text, objdump -M intel).  TEST-INFRASTRUCTURE tool, see tracer.c."""
import numpy as np, subprocess, sys, re, pickle, os, cv2
SO = os.path.join(os.path.dirname(cv2.__file__), "cv2.abi3.so")
DT = np.dtype([("rip", "<u8"), ("gpr", "<u8", 16), ("xmm", "<u8", (16, 2))])


def main(tr):
    T = np.fromfile(tr, dtype=DT)
    segs = []
    for l in open(tr + ".maps").read().splitlines():
        if "cv2.abi3.so" in l:
            a, b = [int(x, 16) for x in l.split()[0].split("-")]
            segs.append((a, b, int(l.split()[2], 16), l.split()[1]))
    a, b, off, _ = [s for s in segs if "x" in s[3]][0]
    va = T["rip"] - (a - off)
    u = np.unique(va)
    ranges = []
    s = p = u[0]
    for x in u[1:]:
        if x - p > 256:
            ranges.append((s, p)); s = x
        p = x
    ranges.append((s, p))
    np.save(tr + ".va.npy", va)
    ins = {}
    for (s, e) in ranges:
        o = subprocess.run(["objdump", "-d", "-M", "intel", "--no-show-raw-insn", f"--start-address={s}",
                            f"--stop-address={e + 16}", SO], capture_output=True, text=True).stdout
        for line in o.splitlines():
            m = re.match(r"\s*([0-9a-f]+):\s+(.*)$", line)
            if m:
                ins[int(m.group(1), 16)] = m.group(2).strip()
    missing = [x for x in u if int(x) not in ins]
    print("steps", len(T), "unique", len(u), "ranges", len(ranges), "missing", len(missing))
    pickle.dump(ins, open(tr + ".ins.pkl", "wb"))


if __name__ == "__main__":
    main(sys.argv[1])
