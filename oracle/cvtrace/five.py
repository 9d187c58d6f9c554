"""Traced workload: ONE cv2.findEssentialMat call on exactly five correspondences (runs the minimal solver once and
returns every candidate) between two SIGUSR1 markers.  usage: tracer out.bin python five.py SEED OUTDIR"""
import os, signal, sys
import numpy as np, cv2
cv2.setNumThreads(1)
seed = int(sys.argv[1]) if len(sys.argv) > 1 else 0
outdir = sys.argv[2] if len(sys.argv) > 2 else "/tmp/tr"
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from tests import util
p0, p1, _, _, _ = util.two_view(5, 1000 + seed, outliers=0.0)
q1 = np.ascontiguousarray(util.normalise(p0)); q2 = np.ascontiguousarray(util.normalise(p1))
np.save(f"{outdir}/q_{seed}.npy", np.stack([q1, q2]))
kw = dict(focal=1.0, pp=(0, 0), method=cv2.RANSAC, prob=0.999, threshold=1.0)
cv2.findEssentialMat(q1, q2, **kw)          # warm-up: lazy initialisation stays out of the trace
os.kill(os.getpid(), signal.SIGUSR1)
E, m = cv2.findEssentialMat(q1, q2, **kw)
os.kill(os.getpid(), signal.SIGUSR1)
np.save(f"{outdir}/E_{seed}.npy", E)
print(E.shape)
