import sys, numpy as np, torch, time
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
from tests import util
from forest_slam_b200 import api
from oracle import cv_literal as lit
for (n0,n1,seed) in [(128,128,1),(256,384,2),(1000,777,3),(2048,2048,4)]:
    d0,d1 = util.float_descriptors(n0,n1,seed)
    ref = lit.float_mutual_nn(d0,d1)
    t=time.time()
    a = api.match_mutual_nn(d0,d1,exact_fp32=True)
    b = api.match_mutual_nn(d0,d1,exact_fp32=False)
    torch.cuda.synchronize()
    keep=~ref['ambiguous0']
    print(n0,n1,'fp32 ok',np.array_equal(a['matches0'][keep],ref['matches0'][keep]),'tc ok',np.array_equal(b['matches0'][keep],ref['matches0'][keep]),
          'tc==fp32', np.array_equal(a['matches0'],b['matches0']), 'maxscorediff', np.abs(a['matching_scores0']-b['matching_scores0']).max(), 'nmatch', (b['matches0']>=0).sum(), (a['matches0']>=0).sum())
