import time, sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
import orientedobjectdetection_b200 as rg
from orientedobjectdetection_b200 import synth, core, _hostpack
patches, offs, shape = synth.c5_patch_results(0)
ts=[]; th=[]
for r in range(8):
    cp = [[a.copy() for a in p] for p in patches]
    torch.cuda.synchronize(); t=time.perf_counter()
    out = rg.merge_results(cp, offs, shape, 0.1)
    ts.append(time.perf_counter()-t)
    cp = [[a.copy() for a in p] for p in patches]
    o64=np.ascontiguousarray(offs,dtype=np.float64); st=core._stage_rows(1).numpy()
    t=time.perf_counter(); _hostpack.gather_shift(cp,o64,st); th.append(time.perf_counter()-t)
print('merge_results ms', 1e3*min(ts), 'helper ms', 1e3*min(th), 'kept', sum(o.shape[0] for o in out))
import cProfile, pstats
cp = [[a.copy() for a in p] for p in patches]
pr=cProfile.Profile(); pr.enable(); rg.merge_results(cp, offs, shape, 0.1); pr.disable()
pstats.Stats(pr).sort_stats('cumulative').print_stats(18)
