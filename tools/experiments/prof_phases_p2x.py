"""Per-warp phase cycle counters of the fused pass 2 + spectrum experiment (pass2x.cuh wired into
search.cu and built with `make EXTRA=-DPS_PROF`; see profiles/r02_summary.md r2b)."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import peasoup_b200 as pb
from peasoup_b200.host import make_search_params
import bench
torch.cuda.set_device(0)
d = bench.synth_trials(1, torch.device("cuda:0"), 0)
s = pb.Search(make_search_params(bench.SIZE, bench.NSAMPS, bench.TSAMP, bench.NFOLDS, 9.0))
s.set_zaplist(bench.ZAP_F, bench.ZAP_W)
s.whiten(d[0]); s.reset_peaks(); s.accel_batch(bench.ACCS[:26]); torch.cuda.synchronize()
s.accel_batch(bench.ACCS[:26]); torch.cuda.synchronize()
buf = np.zeros(148 * 16 * 16, np.uint32)
lib = pb.load()
lib.ps_prof_read.argtypes = [C.c_void_p]
assert lib.ps_prof_read(buf.ctypes.data) == 0
b = buf.reshape(148, 16, 16).astype(np.float64)
pairs = 26 * 256 / 148
names = ["issue L copy", "wait U tile", "U take + phase A", "barrier + U phase B", "wait L tile", "take L + barrier + U phase C",
         "L phase A", "barrier + L phase B", "barrier + L phase C", "barrier + split", "barrier + interbin/stores"]
m = b.mean(axis=(0, 1)) / pairs
for n, v in zip(names, m):
    print(f"{n:32s} {v:8.0f} cycles per pair")
print("sum", int(m.sum()))
