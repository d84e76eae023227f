"""Per-warp phase cycle counters of the warp-specialised pass 1 (build with `make EXTRA=-DPS_PROF`)."""
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
buf = np.zeros(148 * 20 * 8, np.uint32)
lib = pb.load()
lib.ps_prof_read.argtypes = [C.c_void_p]
assert lib.ps_prof_read(buf.ctypes.data) == 0
b = buf.reshape(148, 20, 8).astype(np.float64)
items = 26 * (bench.SIZE // 2 // 2048 // 4) / 148
print("items per CTA %.1f" % items)
names_f = ["wait FULL", "take", "phase A", "barrier 1", "phase B", "barrier 2", "phase C", "-"]
names_l = ["wait EMPTY", "issue", "wait copies", "-", "-", "-", "-", "-"]
f = b[:, :16, :].mean(axis=(0, 1)) / items
l = b[:, 16:, :].mean(axis=(0, 1)) / items
print("FFT warps, cycles per item:", {n: int(v) for n, v in zip(names_f, f) if n != "-"}, "sum", int(f.sum()))
print("loader warps, cycles per item:", {n: int(v) for n, v in zip(names_l, l) if n != "-"}, "sum", int(l.sum()))
