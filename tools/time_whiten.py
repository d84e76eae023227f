"""Times the per-DM whitening stage (ps_search_whiten) with CUDA events on its stream."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import peasoup_b200 as pb
from peasoup_b200.host import make_search_params
import bench
torch.cuda.set_device(0)
d = bench.synth_trials(2, torch.device("cuda:0"), 0)
s = pb.Search(make_search_params(bench.SIZE, bench.NSAMPS, bench.TSAMP, bench.NFOLDS, 9.0))
s.set_zaplist(bench.ZAP_F, bench.ZAP_W)
st = torch.cuda.Stream()
for i in range(3):
    s.whiten(d[i % 2], stream=st)
st.synchronize()
best = 1e9
for rep in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st); s.whiten(d[rep % 2], stream=st); e1.record(st); st.synchronize()
    best = min(best, e0.elapsed_time(e1))
print(f"whiten {best*1e3:.0f} us per DM trial = {best*1e3/len(bench.ACCS):.2f} us per acceleration trial at {len(bench.ACCS)} accelerations")
