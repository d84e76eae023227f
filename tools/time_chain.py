"""Times the acceleration-trial kernel chain (CUDA events on its stream) for quick A/B runs."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import peasoup_b200 as pb
from peasoup_b200.host import make_search_params
import bench
torch.cuda.set_device(0)
d = bench.synth_trials(1, torch.device("cuda:0"), 0)
s = pb.Search(make_search_params(bench.SIZE, bench.NSAMPS, bench.TSAMP, bench.NFOLDS, 9.0))
s.set_zaplist(bench.ZAP_F, bench.ZAP_W)
st = torch.cuda.Stream()
s.whiten(d[0], stream=st); s.reset_peaks(stream=st); s.accel_batch(bench.ACCS[:10], stream=st); st.synchronize()
best = 1e9
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st); s.accel_batch(bench.ACCS, stream=st); e1.record(st); st.synchronize()
    best = min(best, e0.elapsed_time(e1) / len(bench.ACCS))
print(f"chain {best*1e3:.1f} us/trial  ({os.environ.get('PS_HP_VARIANT','-')})")
