"""Small driver for ncu: one DM trial whitened + ONE n-trial ps_search_accel_batch at the bench size.
Only that last batch lies between cudaProfilerStart/Stop, so with `ncu --profile-from-start off` the
report holds exactly one launch of each chain kernel (n trials each): the launch list bench.py reads
its DRAM traffic from (profiles/r*_chain<n>_ncu.csv)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import peasoup_b200 as pb
from peasoup_b200.host import make_search_params
import bench

n_acc = int(sys.argv[1]) if len(sys.argv) > 1 else 26
torch.cuda.set_device(0)
d = bench.synth_trials(1, torch.device("cuda:0"), 0)
s = pb.Search(make_search_params(bench.SIZE, bench.NSAMPS, bench.TSAMP, bench.NFOLDS, 9.0))
s.set_zaplist(bench.ZAP_F, bench.ZAP_W)
s.whiten(d[0]); s.reset_peaks(); s.accel_batch(bench.ACCS[:n_acc])
torch.cuda.synchronize()
s.reset_peaks()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
s.accel_batch(bench.ACCS[:n_acc])
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("peaks", len(s.fetch_peaks()))
