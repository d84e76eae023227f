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
s.whiten(d[0]); torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
s.whiten(d[0]); torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
