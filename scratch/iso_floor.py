"""what an event pair around ONE launch on an idle device costs whatever the kernel does: a 256 x 256 matrix on three kernels"""
import numpy as np, torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from generalsparse_b200 import api, synth
r, c = synth.synth_matrix("uniform", 256, 256, seed=1, density=0.05)
s = api.Session(api.MetaDataSet.from_coo(256, 256, r, c, None)).run_family("thread_total", 8, 8, 1)
eng = api.Engine(0)
for kern, half in (("b200_sp24", True), ("b200_row_smem", False), (None, False)):
    eng.upload_session(s, half=half, as_id=0)
    p = eng.plan(s, 8, half=half, kernel=kern, sub_matrix_id=0)
    p.set_dense(np.ones((256, 8), np.float16 if half else np.float32))
    st = torch.cuda.ExternalStream(eng.stream)
    iso = []
    with torch.cuda.stream(st):
        for i in range(5):
            p.launch()
        for i in range(64):
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(st); p.launch(); b.record(st)
            torch.cuda.synchronize()
            iso.append(a.elapsed_time(b) * 1e3)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record(st); b.record(st); torch.cuda.synchronize()
        empty = a.elapsed_time(b) * 1e3
    print("tiny matrix, one launch alone: %-14s median %.1f min %.1f us (two events with nothing between: %.1f us)  %s" % (kern, np.median(iso), min(iso), empty, p.kernel_name[:60]))
    del p
