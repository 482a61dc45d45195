"""C2 N=8 (2:4 image kernel): one launch alone / a dependent chain / back to back, under the env knobs of kernels_sp24.cu (one process per setting)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from generalsparse_b200 import api, synth

N = int(os.environ.get("ISO_N", 8))
rows, cols = 11008, 4096
r, c = synth.synth_matrix("nm24", rows, cols, seed=3) if os.environ.get("ISO_PATTERN") == "nm24" else synth.synth_matrix("uniform", rows, cols, seed=2, density=float(os.environ.get("ISO_DENSITY", 0.30)))
v = synth.synth_values(len(r), seed=2).astype(np.float16).astype(np.float32)
s = api.Session(api.MetaDataSet.from_coo(rows, cols, r, c, v)).run_family("thread_total", N, 8, 1)
eng = api.Engine(0)
plans = []
for k in range(8):
    eng.upload_session(s, half=True, as_id=k)
    p = eng.plan(s, N, half=True, kernel="auto", sub_matrix_id=k)
    p.set_dense(synth.synth_dense(cols, N, seed=3).astype(np.float16))
    plans.append(p)
stream = torch.cuda.ExternalStream(eng.stream)
def ev():
    return torch.cuda.Event(enable_timing=True)
with torch.cuda.stream(stream):
    for p in plans:
        p.launch()
    torch.cuda.synchronize()
    iso = []
    for i in range(64):
        torch.cuda.synchronize()
        a, b = ev(), ev()
        a.record(stream); plans[i % 8].launch(); b.record(stream)
        torch.cuda.synchronize()
        iso.append(a.elapsed_time(b) * 1e3)
    a, b = ev(), ev()
    a.record(stream)
    for i in range(2000):
        plans[i % 8].launch()
    b.record(stream); torch.cuda.synchronize()
    b2b = a.elapsed_time(b)
    def dep(i):
        plans[i % 8].set_dense_device(plans[(i - 1) % 8].output_device())
        plans[i % 8].launch()
    for i in range(1, 9):
        dep(i)
    torch.cuda.synchronize()
    a, b = ev(), ev()
    a.record(stream)
    for i in range(1, 1001):
        dep(i)
    b.record(stream); torch.cuda.synchronize()
    depc = a.elapsed_time(b)
print("%-40s iso median %.1f min %.1f us | b2b %.2f us | dependent %.2f us | %s" % (os.environ.get("TAG", ""), np.median(iso), min(iso), b2b / 2, depc, plans[0].kernel_name[:110]))
