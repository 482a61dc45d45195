import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from generalsparse_b200 import api, synth
eng = api.Engine(0)
N = int(os.environ.get("N", "32"))
rows, cols = 11008, 4096
r, c = synth.synth_matrix("uniform", rows, cols, seed=7, density=0.3)
vals = synth.synth_values(len(r), seed=7).astype(np.float16).astype(np.float32)
B = synth.synth_dense(cols, N, seed=7).astype(np.float16).astype(np.float32)
ms = api.MetaDataSet.from_coo(rows, cols, r, c, vals)
s = api.Session(ms).run_family("thread_total", N, 8, 1)
eng.upload_session(s, half=True)
plan = eng.plan(s, N, half=True, kernel="b200_tcgen05")
plan.set_dense(B)
print("ms/launch", plan.run(120))
