import numpy as np, json, sys
from generalsparse_b200 import api
p=sys.argv[1]
z=np.load(p); meta=json.loads(str(z["__meta__"]))
rows,cols=[int(x) for x in z["__shape__"]]
ms=api.MetaDataSet.from_coo(rows,cols,z["__coo_row__"],z["__coo_col__"]); s=api.Session(ms)
api.config_set("DENSE_MATRIX_SIZE", meta["N"])
try:
    for op in meta["ops"]: s.add_and_run(op[0], *op[1:])
except Exception as e: print("EXC",e)
ours=set(ms.keys()); theirs=set(k for k in z.files if not k.startswith("__"))
print("only ours",sorted(ours-theirs),"only ref",sorted(theirs-ours))
for k in sorted(ours&theirs):
    if ms[k].shape!=z[k].shape or not np.array_equal(ms[k],z[k]): print("DIFF",k,ms[k].shape,z[k].shape, ms[k][:12], z[k][:12])
    elif ms.info(k)[2]!=meta["dtypes"][k]: print("DTYPE",k,ms.info(k)[2],meta["dtypes"][k])
ref=[l.replace('data_transform-->','') for l in meta["steps"] if '::{' in l and not l.startswith(('input','output','valid'))]
o=s.records()
import difflib
for l in difflib.unified_diff(ref,o,lineterm='',n=0): print(l[:200])
