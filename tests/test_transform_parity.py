"""Bit-exact parity of every transform_step metadata array with the reference pipeline.

tests/golden/*.npz were produced by running the UNMODIFIED reference operator chains
(token_test.cc:1003-1582) through oracle/ref_harness.cc (tests/golden/gen_golden.py); each case
holds every array the reference left in its meta_data_set, the dtype its compression rule chose
(code_source_data.cc:383-416) and the list of arrays its code generator registered as kernel inputs.
"""
import glob
import json
import os

import numpy as np
import pytest

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def _load(path):
    z = np.load(path)
    meta = json.loads(str(z["__meta__"]))
    return z, meta


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[:-4] for p in GOLD])
def test_family_chain_matches_reference(path, built_lib):
    from generalsparse_b200 import api
    z, meta = _load(path)
    rows, cols = [int(x) for x in z["__shape__"]]
    ms = api.MetaDataSet.from_coo(rows, cols, z["__coo_row__"], z["__coo_col__"])  # vals=None -> all ones, like the reference
    s = api.Session(ms)
    s.run_family(meta["family"], meta["N"], meta["S"], meta["C"])
    ours = set(ms.keys())
    theirs = set(k for k in z.files if not k.startswith("__"))
    assert ours == theirs, (sorted(ours - theirs), sorted(theirs - ours))
    for k in sorted(theirs):
        got = ms[k]
        ref = z[k]
        assert got.shape == ref.shape, (k, got.shape, ref.shape)
        assert np.array_equal(got, ref), (k, np.nonzero(got != ref)[0][:5])
        n, mem, comp = ms.info(k)
        assert comp == meta["dtypes"][k], (k, comp, meta["dtypes"][k])
    # kernel inputs: the arrays the reference's code generator registered (its data_source/<id>/ files)
    assert sorted(s.deps()) == sorted(meta["registered"])
    assert int(api.config_get("VECTOR_WIDTH")) == meta["vector_width"]
