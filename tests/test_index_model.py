"""Model-driven index compression (arr_optimization.hpp:13-85): the five closed forms, fitted with the reference's
acceptance rules, on the arrays the chains really produce."""
import numpy as np
import pytest

from generalsparse_b200 import api, synth


def test_models_on_chain_arrays(built_lib):
    r, c = synth.synth_matrix("uniform", 344, 512, seed=4, density=0.3)
    ms = api.MetaDataSet.from_coo(344, 512, r, c)
    api.Session(ms).run_family("thread_total", 8, 4, 1)
    assert ms.fit_model("THREAD_META_first_row_indices_0") == ("linear", 1, 0, 0)          # one row per block: i
    assert ms.fit_model("THREAD_META_first_nz_indices_0")[0] == "none"                     # a real row pointer
    assert ms.fit_model("GLOBAL_META_begin_row_index_0")[0] == "constant"
    ms = api.MetaDataSet.from_coo(344, 512, r, c)
    api.Session(ms).run_family("thread_bit_map", 32, 1, 1)
    t, coef, icpt, _ = ms.fit_model("THREAD_META_first_nz_indices_0")                      # 32 non-zeros per BMT, stream padded: 32 * i
    assert (t, coef, icpt) == ("linear", 32, 0)
    assert ms.fit_model("GLOBAL_META_BMT_size_of_each_blk_0") == ("constant", 0, 32, 0)
    ms = api.MetaDataSet.from_coo(344, 512, r, c)
    s = api.Session(ms)
    s.add_and_run("fixed_interval_row_direction_tblock_blocking_operator", 8, 0)
    s.add_and_run("fixed_interval_row_direction_warp_blocking_operator", 2, 1, 1, 0)
    assert ms.fit_model("WARP_META_first_row_indices_relative_to_BMTB_0") == ("cycle_linear", 2, 0, 4)   # 0 2 4 6 | 0 2 4 6 ...
    assert ms.fit_model("TBLOCK_META_first_BMW_indices_0") == ("linear", 4, 0, 0)


def test_model_rules(built_lib):
    def fit(a):
        n = len(a)
        ms = api.MetaDataSet.from_coo(max(a) + 2, 4, np.zeros(1, np.int64), np.zeros(1, np.int64))
        return ms
    # through the arrays of a hand-made chain is clumsy: the rules are covered on synthetic arrays via the step that stores any u64 item
    r = np.repeat(np.arange(12), 2)
    ms = api.MetaDataSet.from_coo(12, 4, r, np.tile([0, 1], 12))
    s = api.Session(ms)
    s.add_and_run("fixed_interval_nnz_direction_thread_blocking_operator", 4, 0, 0, 0)
    assert ms.fit_model("THREAD_META_first_row_indices_0") == ("linear", 2, 0, 0)          # rows 0 2 4 ... 12
    s2 = api.Session(api.MetaDataSet.from_coo(12, 4, r, np.tile([0, 1], 12)))
    s2.add_and_run("fixed_interval_col_direction_thread_blocking_operator", 1, 0, 0, 0, 0)
    assert s2.ms.fit_model("THREAD_META_first_row_indices_without_ending_0") == ("cycle_increase", 0, 0, 2)   # 0 0 1 1 2 2 ...
    with pytest.raises(api.GsError):
        ms.fit_model("GLOBAL_META_nz_vals_0")


@pytest.mark.gpu
def test_nnz_kernel_computes_linear_first_nz(built_lib, monkeypatch):
    """b200_nnz takes first_nz_indices = 32 * i as a formula; with the model switched off (GS_NO_INDEX_MODEL) the results are the same bits"""
    eng = api.Engine(0)
    r, c = synth.synth_matrix("powerlaw", 3000, 3000, seed=5, avg_deg=20.0, alpha=2.1)
    vals = synth.synth_values(len(r), seed=3)
    B = synth.synth_dense(3000, 32, seed=3)
    outs = []
    for off in (False, True):
        if off:
            monkeypatch.setenv("GS_NO_INDEX_MODEL", "1")
        s = api.Session(api.MetaDataSet.from_coo(3000, 3000, r, c, vals)).run_family("thread_bit_map", 32, 1, 1)
        eng.upload_session(s, as_id=80)
        plan = eng.plan(s, 32, kernel="b200_nnz", sub_matrix_id=80)
        assert plan.model_bytes > 0
        plan.set_dense(B)
        plan.run(1)
        outs.append(plan.output())
    assert np.allclose(outs[0], outs[1], rtol=1e-5, atol=1e-6)
