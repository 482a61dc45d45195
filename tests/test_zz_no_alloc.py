"""gs_launch_no_alloc: the tile kernel with its scratch (stream-K workspace + hand-off flags) in the library's file-scope
device array instead of per-plan allocations, through the C ABI.  Its output is compared with the existing entry points'
(gs_set_dense + gs_run + gs_get_output on the same plan) within the HALF tolerance the other GPU tests use (1e-2 relative,
floor 2^-25 = half the spacing of fp16 subnormals), on two calls in a row that share the array.

Collected last on purpose; passes alone in a fresh process and after the other tests.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _plan(engine, rows, cols, N, seed, density, as_id):
    from generalsparse_b200 import api, synth
    r, c = synth.synth_matrix("uniform", rows, cols, seed=seed, density=density)
    vals = synth.synth_values(len(r), seed=seed).astype(np.float16).astype(np.float32)
    ms = api.MetaDataSet.from_coo(rows, cols, r, c, vals)
    s = api.Session(ms).run_family("thread_total", N, 8, 1)
    engine.upload_session(s, half=True, as_id=as_id)
    return engine.plan(s, N, half=True, kernel="b200_tcgen05", sub_matrix_id=as_id)


def _close(got, want):
    got, want = got.astype(np.float64), want.astype(np.float64)
    err = np.abs(got - want)
    bad = err > 1e-2 * np.abs(want) + 2.0 ** -25
    assert not bad.any(), "max err %g at %s" % (err.max(), np.argwhere(bad)[0])


def test_no_alloc_entry_point_matches_existing_entry_point(built_lib):
    import torch
    from generalsparse_b200 import api, synth
    engine = api.Engine(0)
    # 3000 x 4096: 24 row panels over 148 CTAs, every CTA leaves or collects partials (workspace + flags in use);
    # 1100 x 1000: another grid and panel count, so the flag words of the first call mean something else in the second
    shapes = [(3000, 4096, 32, 0.1), (1100, 1000, 64, 0.2)]
    plans = [_plan(engine, rows, cols, N, 71 + i, d, 40 + i) for i, (rows, cols, N, d) in enumerate(shapes)]
    Bs = [synth.synth_dense(cols, N, seed=80 + i).astype(np.float16) for i, (rows, cols, N, d) in enumerate(shapes)]
    want = []
    for plan, B in zip(plans, Bs):  # the existing entry points
        plan.set_dense(B)
        plan.run(1)
        want.append(plan.output())
    assert all(np.abs(w).max() > 0 for w in want)
    dev = torch.device("cuda:0")
    Bd = [torch.from_numpy(B).to(dev) for B in Bs]
    Cd = [torch.zeros((rows, N), dtype=torch.float16, device=dev) for (rows, cols, N, d) in shapes]
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.synchronize(dev)
    # two calls in a row on one stream, no host wait in between: the second starts from the first one's scratch contents
    for i in (0, 1):
        plans[i].launch_no_alloc(Bd[i].data_ptr(), Cd[i].data_ptr(), stream.cuda_stream)
    stream.synchronize()
    for i in (0, 1):
        _close(Cd[i].cpu().numpy(), want[i])
    # and again in the other order, into fresh outputs, one of them with another B
    B2 = synth.synth_dense(shapes[0][1], shapes[0][2], seed=99).astype(np.float16)
    plans[0].set_dense(B2)
    plans[0].run(1)
    want2 = plans[0].output()
    B2d = torch.from_numpy(B2).to(dev)
    out1 = torch.zeros_like(Cd[1])
    out0 = torch.zeros_like(Cd[0])
    torch.cuda.synchronize(dev)
    plans[1].launch_no_alloc(Bd[1].data_ptr(), out1.data_ptr(), stream.cuda_stream)
    plans[0].launch_no_alloc(B2d.data_ptr(), out0.data_ptr(), stream.cuda_stream)
    stream.synchronize()
    _close(out1.cpu().numpy(), want[1])
    _close(out0.cpu().numpy(), want2)
    # the plan's own launches still work afterwards (its own workspace and flags were not touched)
    plans[0].run(2)
    assert np.array_equal(plans[0].output(), want2)


def test_no_alloc_entry_point_rejects_what_exceeds_the_array(built_lib):
    import torch
    from generalsparse_b200 import api
    engine = api.Engine(0)
    dev = torch.device("cuda:0")
    wide = _plan(engine, 300, 512, 128, 5, 0.2, 45)  # N = 128 > 64
    B = torch.zeros((512, 128), dtype=torch.float16, device=dev)
    Cc = torch.zeros((300, 128), dtype=torch.float16, device=dev)
    with pytest.raises(api.GsError) as ei:
        wide.launch_no_alloc(B.data_ptr(), Cc.data_ptr(), None)
    assert ei.value.status == 1 and "no-alloc maximum" in str(ei.value)
    wide.set_kernel("b200_sp24")  # another kernel: not this entry point's
    with pytest.raises(api.GsError) as ei:
        wide.launch_no_alloc(B.data_ptr(), Cc.data_ptr(), None)
    assert ei.value.status == 1
