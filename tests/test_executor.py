"""The multi-GPU executor (gs_exec_*, north star (3)): nnz-balanced contiguous row blocks from the reference's balanced-interval
rule, one block per GPU, B replicated, C rows disjoint, optional NCCL all-gather of C.

CPU: the host half -- gsh_metadata_row_nnz / gsh_metadata_row_block against numpy, blocks reassemble the matrix.
GPU: the executor on 1 device always, on 2+ devices where the box has them (gpurun --gpus 2): parity with the oracle, the
all-gather replicas equal C bit for bit, and the `token_test --gpus G` driver."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOKEN_TEST = os.path.join(ROOT, "generalsparse_b200", "lib", "token_test")


def _matrix():
    from generalsparse_b200 import synth
    return synth.synth_matrix("powerlaw", 3000, 3000, seed=5, avg_deg=20.0, alpha=2.1)


def test_row_blocks_reassemble_the_matrix(built_lib):
    from generalsparse_b200 import api, shard
    r, c = _matrix()
    vals = np.arange(len(r), dtype=np.float32) % 97 - 48
    ms = api.MetaDataSet.from_coo(3000, 3000, r, c, vals)
    row_nnz = ms.row_nnz()
    assert np.array_equal(row_nnz, np.bincount(r, minlength=3000).astype(np.uint64))
    for G in (1, 2, 3, 8):
        cuts = shard.row_cuts(row_nnz, G)
        assert cuts[0] == 0 and cuts[-1] == 3000 and (np.diff(cuts) >= 0).all()
        got_r, got_c, got_v = [], [], []
        for g in range(G):
            lo, hi = int(cuts[g]), int(cuts[g + 1])
            if lo == hi:
                continue
            blk = ms.row_block(lo, hi)
            assert int(blk["GLOBAL_META_origin_row_num_-1"][0]) == hi - lo and int(blk["GLOBAL_META_origin_col_num_-1"][0]) == 3000
            got_r.append(blk["GLOBAL_META_nz_row_indices_0"].astype(np.int64) + lo)
            got_c.append(blk["GLOBAL_META_nz_col_indices_0"].astype(np.int64))
            got_v.append(blk["GLOBAL_META_nz_vals_0"])
        assert np.array_equal(np.concatenate(got_r), r) and np.array_equal(np.concatenate(got_c), c)
        assert np.array_equal(np.concatenate(got_v), vals.astype(np.float64))
        # the greedy >= rule: no block but the last exceeds ceil(nnz / G) by more than its last row
        per = -(-len(r) // G)
        for g in range(G - 1):
            lo, hi = int(cuts[g]), int(cuts[g + 1])
            if hi > lo:
                assert row_nnz[lo:hi].sum() - row_nnz[hi - 1] < per
    with pytest.raises(api.GsError):
        ms.row_block(10, 10)
    api.Session(ms).add_and_run("sort_operator")
    with pytest.raises(api.GsError):  # blocks are cut from the matrix as loaded
        ms.row_block(0, 10)


def _devices():
    import torch
    return torch.cuda.device_count()


@pytest.mark.gpu
@pytest.mark.parametrize("G", [1, 2, 4, 8])
@pytest.mark.parametrize("family,kernel,N,half", [("thread_total", "auto", 8, True), ("warp_segment", "auto", 32, False), ("thread_bit_map", "chain", 8, False),
                                                 ("warp_total", "chain", 4, False)])
def test_executor_parity_and_allgather(built_lib, G, family, kernel, N, half):
    from generalsparse_b200 import api, synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f64acc
    if _devices() < G:
        pytest.skip("needs %d GPUs" % G)
    r, c = _matrix()
    rows = cols = 3000
    vals = synth.synth_values(len(r), seed=3)
    B = synth.synth_dense(cols, N, seed=3)
    if half:
        vals, B = vals.astype(np.float16).astype(np.float32), B.astype(np.float16).astype(np.float32)
    ms = api.MetaDataSet.from_coo(rows, cols, r, c, vals)
    ex = api.Executor(list(range(G))).build(ms, family, N, 8 if family == "thread_total" else 1, 1, half=half, kernel=kernel)
    cuts, bnnz = ex.row_cuts()
    assert int(bnnz.sum()) == len(r) and bnnz.max() <= -(-len(r) // G) + np.bincount(r).max()
    ex.set_dense(B)
    ms_max, per = ex.run(3)
    assert ms_max > 0 and ms_max == per.max()
    C = ex.output()
    ip = csr_from_coo(rows, r)
    ref = oracle_spmm_f64acc(rows, N, ip, c.astype(np.int32), vals, B)
    absrow = oracle_spmm_absrow(rows, N, ip, c.astype(np.int32), vals, B)
    tol = 1e-2 if half else 1e-5
    loose = half and family in ("warp_segment", "thread_bit_map")  # fp16 partial sums added into an fp16 C (test_spmm_gpu.py): row-L1 floor at the tolerance
    assert (np.abs(C - ref) <= tol * np.abs(ref) + (tol if loose else 1e-6) * absrow + 2.0 ** -25).all(), ex.kernel_name()
    # C replicated on every device: bit-equal to the gathered C
    ex.enable_allgather()
    ex.run(2)
    C2 = ex.output()
    ag_ms = ex.allgather_output()
    assert ag_ms > 0
    for g in range(G):
        assert np.array_equal(ex.replica(g), C2), g
    if "nnz" not in ex.kernel_name() and family not in ("thread_bit_map",):
        assert np.array_equal(C2, C)  # deterministic kernels: writing in place into the full C changes nothing


@pytest.mark.gpu
@pytest.mark.parametrize("G", [1, 2, 8])
@pytest.mark.parametrize("family,kernel,N,half", [("thread_total", "auto", 8, True), ("warp_segment", "chain", 32, False), ("thread_bit_map", "auto", 32, False), ("warp_total", "chain", 4, False)])
def test_executor_blocks_built_on_device(built_lib, G, family, kernel, N, half):
    """gs_exec_build_on_device: every block's chain comes from its GPU's transform passes; same cuts, same kernels, same C as the
    blocks built by the host pipeline (bit-equal where the kernel is deterministic)"""
    from generalsparse_b200 import api, synth
    if _devices() < G:
        pytest.skip("needs %d GPUs" % G)
    r, c = _matrix()
    rows = cols = 3000
    vals = synth.synth_values(len(r), seed=3)
    B = synth.synth_dense(cols, N, seed=3)
    if half:
        vals, B = vals.astype(np.float16).astype(np.float32), B.astype(np.float16).astype(np.float32)
    ms = api.MetaDataSet.from_coo(rows, cols, r, c, vals)
    S = 8 if family == "thread_total" else 1
    eh = api.Executor(list(range(G))).build(ms, family, N, S, 1, half=half, kernel=kernel)
    ed = api.Executor(list(range(G))).build(ms, family, N, S, 1, half=half, kernel=kernel, on_device=True)
    assert np.array_equal(eh.row_cuts()[0], ed.row_cuts()[0]) and np.array_equal(eh.row_cuts()[1], ed.row_cuts()[1])
    assert [eh.kernel_name(g) for g in range(G)] == [ed.kernel_name(g) for g in range(G)]
    eh.set_dense(B), ed.set_dense(B)
    eh.run(2), ed.run(2)
    Ch, Cd = eh.output().astype(np.float32), ed.output().astype(np.float32)
    if family in ("thread_bit_map", "warp_segment") or "merge" in ed.kernel_name(0) or "nnz" in ed.kernel_name(0):
        assert np.allclose(Ch, Cd, rtol=1e-5, atol=1e-5)
    else:
        assert np.array_equal(Ch, Cd)


@pytest.mark.gpu
@pytest.mark.parametrize("G", [2, 8])
def test_token_test_gpus(built_lib, tmp_path, G):
    """`token_test c1.mtx 8 --gpus G --allgather`: the driver's multi-GPU switch is functional (round 1 swallowed it)"""
    from generalsparse_b200 import synth
    if _devices() < G:
        pytest.skip("needs %d GPUs" % G)
    rows, cols, r, c = synth.c1_standin()
    path = str(tmp_path / "c1.mtx")
    synth.write_mtx(path, rows, cols, r, c)
    for extra in ([], ["--half"], ["--family", "warp_segment", "--kernel", "reference"]):
        p = subprocess.run([TOKEN_TEST, path, "8", "--gpus", str(G), "--allgather", "--repeat", "50"] + extra, cwd=str(tmp_path), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
        assert p.returncode == 0, p.stdout[-1500:]
        assert "\ncorrect\n" in p.stdout and "0 of %d replicas differ" % G in p.stdout, p.stdout[-1500:]
        assert p.stdout.count("gpu ") >= G
