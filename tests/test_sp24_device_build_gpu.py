"""The 2:4 image built on the device at plan time (csrc/device/sp24_build.cu) against the host builder (csrc/host/sp24_format.cc, GS_SP24_DEVICE_BUILD=0):
the same kernel (unit size, tiles kept when the empty ones leave the stream, remainder share), the same C bit for bit in reference mode (all-ones: every sum is
exact) and within rounding of the summation order on random values (the host orders a row's tile-local entries against bank conflicts, the device lists them in
column order), and both inside SURVEY 8(d)'s gate against the fp64 oracle."""
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine(built_lib):
    from generalsparse_b200 import api
    return api.Engine(0)


def _cases():
    from generalsparse_b200 import synth
    return [("uniform 30 %", 1000, 4096, synth.synth_matrix("uniform", 1000, 4096, seed=83, density=0.3)),
            ("2:4", 1000, 4096, synth.synth_matrix("nm24", 1000, 4096, seed=81)),
            ("banded (most tiles empty)", 3000, 3000, synth.synth_matrix("banded", 3000, 3000, seed=5, half_bandwidth=40)),
            ("ragged, empty rows, K tail", 300, 1000, synth.synth_matrix("ragged", 300, 1000, seed=84, mean_len=20.0, empty_frac=0.2)),
            ("smaller than two tiles", 130, 72, synth.synth_matrix("nm24", 130, 72, seed=82))]


@pytest.mark.parametrize("N", [8, 16, 32, 64, 128])
@pytest.mark.parametrize("chain", ["thread_total", "warp_total"])
def test_device_built_image_equals_host_built(engine, monkeypatch, N, chain):
    from generalsparse_b200 import api, synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_absrow, oracle_spmm_f32, oracle_spmm_f64acc
    for name, rows, cols, (r, c) in _cases():
        ip = csr_from_coo(rows, r)
        for ones in (True, False):
            vals = None if ones else synth.synth_values(len(r), seed=5).astype(np.float16).astype(np.float32)
            B = np.ones((cols, N), np.float32) if ones else synth.synth_dense(cols, N, seed=5).astype(np.float16).astype(np.float32)
            s = api.Session(api.MetaDataSet.from_coo(rows, cols, r, c, vals)).run_family(chain, N, 8 if chain == "thread_total" else 1, 1)
            engine.upload_session(s, half=True, as_id=90)
            out = {}
            for mode in ("0", "1"):
                monkeypatch.setenv("GS_SP24_DEVICE_BUILD", mode)
                plan = engine.plan(s, N, half=True, kernel="b200_sp24", sub_matrix_id=90)
                plan.set_dense(B)
                plan.run(2)
                out[mode] = (plan.output().astype(np.float32), plan.kernel_name, plan.format_bytes)
                del plan
            (Ch, kh, fh), (Cd, kd, fd) = out["0"], out["1"]
            assert kh == kd and fh == fd, (name, kh, kd)
            if ones:
                ref = oracle_spmm_f32(rows, N, ip, c, np.ones(len(r), np.float32), B)
                if ref.max() <= 2048:
                    assert np.array_equal(Cd, ref), (name, kd)
                assert np.array_equal(Ch, Cd), (name, kd)
            else:
                ref, absrow = oracle_spmm_f64acc(rows, N, ip, c, vals, B), oracle_spmm_absrow(rows, N, ip, c, vals, B)
                for C in (Ch, Cd):
                    assert (np.abs(C - ref) <= 1e-2 * np.abs(ref) + 1e-4 * absrow + 2.0 ** -25).all(), (name, kd)


def test_device_build_is_faster_than_host_build(engine, monkeypatch):
    """the C2 weight: plan creation with the host builder (download, build, upload) against the device passes"""
    from generalsparse_b200 import api, synth
    rows, cols, N = 11008, 4096, 8
    r, c = synth.synth_matrix("uniform", rows, cols, seed=2, density=0.30)
    vals = synth.synth_values(len(r), seed=2).astype(np.float16).astype(np.float32)
    desc, _ = engine.transform_coo("thread_total", rows, cols, r, c, vals, N, 8, 1, half=True, as_id=91)
    t = {}
    for mode in ("0", "1", "1"):
        monkeypatch.setenv("GS_SP24_DEVICE_BUILD", mode)
        t0 = time.perf_counter()
        plan = engine.plan(desc, N, half=True, kernel="b200_sp24")
        t[mode] = time.perf_counter() - t0
        name = plan.kernel_name
        del plan
    print("plan creation, C2 N=8: host builder %.3f s, device builder %.4f s (%s)" % (t["0"], t["1"], name))
    assert t["1"] < t["0"]


def test_failed_device_build_falls_back_to_the_host_builder(engine, monkeypatch):
    """GS_SP24_DEVICE_BUILD=2 makes the fill pass report an allocation failure after the counts were taken: the plan must come out of the host builder"""
    from generalsparse_b200 import api, synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_f32
    rows, cols, N = 1000, 4096, 16
    r, c = synth.synth_matrix("uniform", rows, cols, seed=83, density=0.3)
    s = api.Session(api.MetaDataSet.from_coo(rows, cols, r, c, None)).run_family("thread_total", N, 8, 1)
    engine.upload_session(s, half=True, as_id=92)
    names = {}
    for mode in ("0", "2"):
        monkeypatch.setenv("GS_SP24_DEVICE_BUILD", mode)
        plan = engine.plan(s, N, half=True, kernel="b200_sp24", sub_matrix_id=92)
        plan.set_dense(np.ones((cols, N), np.float32))
        plan.run(1)
        assert np.array_equal(plan.output().astype(np.float32), oracle_spmm_f32(rows, N, csr_from_coo(rows, r), c, np.ones(len(r), np.float32), np.ones((cols, N), np.float32)))
        names[mode] = plan.kernel_name
        del plan
    assert names["0"] == names["2"]
