"""BASELINE.json configs[0] (C1) through the real driver: `token_test <matrix.mtx> 8` on the IG5-18 stand-in.

CPU part (loader + pipeline pin): the stand-in is written as the .mtx TEXT the reference loader reads (1-based, `%`
comments, `rows cols nnz` header, row-sorted; struct.cc:49-261, metadata_set.cc:612-707), gsh_metadata_from_mtx reads it,
each of the seven driver chains runs, and the dump (gsh_metadata_dump, the reference's one-value-per-line writer,
struct.cc:1991) is compared BYTE FOR BYTE -- through sha256 -- with what the unmodified reference left on the same
file (tests/golden/c1_dump_hashes.json, made by tests/golden/gen_c1_golden.py).

GPU part: the `lib/token_test` binary itself on that file: default family (test_spmm_thread_total(..., 4, 1),
token_test.cc:1656-1657), each of the seven families, the B200 kernels, --half, and the host check's verdict."""
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "c1_dump_hashes.json")))
FAMILIES = ["thread_total", "warp_total", "warp_bitmap", "thread_bit_map", "warp_segment", "block_total", "block_bitmap"]
TOKEN_TEST = os.path.join(ROOT, "generalsparse_b200", "lib", "token_test")


@pytest.fixture(scope="module")
def c1_mtx(tmp_path_factory):
    from generalsparse_b200 import synth
    rows, cols, r, c = synth.c1_standin()
    path = str(tmp_path_factory.mktemp("c1") / "c1.mtx")
    synth.write_mtx(path, rows, cols, r, c, comments=("%%MatrixMarket matrix coordinate real general", "% C1 stand-in: shape and nnz of Meszaros/IG5-18"))
    assert hashlib.sha256(open(path, "rb").read()).hexdigest() == GOLD["matrix"]["mtx_sha256"], "the stand-in generator drifted from the golden file"
    return path, rows, cols, r, c


def _compare_dump(dump_dir, fam):
    want = GOLD["chains"][fam]["arrays"]
    got = {f for f in os.listdir(dump_dir) if "_META_" in f}
    assert got == set(want), (sorted(got ^ set(want)))
    for key, w in want.items():
        h = hashlib.sha256(open(os.path.join(dump_dir, key), "rb").read()).hexdigest()
        assert h == w["sha256"], "%s differs from the reference's dump (%s)" % (key, fam)


@pytest.mark.parametrize("fam", FAMILIES)
def test_c1_loader_and_pipeline_dump_equal_the_reference(built_lib, c1_mtx, tmp_path, fam):
    from generalsparse_b200 import api
    path, rows, cols, r, c = c1_mtx
    spec = GOLD["chains"][fam]
    ms = api.MetaDataSet.from_mtx(path)  # reference mode: every value becomes 1 (struct.cc:188-195)
    assert int(ms["GLOBAL_META_origin_nnz_num_-1"][0]) == len(r)
    s = api.Session(ms)
    if "aborted" in spec:
        with pytest.raises(api.GsError):
            s.run_family(fam, spec["N"], spec["S"], spec["C"])
        return
    s.run_family(fam, spec["N"], spec["S"], spec["C"])
    for key, w in spec["arrays"].items():  # lengths and both dtypes first: a readable failure
        n, mem, comp = ms.info(key)
        assert (n, mem, comp) == (w["len"], w["mem"], w["compressed"]), key
    assert sorted(k for k in s.deps()) == sorted(spec["registered"])
    d = str(tmp_path / "dump")
    ms.dump(d)
    _compare_dump(d, fam)


def test_mtx_loader_rules(built_lib, tmp_path):
    """the loader's own rules (struct.cc:97-125,188-195): `%` comments skipped, 1-based indices, a header without the
    nnz field, values read but replaced by 1 in reference mode and kept with reference_ones=0, unsorted rows refused"""
    from generalsparse_b200 import api
    p = tmp_path / "m.mtx"
    p.write_text("%%MatrixMarket matrix coordinate real general\n% a comment\n%another\n3 4\n1 2 0.5\n1 4 -2\n3 1 7.25\n")
    ms = api.MetaDataSet.from_mtx(str(p))
    assert ms["GLOBAL_META_nz_row_indices_0"].tolist() == [0, 0, 2] and ms["GLOBAL_META_nz_col_indices_0"].tolist() == [1, 3, 0]
    assert ms["GLOBAL_META_nz_vals_0"].tolist() == [1.0, 1.0, 1.0]
    assert int(ms["GLOBAL_META_origin_row_num_-1"][0]) == 3 and int(ms["GLOBAL_META_origin_col_num_-1"][0]) == 4
    ms = api.MetaDataSet.from_mtx(str(p), reference_ones=False)
    assert ms["GLOBAL_META_nz_vals_0"].tolist() == [0.5, -2.0, 7.25]
    p.write_text("3 4 2\n2 1 1\n1 1 1\n")
    with pytest.raises(api.GsError):
        api.MetaDataSet.from_mtx(str(p))
    with pytest.raises(api.GsError):
        api.MetaDataSet.from_mtx(str(tmp_path / "missing.mtx"))


def _run(args, cwd):
    p = subprocess.run([TOKEN_TEST] + args, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:]
    return p.stdout


@pytest.mark.gpu
def test_token_test_binary_default_family_on_c1(built_lib, c1_mtx, tmp_path):
    """`token_test c1.mtx 8`, nothing else: the reference driver's command line, fp32, its default chain; the dump equals the
    reference's, the host check says `correct`, the two result lines have the reference's form"""
    path = c1_mtx[0]
    d = str(tmp_path / "dump")
    out = _run([path, "8", "--repeat", "200", "--dump", d], str(tmp_path))
    assert "\ncorrect\n" in out and "wrong number" not in out, out[-1500:]
    _compare_dump(d, "thread_total")
    lines = out.strip().splitlines()
    assert lines[-2].startswith(path + " ") and float(lines[-2].split()[-1]) > 0          # "<matrix> <gflops>"
    assert lines[-1].startswith(path + " : min time: ") and float(lines[-1].split()[-1]) > 0
    assert "get_row_order_by_length::{" in out  # the step records the reference driver prints


@pytest.mark.gpu
@pytest.mark.parametrize("fam", FAMILIES)
@pytest.mark.parametrize("half", [False, True], ids=["fp32", "half"])
def test_token_test_binary_every_family_on_c1(built_lib, c1_mtx, tmp_path, fam, half):
    path = c1_mtx[0]
    args = [path, "8", "--family", fam, "--kernel", "reference", "--sparse-coarsen", str(GOLD["chains"][fam]["S"]), "--repeat", "20"] + (["--half"] if half else [])
    if "aborted" in GOLD["chains"][fam]:
        p = subprocess.run([TOKEN_TEST] + args, cwd=str(tmp_path), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
        assert p.returncode != 0 and "padding rate" in p.stdout
        return
    out = _run(args, str(tmp_path))
    assert "\ncorrect\n" in out and "wrong number" not in out, out[-1500:]


@pytest.mark.gpu
@pytest.mark.parametrize("args", [["--kernel", "b200_row_smem", "--sparse-coarsen", "8"], ["--kernel", "b200_row_smem", "--sparse-coarsen", "8", "--half"],
                                  ["--kernel", "auto"], ["--kernel", "auto", "--half"], ["--family", "warp_segment", "--kernel", "b200_nnz", "--n32"]],
                         ids=["row_smem", "row_smem_half", "auto", "auto_half", "nnz_N32"])
def test_token_test_binary_b200_kernels_on_c1(built_lib, c1_mtx, tmp_path, args):
    path = c1_mtx[0]
    N = "8"
    if "--n32" in args:
        args = [a for a in args if a != "--n32"]
        N = "32"
    out = _run([path, N, "--repeat", "50"] + args, str(tmp_path))
    assert "\ncorrect\n" in out and "wrong number" not in out, out[-1500:]


@pytest.mark.gpu
@pytest.mark.parametrize("args", [[], ["--half"], ["--family", "warp_total"], ["--family", "thread_bit_map", "--kernel", "reference"],
                                  ["--family", "thread_bit_map", "--kernel", "b200_nnz", "--n32"], ["--family", "warp_segment", "--kernel", "reference"]],
                         ids=["default", "default_half", "warp_total", "thread_bit_map", "nnz_N32", "warp_segment"])
def test_token_test_binary_device_transform_on_c1(built_lib, c1_mtx, tmp_path, args):
    """--device-transform: the loader's COO goes to the device and the chain is built there; same kernel line, same check"""
    path = c1_mtx[0]
    N = "8"
    if "--n32" in args:
        args = [a for a in args if a != "--n32"]
        N = "32"
    host = _run([path, N, "--repeat", "20", "--sparse-coarsen", "4"] + args, str(tmp_path))
    dev = _run([path, N, "--repeat", "20", "--sparse-coarsen", "4", "--device-transform"] + args, str(tmp_path))
    assert "\ncorrect\n" in dev and "wrong number" not in dev and "device transform: " in dev, dev[-1500:]
    kernel = [l for l in host.splitlines() if l.startswith("kernel: ")]
    assert kernel and kernel == [l for l in dev.splitlines() if l.startswith("kernel: ")], (kernel, dev[-800:])


@pytest.mark.gpu
@pytest.mark.parametrize("half", [False, True], ids=["fp32", "half"])
def test_token_test_check_can_fail(built_lib, c1_mtx, tmp_path, half):
    """--corrupt-check flips one element of the host copy of C: the check must say `wrong number:1`, with HALF too
    (round 1's HALF check compared nothing)"""
    out = _run([c1_mtx[0], "8", "--repeat", "2", "--corrupt-check"] + (["--half"] if half else []), str(tmp_path))
    assert "wrong number:1\n" in out and "\ncorrect\n" not in out, out[-1500:]


@pytest.mark.gpu
def test_token_test_reads_half_from_global_config(built_lib, c1_mtx, tmp_path):
    """CWD convention of the reference: ./global_config.json with "HALF": true switches the run to fp16 storage"""
    cfg = {"HALF": True, "PADDING_RATE_UP_BOUND": 4, "FLOAT_RATE": 2, "VECTOR_WIDTH": 32, "DENSE_MATRIX_SIZE": 8, "OPERATOR_RUNTIME_CHECK": True}
    (tmp_path / "global_config.json").write_text(json.dumps(cfg))
    out = _run([c1_mtx[0], "8", "--repeat", "2", "--corrupt-check"], str(tmp_path))
    assert "wrong number:1\n" in out  # the HALF check ran (fp32 storage would flip another byte: still 1)
    out = _run([c1_mtx[0], "8", "--repeat", "2"], str(tmp_path))
    assert "\ncorrect\n" in out
