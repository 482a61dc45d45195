"""The C-ABI library on a machine without a GPU: it loads, exports every symbol include/gsparse_b200.h declares,
the host half (metadata store, config, operator graph, partition) behaves like the reference's, and the device half
fails loudly -- never silently -- when there is no CUDA device.  No compute calls here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from generalsparse_b200 import _lib, api, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    """every function name declared in include/gsparse_b200.h"""
    src = open(os.path.join(ROOT, "include", "gsparse_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gsh?_[a-z0-9_]+)\s*\(", src)))


def test_header_and_ctypes_agree(built_lib):
    declared = _declared_symbols()
    assert len(declared) >= 45
    assert sorted(_lib.SYMBOLS) == declared, "ctypes SYMBOLS and the header drifted apart"


def test_library_exports_every_declared_symbol(built_lib):
    raw = C.CDLL(_lib.LIB_PATH)
    missing = [s for s in _declared_symbols() if not hasattr(raw, s)]
    assert not missing, missing


def test_header_compiles_as_plain_c(tmp_path):
    """the boundary is a C ABI: the header must be valid C (no C++ types in the signatures)"""
    import subprocess
    src = tmp_path / "t.c"
    src.write_text('#include "gsparse_b200.h"\nint main(void) { gs_plan_desc d; (void)d; return (int)GS_OK; }\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(src), "-o", str(tmp_path / "t.o")], check=True)


def test_version_and_family_names(built_lib):
    L = built_lib
    assert L.gs_version().decode()
    names = [L.gs_family_name(i).decode() for i in range(len(_lib.FAMILIES))]
    assert names == _lib.FAMILIES


def _small():
    r, c = synth.synth_matrix("ragged", 200, 150, seed=2, mean_len=6.0, empty_frac=0.25)
    return 200, 150, r, c


def test_metadata_store_roundtrip(built_lib):
    rows, cols, r, c = _small()
    ms = api.MetaDataSet.from_coo(rows, cols, r, c)
    keys = ms.keys()
    for k in ("GLOBAL_META_nz_row_indices_0", "GLOBAL_META_nz_col_indices_0", "GLOBAL_META_nz_vals_0", "GLOBAL_META_origin_nnz_num_-1"):
        assert k in keys
    assert int(ms["GLOBAL_META_origin_nnz_num_-1"][0]) == len(r)
    assert np.array_equal(ms["GLOBAL_META_nz_col_indices_0"].astype(np.int64), c[np.lexsort((c, r))])
    assert np.all(ms["GLOBAL_META_nz_vals_0"] == 1.0)  # the reference forces every value to 1 (struct.cc:188-195)
    ms.check()
    with pytest.raises(_lib.GsError) as e:
        ms["GLOBAL_META_no_such_array_0"]
    assert e.value.status == 4  # GS_ERR_NOT_FOUND


def test_dump_is_one_decimal_line_per_element(built_lib, tmp_path):
    """print_arr_to_file_with_data_type (struct.cc:1991): what the generated program used to read back"""
    rows, cols, r, c = _small()
    ms = api.MetaDataSet.from_coo(rows, cols, r, c)
    api.Session(ms).run_family("thread_total", 8, 4, 1)
    ms.dump(str(tmp_path))
    f = tmp_path / "THREAD_META_first_nz_indices_0"
    assert f.exists()
    lines = f.read_text().split("\n")
    want = ms["THREAD_META_first_nz_indices_0"]
    assert [int(x) for x in lines if x] == [int(x) for x in want]
    vals = (tmp_path / "GLOBAL_META_nz_vals_0").read_text().split()
    assert set(vals) <= {"1", "0"}  # ones, and the padding zeros


def test_config_set_get(built_lib):
    api.config_set("VECTOR_WIDTH", 16)
    assert api.config_get("VECTOR_WIDTH") == "16"
    api.config_set("VECTOR_WIDTH", 32)
    with pytest.raises(_lib.GsError):
        api.config_get("NO_SUCH_KEY")


def test_operator_preconditions_are_statuses_not_aborts(built_lib):
    rows, cols, r, c = _small()
    s = api.Session(api.MetaDataSet.from_coo(rows, cols, r, c))
    with pytest.raises(_lib.GsError) as e:
        s.add_and_run("no_such_operator")
    assert e.value.status in (1, 3, 4)
    with pytest.raises(_lib.GsError):
        s.add_and_run("fixed_interval_row_direction_thread_blocking_operator", 0)  # block size 0: the reference asserts
    with pytest.raises(_lib.GsError):
        s.plan_desc()  # nothing compiled yet


def test_padding_rate_abort_is_a_status(built_lib):
    """modify_*_by_*_pad aborts when padded/original >= PADDING_RATE_UP_BOUND; here: GS_ERR_PADDING_RATE"""
    rows, cols = 64, 64
    r = np.arange(rows, dtype=np.int64)
    c = np.arange(rows, dtype=np.int64)  # one non-zero per row
    s = api.Session(api.MetaDataSet.from_coo(rows, cols, r, c))
    with pytest.raises(_lib.GsError) as e:
        s.run_family("thread_total", 8, 8, 1)  # every row padded 1 -> 8
    assert e.value.status == 5


def test_partition_rows_balanced_matches_reference_rule(built_lib):
    rng = np.random.default_rng(3)
    row_nnz = rng.integers(0, 50, size=1000).astype(np.uint64)
    for G in (1, 2, 3, 4, 8):
        cuts = api.partition_rows_balanced(row_nnz, G)
        assert len(cuts) == G + 1 and cuts[0] == 0 and cuts[-1] == 1000 and np.all(np.diff(cuts.astype(np.int64)) >= 0)
        # greedy rule of data_transform_common.cc:960-988: close a part when its nnz reaches ceil(total/G)
        per = -(-int(row_nnz.sum()) // G)
        want, acc = [0], 0
        for i, n in enumerate(row_nnz):
            acc += int(n)
            if acc >= per and len(want) < G:
                want.append(i + 1)
                acc = 0
        while len(want) < G:
            want.append(1000)
        want.append(1000)
        assert list(cuts) == want


def test_device_half_fails_loudly_without_a_gpu(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(_lib.GsError) as e:
        api.Engine(0)
    assert e.value.status == 2 and "no CPU fallback" in str(e.value)
