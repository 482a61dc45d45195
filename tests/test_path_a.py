"""The process boundary of the reference, both ways round.

  * façade: our operators -> code_generator::generate_final_program(repeat) writes ROOT_PATH_STR/data_source/<id>/ (one text file per
    array the kernel reads, the reference's hand-off: metadata_set.cc:517-571) -> execute_binary(path, matrix, &ms, &gflops)
    (executor.hpp:15) runs the B200 kernels on those files;
  * INTEGRATION.md path A compiled for real: oracle/_ref/ref_path_a = the UNMODIFIED reference (its driver, operators, transform steps
    and code generator, compiled from /root/reference) linked with integration/executor_b200.cc in place of its executor.cc.  The
    reference writes the hand-off itself; the same gs_execute_binary reads it.  Present only where the dev container built it."""
import json
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATH_A = os.path.join(ROOT, "oracle", "_ref", "ref_path_a")
GOLD = os.path.join(ROOT, "tests", "golden")
# the keys the reference reads from ./global_config.json (values of its shipped global_config.json.bak, HALF off, paths set per test)
REF_CONFIG = {"BRANCH_COMPRESS_MAX_SIZE": 5, "DATA_SET": "", "DATA_TYPE_COMPRESS": True, "DEFAULT_DEVICE_ID": 0, "DEFAULT_THREAD_BLOCK_NUM": 120,
              "DEFAULT_THREAD_NUM_IN_BLOCK": 512, "DENSE_MATRIX_SIZE": 8, "FIXED_THREAD_COMBINE_SIZE": 32, "FIXED_WARP_COMBINE_SIZE": 4, "FLOAT_RATE": 2,
              "FORMAT_OF_MTX": "COO", "GFLOPS_UP_BOUND": 1000000, "GRAPH_MAX_ITER": 1000, "Graph_Algorithm": "PR", "HALF": False,
              "HALF_MAX_ROW_REDUCE_THREAD": 999998, "KERNAL_REPEAT_TIME": 1200, "MAX_DIV_TIMES_OF_DIV": 12, "MAX_ROW_REDUCE_THREAD": 999999,
              "MAX_TBLOCK_NUM": 65535, "MAX_THREAD_NUM_IN_BLOCK": 1024, "MODEL_DRIVEN_INDEX_COMPRESS": True, "MODEL_DRIVEN_INDEX_COMPRESS_WITH_ERROR": True,
              "OPERATOR_RUNTIME_CHECK": True, "PADDING_RATE_UP_BOUND": 4, "PERFORMANCE_FLAG": "throughput", "PRECISE_OF_FLOAT": "float",
              "SHARED_MEM_BROADCASR_OPTIMIZATION": True, "SHARED_MEM_TOTAL_SIZE": 48000, "SORT_THREAD_NUM": 1, "SYNC_IN_KERNEL_ITER": False,
              "Semiring": "MultipleAndAdd", "VECTOR_WIDTH": 8, "shared_sparse_flag": True}


def _scratch(tmp_path, half=False):
    d = str(tmp_path)
    cfg = dict(REF_CONFIG, ROOT_PATH_STR=d, spmv_header_file=d + "/spmm_header_top.code", HALF=half)
    json.dump(cfg, open(os.path.join(d, "global_config.json"), "w"))
    os.makedirs(os.path.join(d, "data_source"), exist_ok=True)
    os.makedirs(os.path.join(d, "arm_code"), exist_ok=True)
    # generate_final_program copies these two into the hand-off directory and asserts that the first exists (code_generator.cc:686-693);
    # nobody compiles them any more
    open(os.path.join(d, "arm_code", "kernel_lib.hpp"), "w").write("// placeholder: the generated program is not compiled on this path\n")
    open(os.path.join(d, "arm_code", "make_kernel.sh"), "w").write("#!/bin/sh\nexit 0\n")
    return d


def _matrix(tmp_path, kind="ig5like"):
    from generalsparse_b200 import synth
    rows, cols = 1500, 1300
    r, c = synth.synth_matrix("lognormal", rows, cols, seed=3, mean_len=37.4)
    path = os.path.join(str(tmp_path), "m.mtx")
    synth.write_mtx(path, rows, cols, r, c)
    return path, rows, cols, r, c


@pytest.mark.parametrize("fam", ["thread_total", "warp_segment", "block_bitmap"])
def test_generate_final_program_writes_the_reference_handoff(built_lib, tmp_path, fam):
    """the hand-off directory holds exactly the arrays the reference's generator registers for the chain (golden `registered` lists),
    byte-equal to the metadata dump, plus the repeat count"""
    from generalsparse_b200 import api
    path, rows, cols, r, c = _matrix(tmp_path)
    api.config_set("ROOT_PATH_STR", str(tmp_path))
    ms = api.MetaDataSet.from_mtx(path)
    s = api.Session(ms).run_family(fam, 8, 4 if fam == "thread_total" else 1, 1)
    ident, d = s.generate_final_program(123)
    assert d == os.path.join(str(tmp_path), "data_source", str(ident)) and os.path.isdir(d)
    z = np.load(os.path.join(GOLD, "ig5like__%s__N8_S%d_C1.npz" % (fam, 4 if fam == "thread_total" else 1)), allow_pickle=False)
    registered = json.loads(str(z["__meta__"]))["registered"]
    files = sorted(f for f in os.listdir(d) if "_META_" in f)
    assert files == sorted(registered)
    dump = os.path.join(str(tmp_path), "dump")
    ms.dump(dump)
    for f in files:
        assert open(os.path.join(d, f), "rb").read() == open(os.path.join(dump, f), "rb").read(), f
    assert open(os.path.join(d, "gs_repeat")).read().strip() == "123"
    ident2, d2 = s.generate_final_program(5)  # another call, another directory
    assert d2 != d and os.path.isdir(d2)


@pytest.mark.gpu
@pytest.mark.parametrize("fam,kernel", [("thread_total", "auto"), ("thread_total", "chain"), ("warp_total", "chain"), ("warp_bitmap", "chain"), ("thread_bit_map", "chain"),
                                        ("warp_segment", "chain"), ("block_total", "chain"), ("block_bitmap", "chain")])
def test_facade_generate_then_execute_binary(built_lib, tmp_path, capfd, fam, kernel):
    from generalsparse_b200 import api
    path, rows, cols, r, c = _matrix(tmp_path)
    d0 = _scratch(tmp_path)
    api.config_set("ROOT_PATH_STR", d0)
    cwd = os.getcwd()
    os.chdir(d0)  # ./global_config.json is read from the CWD, as in the reference
    try:
        s = api.Session(api.MetaDataSet.from_mtx(path)).run_family(fam, 8, 4 if fam == "thread_total" else 1, 1)
        json.dump(dict(json.load(open("global_config.json")), VECTOR_WIDTH=int(float(api.config_get("VECTOR_WIDTH"))), DENSE_MATRIX_SIZE=8), open("global_config.json", "w"))
        ident, d = s.generate_final_program(50)
        ms_total, gflops = api.execute_binary(d, path, kernel=kernel)
    finally:
        os.chdir(cwd)
    out = capfd.readouterr().out
    assert "\ncorrect\n" in "\n" + out and ms_total > 0 and gflops > 0
    assert abs(gflops - 2.0 * len(r) * 8 * 50 / (ms_total * 1e-3) / 1e9) / gflops < 1e-3  # repeat came from the directory
    assert [float(x) for x in open(os.path.join(d, "perf_result")).read().split()] == pytest.approx([ms_total, gflops], rel=1e-4)


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(PATH_A), reason="oracle/_ref/ref_path_a is built only where /root/reference exists")
@pytest.mark.parametrize("args,kernel", [([], "auto"), ([], "chain"), (["warp_total", "1", "1"], "chain"), (["thread_bit_map", "1", "1"], "chain"),
                                         (["warp_segment", "1", "1"], "auto"), (["block_total", "1", "1"], "chain"), (["warp_bitmap", "1", "1"], "chain"),
                                         (["block_bitmap", "1", "1"], "chain")], ids=lambda v: "-".join(v) if isinstance(v, list) else v)
def test_unmodified_reference_driver_with_our_executor(built_lib, tmp_path, args, kernel):
    """INTEGRATION.md path A for real: the reference's own token_test main / test_spmm_* chains build the metadata and write
    data_source/<id>/; its execute_binary call lands in integration/executor_b200.cc -> gs_execute_binary -> B200"""
    path, rows, cols, r, c = _matrix(tmp_path)
    d0 = _scratch(tmp_path)
    env = dict(os.environ, GS_PATH_A_KERNEL=kernel, GS_PATH_A_REPEAT="20")
    p = subprocess.run([PATH_A, path, "8"] + args, cwd=d0, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-3000:]
    assert "\ncorrect\n" in p.stdout and "wrong number" not in p.stdout, p.stdout[-3000:]
    line = [ln for ln in p.stdout.splitlines() if ln.startswith(path + " ") and "min time" not in ln][-1]  # "<matrix> <gflops>" (token_test.cc:1086)
    assert float(line.split()[-1]) > 0
    if kernel == "auto" and not args:
        assert "kernel: k_b200_" in p.stdout, p.stdout[-1500:]
    hand = [x for x in os.listdir(os.path.join(d0, "data_source"))]
    assert len(hand) == 1 and os.path.exists(os.path.join(d0, "data_source", hand[0], "kernel_file.cpp"))  # the reference wrote its program text; nobody compiled it
