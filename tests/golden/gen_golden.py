#!/usr/bin/env python3
"""Generate golden transform_step fixtures by running the UNMODIFIED reference pipeline.

TEST INFRASTRUCTURE ONLY.  Runs in the dev container (needs /root/reference and
oracle/_ref/ref_harness, built by `make -C oracle ref`).  For every (matrix, family, N, S, C)
case below it
  1. writes the synthetic matrix as a 1-based, row-major-sorted COO text file
     (the only input format the reference loader accepts, struct.cc:97-125),
  2. runs the reference's own test_spmm_<family> operator chain (token_test.cc:1003-1582)
     through oracle/ref_harness.cc, which dumps EVERY metadata array left in the
     meta_data_set with the reference's own writer,
  3. packs the dump into tests/golden/<case>.npz: one u64 (or f64 for nz_vals) array per
     metadata key, plus "__dtypes__" (key -> compressed dtype chosen by the reference,
     code_source_data.cc:383-416), "__registered__" (keys the reference's code generator
     registered as kernel inputs = files in its data_source/<id>/ hand-off dir) and the
     input COO itself.

The .npz files are committed; /root/reference is not needed to run the tests.
"""
import json
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("GS_REFERENCE", "/root/reference")
HARNESS = os.path.join(REPO, "oracle", "_ref", "ref_harness")

sys.path.insert(0, REPO)
from generalsparse_b200.synth import synth_matrix  # noqa: E402

FAMILIES = ["thread_total", "warp_total", "warp_bitmap", "thread_bit_map", "warp_segment",
            "block_total", "block_bitmap"]

# (case name, generator kind, rows, cols, params, seed)
MATRICES = [
    ("tiny", "uniform", 64, 48, dict(density=0.085), 1),
    ("ragged", "ragged", 200, 150, dict(mean_len=9.0, empty_frac=0.15), 2),
    ("ig5like", "lognormal", 1500, 1300, dict(mean_len=37.4), 3),
    ("ffnlike", "uniform", 344, 512, dict(density=0.30), 4),
    ("powerlaw", "powerlaw", 3000, 3000, dict(avg_deg=20.0, alpha=2.1), 5),
]

# (family, N, sparse_coarsen, coarsen)
def cases_for(mname):
    out = []
    for fam in FAMILIES:
        out.append((fam, 8, 4 if fam == "thread_total" else 1, 1))
    if mname in ("tiny", "ffnlike"):
        out.append(("thread_total", 32, 8, 1))
        out.append(("thread_total", 8, 2, 2))
        out.append(("warp_segment", 32, 1, 1))
        out.append(("warp_bitmap", 32, 1, 2))
        out.append(("block_bitmap", 16, 1, 1))
        out.append(("thread_bit_map", 4, 1, 1))
    return out


# operator sequences none of the seven chains reaches, run through the harness' "ops=" mode
# (name, [(operator, args...), ...], N) -- arguments are the reference constructors' own, in declaration order
OPS_CASES = [
    ("mp_thread4", [("merge_path_thread_operator", 4)], 8),
    ("mp_thread33", [("merge_path_thread_operator", 33)], 8),
    ("mp_warp64", [("merge_path_warp_operator", 64)], 8),
    ("mp_tblock256", [("merge_path_tblock_operator", 256)], 8),
    ("sort_mp_thread7", [("sort_operator",), ("merge_path_thread_operator", 7)], 8),
    ("mp_tblock1", [("merge_path_tblock_operator", 1)], 8),
    # operators the factory already had, pinned here for the first time
    ("bal_tblock300", [("balanced_interval_row_direction_tblock_blocking_operator", 300)], 8),
    ("bal_warp64", [("balanced_interval_row_direction_warp_blocking_operator", 64, 0, 0)], 8),
    ("bal_thread16", [("balanced_interval_row_direction_thread_blocking_operator", 16, 0, 0)], 8),
    ("sort_bal_thread9", [("sort_operator",), ("balanced_interval_row_direction_thread_blocking_operator", 9, 0, 0)], 8),
    ("col_thread8_pad", [("fixed_interval_col_direction_thread_blocking_operator", 8, 0, 0, 1, 0)], 8),
    ("col_thread4_maxpad", [("fixed_interval_col_direction_thread_blocking_operator", 4, 0, 0, 0, 1)], 8),
    ("nnz_thread32_pad", [("fixed_interval_nnz_direction_thread_blocking_operator", 32, 0, 0, 1)], 8),
    ("nnz_thread5", [("fixed_interval_nnz_direction_thread_blocking_operator", 5, 0, 0, 0)], 8),
    ("empty_pad", [("empty_row_pad_operator",)], 8),
    ("tb16_w4_t1", [("fixed_interval_row_direction_tblock_blocking_operator", 16, 0),
                    ("fixed_interval_row_direction_warp_blocking_operator", 4, 0, 0, 0),
                    ("fixed_interval_row_direction_thread_blocking_operator", 1, 0, 0, 0, 0, 0, 0)], 8),
    ("tb16_w4rel_t1rel", [("fixed_interval_row_direction_tblock_blocking_operator", 16, 0),
                          ("fixed_interval_row_direction_warp_blocking_operator", 4, 1, 1, 0),
                          ("fixed_interval_row_direction_thread_blocking_operator", 1, 1, 1, 0, 0, 0, 0)], 8),
    ("tb32pad_w8pad", [("fixed_interval_row_direction_tblock_blocking_operator", 32, 1),
                       ("fixed_interval_row_direction_warp_blocking_operator", 8, 0, 0, 1)], 8),
    ("w2_t1_maxpad", [("fixed_interval_row_direction_warp_blocking_operator", 2, 0, 0, 0),
                      ("fixed_interval_row_direction_thread_blocking_operator", 1, 0, 0, 0, 1, 0, 0)], 8),
    ("t4_rowpad", [("fixed_interval_row_direction_thread_blocking_operator", 4, 0, 0, 1, 0, 0, 0)], 8),
    ("tb8_colthread4", [("fixed_interval_row_direction_tblock_blocking_operator", 8, 0),
                        ("fixed_interval_col_direction_thread_blocking_operator", 4, 0, 0, 1, 0)], 8),
    ("tb8_colthread4_rel", [("fixed_interval_row_direction_tblock_blocking_operator", 8, 0),
                            ("fixed_interval_col_direction_thread_blocking_operator", 4, 1, 1, 1, 0)], 8),
    # nnz-direction blocking above the thread level (fixed_interval_nnz_direction_{tblock,warp}_blocking_operator)
    ("nnz_tblock64", [("fixed_interval_nnz_direction_tblock_blocking_operator", 64, 0)], 8),
    ("nnz_tblock64_pad", [("fixed_interval_nnz_direction_tblock_blocking_operator", 64, 1)], 8),
    ("nnz_warp32", [("fixed_interval_nnz_direction_warp_blocking_operator", 32, 0, 0, 0)], 8),
    ("nnz_warp32_pad", [("fixed_interval_nnz_direction_warp_blocking_operator", 32, 0, 0, 1)], 8),
    ("nnz_tb128pad_w32", [("fixed_interval_nnz_direction_tblock_blocking_operator", 128, 1),
                          ("fixed_interval_nnz_direction_warp_blocking_operator", 32, 0, 0, 0)], 8),
    ("nnz_tb128pad_w32_t8", [("fixed_interval_nnz_direction_tblock_blocking_operator", 128, 1),
                             ("fixed_interval_nnz_direction_warp_blocking_operator", 32, 0, 0, 0),
                             ("fixed_interval_nnz_direction_thread_blocking_operator", 8, 0, 0, 0)], 8),
    ("nnz_tb128pad_w32rel", [("fixed_interval_nnz_direction_tblock_blocking_operator", 128, 1),
                             ("fixed_interval_nnz_direction_warp_blocking_operator", 32, 1, 1, 0)], 8),
    ("nnz_w32pad_t8", [("fixed_interval_nnz_direction_warp_blocking_operator", 32, 0, 0, 1),
                       ("fixed_interval_nnz_direction_thread_blocking_operator", 8, 0, 0, 0)], 8),
]

# round 2: the operators that were missing (column-direction blocking above the thread level, interleaved storage, the two sub-matrix
# divisions, balanced blocking inside a parent level, nnz-direction BMT indices relative to their parent)
OPS_CASES_R2 = [
    ("col_tblock16", [("fixed_interval_col_direction_tblock_blocking_operator", 16, 0, 0)], 8),
    ("col_tblock16_pad", [("fixed_interval_col_direction_tblock_blocking_operator", 16, 1, 0)], 8),
    ("col_tblock4_maxpad", [("fixed_interval_col_direction_tblock_blocking_operator", 4, 0, 1)], 8),
    ("col_warp8", [("fixed_interval_col_direction_warp_blocking_operator", 8, 0, 0, 0, 0)], 8),
    ("col_warp8_pad", [("fixed_interval_col_direction_warp_blocking_operator", 8, 0, 0, 1, 0)], 8),
    ("tb8_colwarp4_rel", [("fixed_interval_row_direction_tblock_blocking_operator", 8, 0),
                          ("fixed_interval_col_direction_warp_blocking_operator", 4, 1, 1, 0, 0)], 8),
    ("tb8_colwarp4_pad", [("fixed_interval_row_direction_tblock_blocking_operator", 8, 0),
                          ("fixed_interval_col_direction_warp_blocking_operator", 4, 1, 1, 1, 0)], 8),
    ("tb8_colwarp4_maxpad", [("fixed_interval_row_direction_tblock_blocking_operator", 8, 0),
                             ("fixed_interval_col_direction_warp_blocking_operator", 4, 0, 0, 0, 1)], 8),
    ("colwarp8_colthread", [("fixed_interval_col_direction_warp_blocking_operator", 8, 0, 0, 1, 0),
                            ("fixed_interval_col_direction_thread_blocking_operator", 4, 0, 0, 1, 0)], 8),
    ("col_thread8_pad_interlance", [("fixed_interval_col_direction_thread_blocking_operator", 8, 0, 0, 1, 0), ("interlance_storage_operator",)], 8),
    ("tb8_colthread4_pad_interlance", [("fixed_interval_row_direction_tblock_blocking_operator", 8, 0),
                                       ("fixed_interval_col_direction_thread_blocking_operator", 4, 0, 0, 1, 0), ("interlance_storage_operator",)], 8),
    ("w2_t1_maxpad_interlance", [("fixed_interval_row_direction_warp_blocking_operator", 2, 0, 0, 0),
                                 ("fixed_interval_row_direction_thread_blocking_operator", 1, 0, 0, 0, 1, 0, 0), ("interlance_storage_operator",)], 8),
    ("t1_maxpad_interlance", [("fixed_interval_row_direction_thread_blocking_operator", 1, 0, 0, 0, 1, 0, 0), ("interlance_storage_operator",)], 8),
    ("interlance_without_padding", [("fixed_interval_col_direction_thread_blocking_operator", 8, 0, 0, 0, 0), ("interlance_storage_operator",)], 8),
    ("fixed_div16", [("fixed_interval_row_matrix_div_operator", 16)], 8),
    ("fixed_div40", [("fixed_interval_row_matrix_div_operator", 40)], 8),
    ("fixed_div_too_many", [("fixed_interval_row_matrix_div_operator", 3)], 8),
    ("sort_rownz_div", [("sort_operator",), ("row_nz_matrix_div_operator", 2, 16, 2)], 8),
    ("sort_rownz_div_4_64_4", [("sort_operator",), ("row_nz_matrix_div_operator", 4, 64, 4)], 8),
    ("rownz_div_unsorted", [("row_nz_matrix_div_operator", 8, 64, 4)], 8),
    ("tb16_balwarp40", [("fixed_interval_row_direction_tblock_blocking_operator", 16, 0),
                        ("balanced_interval_row_direction_warp_blocking_operator", 40, 0, 0)], 8),
    ("tb16_balwarp40_rel", [("fixed_interval_row_direction_tblock_blocking_operator", 16, 0),
                            ("balanced_interval_row_direction_warp_blocking_operator", 40, 1, 1)], 8),
    ("tb16_balwarp40_balthread9_rel", [("fixed_interval_row_direction_tblock_blocking_operator", 16, 0),
                                       ("balanced_interval_row_direction_warp_blocking_operator", 40, 0, 0),
                                       ("balanced_interval_row_direction_thread_blocking_operator", 9, 1, 1)], 8),
    ("tb16_balthread12_rel", [("fixed_interval_row_direction_tblock_blocking_operator", 16, 0),
                              ("balanced_interval_row_direction_thread_blocking_operator", 12, 1, 1)], 8),
    ("baltb300_balthread30", [("balanced_interval_row_direction_tblock_blocking_operator", 300),
                              ("balanced_interval_row_direction_thread_blocking_operator", 30, 0, 1)], 8),
    ("nnz_tb128pad_w32_t8_rel", [("fixed_interval_nnz_direction_tblock_blocking_operator", 128, 1),
                                 ("fixed_interval_nnz_direction_warp_blocking_operator", 32, 0, 0, 0),
                                 ("fixed_interval_nnz_direction_thread_blocking_operator", 8, 1, 1, 0)], 8),
    ("nnz_tb64pad_t8_rel", [("fixed_interval_nnz_direction_tblock_blocking_operator", 64, 1),
                            ("fixed_interval_nnz_direction_thread_blocking_operator", 8, 1, 1, 0)], 8),
]
OPS_CASES = OPS_CASES + OPS_CASES_R2


def ops_cases_for(mname):
    if "--r2-only" in sys.argv:
        return OPS_CASES_R2 if mname in ("tiny", "ragged") else []
    return OPS_CASES if mname in ("tiny", "ragged") else OPS_CASES[:2] + OPS_CASES[6:10] if mname == "ig5like" else OPS_CASES[:2]


def write_mtx(path, rows, cols, r, c):
    with open(path, "w") as f:
        f.write("%%MatrixMarket matrix coordinate real general\n")
        f.write("%d %d %d\n" % (rows, cols, len(r)))
        for i in range(len(r)):
            f.write("%d %d 1\n" % (r[i] + 1, c[i] + 1))


def make_scratch():
    d = tempfile.mkdtemp(prefix="gsref_")
    cfg = json.load(open(os.path.join(REF, "global_config.json.bak")))
    cfg.update(ROOT_PATH_STR=d, spmv_header_file=os.path.join(d, "spmm_header_top.code"), HALF=False, DATA_SET="")
    json.dump(cfg, open(os.path.join(d, "global_config.json"), "w"))
    shutil.copytree(os.path.join(REF, "arm_code"), os.path.join(d, "arm_code"))
    os.makedirs(os.path.join(d, "data_source"))
    return d


def read_dump(dump_dir):
    out, dtypes = {}, {}
    for line in open(os.path.join(dump_dir, "INDEX")):
        key, n, mem_t, comp_t = line.split()
        txt = open(os.path.join(dump_dir, key)).read().split()
        assert len(txt) == int(n), (key, len(txt), n)
        if mem_t in ("FLOAT", "DOUBLE"):
            out[key] = np.array([float(x) for x in txt], dtype=np.float64)
        else:
            out[key] = np.array([int(x) for x in txt], dtype=np.uint64)
        dtypes[key] = comp_t
    return out, dtypes


def run_case(scratch, mtx, fam, n, s, c, tag=None):
    dump = os.path.join(scratch, "dump_%s_%d_%d_%d" % (tag or fam, n, s, c))
    shutil.rmtree(dump, ignore_errors=True)
    p = subprocess.run([HARNESS, fam, mtx, str(n), str(s), str(c), dump], cwd=scratch,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if p.returncode != 0 or not os.path.isdir(dump):
        return None, p.stdout[-2000:]
    arrs, dtypes = read_dump(dump)
    registered = []
    if os.path.exists(dump + ".handoff"):
        hand = open(dump + ".handoff").read().strip()
        if os.path.isdir(hand):
            registered = sorted(x for x in os.listdir(hand) if "_META_" in x)
    cfg = json.load(open(os.path.join(scratch, "global_config.json")))
    # operator / step records the reference prints (observable behaviour)
    steps = [ln for ln in p.stdout.splitlines() if "::{" in ln or "operator" in ln]
    return dict(arrs=arrs, dtypes=dtypes, registered=registered, vector_width=int(cfg["VECTOR_WIDTH"]),
                steps=steps), p.stdout[-500:]


def main():
    only = sys.argv[1:]
    scratch = make_scratch()
    summary = {}
    ops_only = "--ops-only" in only
    only = [x for x in only if not x.startswith("--")]
    if ops_only or only:
        summary = json.load(open(os.path.join(HERE, "SUMMARY.json")))
    for mname, kind, rows, cols, params, seed in MATRICES:
        if only and mname not in only:
            continue
        r, c = synth_matrix(kind, rows, cols, seed=seed, **params)
        mtx = os.path.join(scratch, mname + ".mtx")
        write_mtx(mtx, rows, cols, r, c)
        for oname, ops, n in ops_cases_for(mname):
            spec = "ops=" + ";".join(",".join(str(int(x)) if not isinstance(x, str) else x for x in op) for op in ops)
            res, log = run_case(scratch, mtx, spec, n, 0, 0, tag=oname)
            name = "%s__ops_%s" % (mname, oname)
            if res is None:
                # the reference asserted: keep the input and the assertion text, ours has to refuse the sequence too
                why = [ln for ln in log.splitlines() if "Assertion" in ln or "padding rate" in ln]
                print("ABORTED", name, why[-1][-160:] if why else log[-200:])
                summary[name] = "reference aborted"
                np.savez_compressed(os.path.join(HERE, name + ".npz"), __coo_row__=r.astype(np.int64), __coo_col__=c.astype(np.int64),
                                    __shape__=np.array([rows, cols], dtype=np.int64),
                                    __meta__=np.array(json.dumps(dict(ops=[list(op) for op in ops], N=n, aborted=[w[-300:] for w in why]))))
                continue
            payload = {k: v for k, v in res["arrs"].items()}
            payload["__coo_row__"] = r.astype(np.int64)
            payload["__coo_col__"] = c.astype(np.int64)
            payload["__shape__"] = np.array([rows, cols], dtype=np.int64)
            payload["__meta__"] = np.array(json.dumps(dict(
                dtypes=res["dtypes"], ops=[list(op) for op in ops], N=n, steps=res["steps"],
                matrix=dict(kind=kind, rows=rows, cols=cols, params=params, seed=seed))))
            np.savez_compressed(os.path.join(HERE, name + ".npz"), **payload)
            summary[name] = "%d arrays" % len(res["arrs"])
            print(name, summary[name])
        if ops_only:
            continue
        for fam, n, s, co in cases_for(mname):
            res, log = run_case(scratch, mtx, fam, n, s, co)
            name = "%s__%s__N%d_S%d_C%d" % (mname, fam, n, s, co)
            if res is None:
                print("FAILED", name, "\n", log)
                summary[name] = "reference aborted"
                continue
            payload = {k: v for k, v in res["arrs"].items()}
            payload["__coo_row__"] = r.astype(np.int64)
            payload["__coo_col__"] = c.astype(np.int64)
            payload["__shape__"] = np.array([rows, cols], dtype=np.int64)
            payload["__meta__"] = np.array(json.dumps(dict(
                dtypes=res["dtypes"], registered=res["registered"], vector_width=res["vector_width"],
                family=fam, N=n, S=s, C=co, steps=res["steps"], matrix=dict(kind=kind, rows=rows, cols=cols,
                                                                              params=params, seed=seed))))
            np.savez_compressed(os.path.join(HERE, name + ".npz"), **payload)
            summary[name] = "%d arrays" % len(res["arrs"])
            print(name, summary[name], "registered:", len(res["registered"]))
    json.dump(summary, open(os.path.join(HERE, "SUMMARY.json"), "w"), indent=1, sort_keys=True)
    shutil.rmtree(scratch, ignore_errors=True)


if __name__ == "__main__":
    main()
