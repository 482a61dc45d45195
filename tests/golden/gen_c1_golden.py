#!/usr/bin/env python3
"""Golden hashes of the reference's own pipeline on the C1 stand-in (BASELINE.json configs[0]).

TEST INFRASTRUCTURE ONLY; runs in the dev container (needs /root/reference and oracle/_ref/ref_harness).
The stand-in (generalsparse_b200.synth.c1_standin: IG5-18's shape and non-zero count) is written as the
1-based row-sorted .mtx text the reference loader reads (struct.cc:49-261, metadata_set.cc:612-707), the
UNMODIFIED reference runs each of its seven driver chains on that FILE at N=8 with the driver's own
coarsen factors, and every metadata array it leaves is dumped with its own writer.  The dumps are ~25 MB of
text per chain, so what is committed is tests/golden/c1_dump_hashes.json: per chain, per array, the sha256 of
the dump file, its length and both dtypes (or "aborted" with the assertion when the reference refuses the input).
tests/test_c1_token_test.py compares `token_test --dump` / gsh_metadata_dump byte for byte through these hashes.
"""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REPO)
sys.path.insert(0, HERE)
import gen_golden as gg  # noqa: E402
from generalsparse_b200 import synth  # noqa: E402


def main():
    scratch = gg.make_scratch()
    rows, cols, r, c = synth.c1_standin()
    mtx = os.path.join(scratch, "c1.mtx")
    synth.write_mtx(mtx, rows, cols, r, c, comments=("%%MatrixMarket matrix coordinate real general", "% C1 stand-in: shape and nnz of Meszaros/IG5-18"))
    out = {"matrix": dict(rows=rows, cols=cols, nnz=int(len(r)), mtx_sha256=hashlib.sha256(open(mtx, "rb").read()).hexdigest()), "chains": {}}
    import subprocess
    for fam in gg.FAMILIES:
        S = 4 if fam == "thread_total" else 1
        dump = os.path.join(scratch, "dump_" + fam)
        p = subprocess.run([gg.HARNESS, fam, mtx, "8", str(S), "1", dump], cwd=scratch, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if p.returncode != 0 or not os.path.isdir(dump):
            why = [ln for ln in p.stdout.splitlines() if "Assertion" in ln or "padding rate" in ln]
            out["chains"][fam] = dict(aborted=[w[-300:] for w in why] or [p.stdout[-300:]], S=S, C=1, N=8)
            print(fam, "ABORTED", why[-1][-120:] if why else "")
            continue
        arrays = {}
        for line in open(os.path.join(dump, "INDEX")):
            key, n, mem_t, comp_t = line.split()
            arrays[key] = dict(sha256=hashlib.sha256(open(os.path.join(dump, key), "rb").read()).hexdigest(), len=int(n), mem=mem_t, compressed=comp_t)
        registered = []
        if os.path.exists(dump + ".handoff"):
            hand = open(dump + ".handoff").read().strip()
            if os.path.isdir(hand):
                registered = sorted(x for x in os.listdir(hand) if "_META_" in x)
        out["chains"][fam] = dict(arrays=arrays, registered=registered, S=S, C=1, N=8)
        print(fam, len(arrays), "arrays")
        import shutil
        shutil.rmtree(dump, ignore_errors=True)
    json.dump(out, open(os.path.join(HERE, "c1_dump_hashes.json"), "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
