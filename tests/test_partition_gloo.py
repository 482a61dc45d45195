"""Multi-process path on CPU: two ranks over gloo shard one matrix by the reference's nnz-balanced rule, run the host
operator chain on their own row block, and agree -- without any data-path collective -- on a partition that is disjoint,
covering and balanced; the per-rank oracle results stitched together equal the single-process result bit for bit.
(The device launch itself is covered by the -m gpu tests; bench.py --gpus N uses exactly this sharding.)"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import sys
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from generalsparse_b200 import api, shard, synth
    from oracle.oracle_check import csr_from_coo, oracle_spmm_f32

    rows, cols, N = 3000, 3000, 8
    r, c = synth.synth_matrix("powerlaw", rows, cols, seed=7, avg_deg=8.0)  # skewed rows: the case balancing is for
    vals = synth.synth_values(len(r), seed=7)
    B = synth.synth_dense(cols, N, seed=7)
    lo, hi, rl, cl, vl = shard.shard_coo(rows, r, c, vals, rank, world)

    # every rank derived the same cuts on its own; gather them only to check that
    mine = torch.tensor([lo, hi, len(rl)], dtype=torch.int64)
    allv = [torch.zeros(3, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(allv, mine)
    los, his, nnzs = [int(t[0]) for t in allv], [int(t[1]) for t in allv], [int(t[2]) for t in allv]
    assert los[0] == 0 and his[-1] == rows and all(his[i] == los[i + 1] for i in range(world - 1))
    assert sum(nnzs) == len(r)
    assert max(nnzs) <= 1.2 * len(r) / world + 0.05 * len(r)  # nnz-balanced (one hub row may sit on a cut)

    # the reference's operator chain on the local block only (host pipeline, no GPU)
    ms = api.MetaDataSet.from_coo(hi - lo, cols, rl, cl, vl)
    sess = api.Session(ms).run_family("warp_segment", N, 1, 1)
    desc = sess.plan_desc()
    assert desc.rows == hi - lo and desc.origin_nnz == len(rl)

    # local rows of C from the oracle; stitched together they must equal the unsharded result exactly
    order = np.lexsort((cl, rl))
    C_local = oracle_spmm_f32(hi - lo, N, csr_from_coo(hi - lo, rl), cl[order], vl[order], B)
    np.save(os.path.join(out_dir, "C_%d.npy" % rank), C_local)
    dist.barrier()
    if rank == 0:
        order = np.lexsort((c, r))
        C_all = oracle_spmm_f32(rows, N, csr_from_coo(rows, r), c[order], vals[order], B)
        parts = np.concatenate([np.load(os.path.join(out_dir, "C_%d.npy" % k)) for k in range(world)])
        assert np.array_equal(parts, C_all)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_ranks_gloo(built_lib, tmp_path):
    import __graft_entry__ as g
    if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle_spmm.so")):
        g.build_oracle()
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
