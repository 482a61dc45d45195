"""Order rules of the implementing operators (is_valid_according_to_operator, operator/*_reduce_operator.cc,
operator/*_bit_map_operator.cc): add_and_run must refuse what the reference refuses -- a total reduction after an
nnz-direction blocking would store one nnz chunk's sum as a whole row."""
import numpy as np
import pytest

from generalsparse_b200 import api, synth


def _session():
    r, c = synth.synth_matrix("uniform", 64, 48, seed=1, density=0.2)
    return api.Session(api.MetaDataSet.from_coo(64, 48, r, c))


REFUSED = [
    # (operators run first, the operator that must be refused)
    ([("fixed_interval_nnz_direction_thread_blocking_operator", (32, 0, 0, 1))], ("thread_total_reduce_operator", (0, 1, 1))),
    ([("fixed_interval_nnz_direction_warp_blocking_operator", (64, 0, 0, 1))], ("warp_total_reduce_operator", (1,))),
    ([("fixed_interval_nnz_direction_tblock_blocking_operator", (128, 1))], ("tblock_total_reduce_operator", (1,))),
    # thread_bit_map needs an nnz-direction operator AND a thread-level one
    ([("fixed_interval_row_direction_thread_blocking_operator", (1, 0, 0, 0, 0, 0, 1))], ("thread_bit_map_operator", (3, 8, 1, 1))),
    # nothing may have been implemented before a total reduction
    ([("fixed_interval_row_direction_thread_blocking_operator", (1, 0, 0, 0, 0, 0, 1)), ("thread_total_reduce_operator", (0, 1, 1))],
     ("thread_total_reduce_operator", (0, 1, 1))),
    ([("fixed_interval_row_direction_warp_blocking_operator", (1, 0, 0, 0)), ("warp_total_reduce_operator", (1,))], ("warp_total_reduce_operator", (1,))),
    ([("fixed_interval_row_direction_tblock_blocking_operator", (1, 0)), ("tblock_total_reduce_operator", (1,))], ("tblock_total_reduce_operator", (1,))),
    # a total reduction needs its own level
    ([("fixed_interval_row_direction_warp_blocking_operator", (1, 0, 0, 0))], ("thread_total_reduce_operator", (0, 1, 1))),
    ([("fixed_interval_row_direction_thread_blocking_operator", (1, 0, 0, 0, 0, 0, 1))], ("warp_total_reduce_operator", (1,))),
    # warp_segment needs the nnz direction, warp_bit_map refuses it
    ([("fixed_interval_nnz_direction_thread_blocking_operator", (32, 0, 0, 1)), ("thread_bit_map_operator", (2, 8, 1, 1))], ("warp_bit_map_operator", (1, 0, 0))),
    ([("fixed_interval_row_direction_thread_blocking_operator", (1, 0, 0, 0, 0, 0, 1)), ("thread_total_reduce_operator", (1, 1, 1))],
     ("warp_segment_reduce_operator", (1, 0, 0))),
]


@pytest.mark.parametrize("before,refused", REFUSED, ids=["%s<-%s" % (r[0].replace("_operator", ""), "+".join(b[0].split("_direction_")[-1].replace("_operator", "") for b in bf)) for bf, r in REFUSED])
def test_add_and_run_refuses_what_the_reference_refuses(built_lib, before, refused):
    api.config_set("VECTOR_WIDTH", 8)
    s = _session()
    for name, args in before:
        s.add_and_run(name, *args)
    with pytest.raises(api.GsError) as ei:
        s.add_and_run(refused[0], *refused[1])
    assert ei.value.status == 1 and "not valid after" in str(ei.value)


def test_the_driver_chains_still_pass_the_rules(built_lib):
    r, c = synth.synth_matrix("uniform", 344, 512, seed=4, density=0.3)
    for fam in ("thread_total", "warp_total", "warp_bitmap", "thread_bit_map", "warp_segment", "block_total", "block_bitmap"):
        s = api.Session(api.MetaDataSet.from_coo(344, 512, r, c)).run_family(fam, 8, 4 if fam == "thread_total" else 1, 1)
        assert s.plan_desc().n_deps >= 4
