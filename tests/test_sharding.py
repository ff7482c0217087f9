"""Placement policies vs the reference's arithmetic (SS/worker.py:19-36, SG/worker.py:15-37)."""
import pytest
from hypothesis import given, settings, strategies as st

from ddl_b200.models.spec import REFERENCE_SPEC, make_spec
from ddl_b200.parallel import sharding as S

SIZES = [v.numel for v in REFERENCE_SPEC.vars]


def test_reference_param_count():
    assert len(SIZES) == 14 and sum(SIZES) == 2_656_010
    assert REFERENCE_SPEC.spatial() == [28, 14, 7, 4, 2]


def test_greedy_interleave_order_matches_reference():
    assert S.reference_greedy_order(SIZES) == [13, 8, 1, 6, 3, 10, 5, 4, 7, 2, 11, 12, 0, 9]


@pytest.mark.parametrize("P,block,greedy", [
    (2, [1076224, 1579786], [2392298, 263712]),
    (4, [52032, 204992, 1868032, 530954], [1048618, 1343552, 205184, 58656]),
    (8, [800, 32, 51200, 64, 204800, 128, 819200, 1579786],
        [10, 1048576, 32, 819200, 64, 524288, 128, 263712]),
])
def test_shard_size_table(P, block, greedy):
    assert S.reference_shard_sizes("block", SIZES, P) == block
    assert S.reference_shard_sizes("greedy_interleave", SIZES, P) == greedy


def test_block_owner_matches_reference_formula():
    V, P = 14, 4
    avg = V // P
    local_last = avg + V % P
    for num in range(V):
        ind = num // avg
        if num >= V - local_last:
            ind = P - 1
        assert S.block_owner(num, V, P) == (ind, num - ind * avg)


def test_p8_var_lists():
    plan = S.make_plan("block", SIZES, 8)
    assert [plan.shard_vars(r) for r in range(8)] == [[0], [1], [2], [3], [4], [5], [6], list(range(7, 14))]
    plan = S.make_plan("greedy_interleave", SIZES, 8)
    assert [plan.shard_vars(r) for r in range(8)] == [[13], [8], [1], [6], [3], [10], [5], [4, 7, 2, 11, 12, 0, 9]]


@settings(max_examples=60, deadline=None)
@given(st.lists(st.integers(1, 5000), min_size=1, max_size=24), st.integers(1, 9),
       st.sampled_from(["block", "round_robin", "greedy", "greedy_interleave", "flat"]))
def test_every_var_owned_once_and_layout_sound(sizes, P, policy):
    plan = S.make_plan(policy, sizes, P)
    assert sorted(plan.order) == list(range(len(sizes)))
    # arena ranges: aligned, disjoint, inside the arena
    spans = sorted((plan.var_offset[v], plan.var_offset[v] + sizes[v]) for v in range(len(sizes)))
    for (a, b), (c, d) in zip(spans, spans[1:]):
        assert b <= c
    assert all(o % S.ALIGN_ELEMS == 0 for o in plan.var_offset)
    assert spans[-1][1] <= plan.total
    # shard ranges tile the arena
    assert plan.shard_range[0][0] == 0 and plan.shard_range[-1][1] == plan.total
    for (a, b), (c, d) in zip(plan.shard_range, plan.shard_range[1:]):
        assert b == c
    if policy != "flat":
        for v in range(len(sizes)):
            r = plan.owner[v]
            assert 0 <= r < P
            b, e = plan.shard_range[r]
            assert b <= plan.var_offset[v] and plan.var_offset[v] + sizes[v] <= e
        # tags are the local index inside the shard
        for r in range(P):
            assert [plan.tag[v] for v in plan.shard_vars(r)] == list(range(len(plan.shard_vars(r))))
    else:
        assert len({e - b for b, e in plan.shard_range}) == 1


@settings(max_examples=40, deadline=None)
@given(st.lists(st.integers(1, 10 ** 6), min_size=2, max_size=20), st.integers(2, 8))
def test_lpt_no_worse_than_interleave(sizes, P):
    def mx(policy):
        return max(S.reference_shard_sizes(policy, sizes, P))
    assert mx("greedy") <= mx("greedy_interleave")


def test_lpt_balance_on_reference_model():
    for P in (2, 4, 8):
        lpt = S.reference_shard_sizes("greedy", SIZES, P)
        ref = S.reference_shard_sizes("greedy_interleave", SIZES, P)
        assert max(lpt) <= max(ref)
    assert S.make_plan("flat", SIZES, 8).imbalance() == pytest.approx(1.0)


def test_widened_model_is_about_100m():
    spec = make_spec("cnn", 6)
    assert 80e6 < spec.num_params < 120e6
