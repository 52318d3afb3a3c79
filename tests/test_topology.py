import pytest

from simple_distributed_machine_learning_b200.parallel.topology import Topology


def test_coords_roundtrip_and_neighbours():
    t = Topology(8, pp=2, dp=2, tp=2)
    seen = set()
    for r in range(8):
        s, d, k = t.coords(r)
        assert t.rank_of(s, d, k) == r
        seen.add((s, d, k))
    assert len(seen) == 8
    assert t.coords(0) == (0, 0, 0) and t.coords(1) == (0, 0, 1) and t.coords(2) == (0, 1, 0)
    assert t.next_stage_rank(0) == 4 and t.prev_stage_rank(4) == 0
    assert t.prev_stage_rank(3) is None and t.next_stage_rank(7) is None
    assert t.dp_group_ranks(1) == [1, 3]
    assert t.tp_group_ranks(6) == [6, 7]
    assert t.pp_group_ranks(3) == [3, 7]


def test_reference_layout_and_names():
    t = Topology(2, pp=2)
    assert t.worker_name(0) == "master" and t.worker_name(1) == "worker1"
    assert t.is_first_stage(0) and t.is_last_stage(1)
    assert t.describe() == "pp2"
    assert Topology(8, 2, 4, 1).describe() == "pp2xdp4"
    assert Topology(8, 4, 1, 2).describe() == "pp4xtp2"


def test_groups_partition_the_world():
    t = Topology(8, pp=2, dp=4)
    for groups in (t.all_dp_groups(), t.all_tp_groups(), t.all_pp_groups()):
        flat = sorted(r for g in groups for r in g)
        assert flat == list(range(8))


def test_bad_sizes():
    with pytest.raises(ValueError):
        Topology(8, pp=3)
    with pytest.raises(ValueError):
        Topology(2, pp=2).coords(2)
