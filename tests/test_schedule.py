"""Schedule generator invariants (pure C++/Python, no devices)."""
import pytest
from hypothesis import given, settings, strategies as st

from simple_distributed_machine_learning_b200.parallel import schedule as S
from simple_distributed_machine_learning_b200.parallel.schedule import OpKind

KINDS = ["gpipe", "1f1b", "zb-h1", "zb-h2"]


@settings(max_examples=60, deadline=None)
@given(kind=st.sampled_from(KINDS), p=st.integers(1, 8), m=st.integers(1, 24))
def test_every_microbatch_once_and_ordered(kind, p, m):
    for rank in range(p):
        ops = S.generate(kind, p, m, rank)
        S.validate(ops, m)
        f = [o.mb for o in ops if o.kind == OpKind.F]
        b = [o.mb for o in ops if o.kind in (OpKind.B, OpKind.BW)]
        w = [o.mb for o in ops if o.kind in (OpKind.W, OpKind.BW)]
        assert f == list(range(m)) and b == list(range(m)) and sorted(w) == list(range(m))
        if S.splits_backward(kind):
            assert all(o.kind != OpKind.BW for o in ops)
        else:
            assert all(o.kind not in (OpKind.B, OpKind.W) for o in ops)


@settings(max_examples=60, deadline=None)
@given(kind=st.sampled_from(KINDS), p=st.integers(1, 8), m=st.integers(1, 24))
def test_no_deadlock_and_memory_bound(kind, p, m):
    r = S.simulate(kind, p, m)   # raises if the cross-rank dependencies dead-lock
    assert r.makespan >= 3 * m - 1e-9
    for rank, peak in enumerate(r.peak_inflight):
        if kind in ("1f1b", "zb-h1"):
            assert peak <= min(m, p - rank)          # the 1F1B activation footprint
        elif kind == "zb-h2":
            assert peak <= min(m, max(1, 2 * (p - rank) - 1))
        else:
            assert peak == m


def test_bubble_formulas():
    # GPipe / 1F1B with F=B=W=1: the fused BW op releases dX after its dgrad half, so the
    # ramp costs (p-1)*(F+B): bubble = 2(p-1) / (3m + 2(p-1))
    for p, m in ((2, 8), (4, 8), (8, 8)):
        for kind in ("gpipe", "1f1b"):
            r = S.simulate(kind, p, m)
            assert r.bubble_fraction == pytest.approx(2 * (p - 1) / (3 * m + 2 * (p - 1)), abs=1e-9)
    # zero-bubble H1 at two stages and 8 micro-batches: under the 10 % target of BASELINE.json
    assert S.simulate("zb-h1", 2, 8).bubble_fraction < 0.10
    for p in (2, 4, 8):
        assert S.simulate("zb-h1", p, 8).makespan < S.simulate("1f1b", p, 8).makespan


def test_lockstep_is_the_reference_loop():
    assert [repr(o) for o in S.generate("lockstep", 2, 1, 0)] == ["F0", "BW0"]
    with pytest.raises(ValueError):
        S.generate("lockstep", 2, 4, 0)
    with pytest.raises(ValueError):
        S.generate("nope", 2, 4, 0)


def test_send_recv_pairs_match_across_neighbours():
    for kind in KINDS:
        p, m = 4, 6
        all_ops = S.generate_all(kind, p, m)
        for s in range(p - 1):
            sends = [o.mb for o in all_ops[s] if o.kind == OpKind.F]
            recvs = [o.mb for o in all_ops[s + 1] if o.kind == OpKind.F]
            assert sends == recvs
            gsend = [o.mb for o in all_ops[s + 1] if o.kind in (OpKind.B, OpKind.BW)]
            grecv = [o.mb for o in all_ops[s] if o.kind in (OpKind.B, OpKind.BW)]
            assert gsend == grecv


def test_zero_bubble_schedule_meets_the_single_step_lower_bound():
    """With unit F / B / W costs and one model chunk per rank, the last stage cannot start before
    p - 1 forward hops and then has 3m units of work, so a step cannot finish before p - 1 + 3m:
    bubble >= (p - 1) / (3m + p - 1).  The generator's zero-bubble lists reach exactly that (so at
    8 micro-batches p = 4 cannot go below 11.1 % and p = 8 not below 22.6 % without interleaving;
    BASELINE.md section 4 lists the micro-batch counts that get under 10 %)."""
    from simple_distributed_machine_learning_b200.parallel import schedule as S
    for p in (2, 4, 8):
        for m in (8, 16, 32):
            bound = (p - 1) / (3 * m + p - 1)
            for kind in ("zb-h1", "zb-h2"):
                got = S.simulate(kind, p, m).mean_bubble_fraction
                assert abs(got - bound) < 1e-9, (kind, p, m, got, bound)
            assert S.simulate("1f1b", p, m).mean_bubble_fraction >= bound - 1e-9


def test_interleaved_zero_bubble_lists_are_valid_and_reach_the_chunked_bound():
    """v model chunks per rank: every (chunk, micro-batch) gets F < B < W, chunk-wise in order, the
    lists never dead-lock in the simulator, and with unit STAGE costs (1 / v per chunk op) the bubble is
    exactly ((p - 1) / v) / (3m + (p - 1) / v): under 10 % for p = 4 at 8 micro-batches with 2 chunks."""
    from simple_distributed_machine_learning_b200.parallel import schedule as S
    for p in (2, 3, 4, 8):
        for v in (1, 2, 4):
            for m in (1, 4, 8, 16):
                for r in range(p):
                    ops = S.generate_interleaved(p, v, m, r)
                    assert len(ops) == 3 * m * v
                    S.validate_interleaved(ops, m, v)
                c = S.Costs(1.0 / v, 1.0 / v, 1.0 / v, 0.0)
                got = S.simulate_interleaved(p, v, m, c).mean_bubble_fraction
                bound = ((p - 1) / v) / (3 * m + (p - 1) / v)
                assert got >= bound - 1e-9
                if m >= p:      # fewer micro-batches than stages: the first rank's round trip is the floor instead
                    assert got <= bound + 1e-9, (p, v, m, got, bound)
    assert S.simulate_interleaved(4, 2, 8, S.Costs(0.5, 0.5, 0.5, 0.0)).mean_bubble_fraction < 0.06
    # v = 1 is the plain zero-bubble schedule
    assert abs(S.simulate_interleaved(8, 1, 8).mean_bubble_fraction - S.simulate("zb-h1", 8, 8).mean_bubble_fraction) < 1e-9


def test_interleaved_lists_with_uneven_costs_and_transfer_time_do_not_deadlock():
    from simple_distributed_machine_learning_b200.parallel import schedule as S
    for costs in (S.Costs(1.0, 2.0, 1.0, 0.1), S.Costs(0.7, 0.9, 1.3, 0.0), S.Costs(1.0, 1.0, 0.5, 0.5)):
        for p, v, m in ((4, 2, 8), (3, 3, 5), (8, 2, 4)):
            r = S.simulate_interleaved(p, v, m, costs)
            assert r.makespan > 0 and all(b > 0 for b in r.busy)
            for rank in range(p):
                S.validate_interleaved(S.generate_interleaved(p, v, m, rank, costs), m, v)
