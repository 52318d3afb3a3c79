"""Exhaustive interleaving check of the two hand-rolled cross-GPU protocols, as small models.

compute-sanitizer's racecheck is closed on the GPU pool, and the hardware stress tests
(tests/test_stress_gpu.py) can only sample schedules.  Here the PROTOCOLS -- not the kernels -- are
written down as two little programs of atomic steps over shared words, exactly as csrc/boundary.cu and
csrc/dp_allreduce_sgd.cu use them (32-bit counters compared by signed difference, flags per slot, one
acknowledgement per round; (value, epoch) pairs in parity-selected buffers), and EVERY interleaving of the
two programs is explored: no slot may be read before it holds this round's data, none overwritten before
its reader is done, and no reachable state may be stuck.  The same checker must FIND the bug when the
guard is removed (otherwise it proves nothing).  The reference's only guard is a mutex around its forward
(/root/reference/simple_distributed.py:66, :72).
"""
from collections import deque

M32 = 1 << 32


def _ge(a, b):
    """a >= b on the signed 32-bit difference (wrap-safe, as the kernels compare)."""
    d = (a - b) % M32
    return d < (1 << 31)


def _explore(programs, init, step):
    """BFS over all interleavings.  `programs`: number of steps of each process; `step(state, who, pc)`
    returns None when that step is blocked, else the new shared state (or raises AssertionError)."""
    start = (tuple(0 for _ in programs), init)
    seen = {start}
    todo = deque([start])
    finished = False
    while todo:
        pcs, shared = todo.popleft()
        if all(pc == n for pc, n in zip(pcs, programs)):
            finished = True
            continue
        moved = False
        for who, n in enumerate(programs):
            if pcs[who] == n:
                continue
            nxt = step(shared, who, pcs[who])
            if nxt is None:
                continue
            moved = True
            state = (tuple(pc + (1 if i == who else 0) for i, pc in enumerate(pcs)), nxt)
            if state not in seen:
                seen.add(state)
                todo.append(state)
        assert moved, f"dead-lock at {pcs} {shared}"
    assert finished
    return len(seen)


# ---------------------------------------------------------------------------------------------------
# stage boundary: push + flag per slot, one acknowledgement per round (csrc/boundary.cu)
# ---------------------------------------------------------------------------------------------------
def _boundary_model(slots, rounds, start, with_ack=True):
    # sender program per round: [wait ack] (write slot i, flag i)*slots ; receiver: (wait flag i, read slot i)*slots, ack
    s_len = rounds * (1 + 2 * slots)
    r_len = rounds * (2 * slots + 1)
    init = (tuple([None] * slots), tuple([start] * slots), start)       # slot data, flags, ack

    def step(shared, who, pc):
        data, flags, ack = shared
        if who == 0:
            k, j = divmod(pc, 1 + 2 * slots)
            cnt = (start + k) % M32
            if j == 0:
                if with_ack and not _ge(ack, cnt):
                    return None                                         # receiver still reads last round's slots
                return shared
            i, phase = divmod(j - 1, 2)
            if phase == 0:
                d = list(data)
                d[i] = (cnt + 1) % M32
                return (tuple(d), flags, ack)
            f = list(flags)
            f[i] = (cnt + 1) % M32
            return (data, tuple(f), ack)
        k, j = divmod(pc, 2 * slots + 1)
        cnt = (start + k) % M32
        if j == 2 * slots:
            return (data, flags, (cnt + 1) % M32)                       # acknowledge the round
        i, phase = divmod(j, 2)
        if phase == 0:
            return shared if _ge(flags[i], (cnt + 1) % M32) else None   # wait for the slot's flag
        assert data[i] == (cnt + 1) % M32, f"round {k}: slot {i} holds {data[i]}, expected {(cnt + 1) % M32}"
        return shared

    return [s_len, r_len], init, step


def test_boundary_push_flag_ack_protocol_all_interleavings_incl_counter_wrap():
    for start in (0, M32 - 2):                      # counters wrap inside the run
        n = _explore(*_boundary_model(slots=2, rounds=3, start=start))
        assert n >= 40          # (the two programs are tightly coupled: few distinct states, all visited)


def test_boundary_model_finds_the_overwrite_when_the_acknowledgement_is_removed():
    try:
        _explore(*_boundary_model(slots=2, rounds=2, start=0, with_ack=False))
    except AssertionError as e:
        assert "expected" in str(e)
    else:
        raise AssertionError("the checker did not notice a sender overwriting a slot that was still being read")


# ---------------------------------------------------------------------------------------------------
# low-latency all-reduce: every rank pushes (value, epoch) into the parity buffer of its peer, polls its own
# ---------------------------------------------------------------------------------------------------
def _ll_model(steps, start, parities=2):
    # per step each rank: push (its value, epoch) to the peer's buffer[epoch % parities]; poll own buffer until epoch matches; read
    n = steps * 2

    def step(shared, who, pc):
        bufs = shared                                   # bufs[rank][parity] = (value, epoch)
        k, phase = divmod(pc, 2)
        epoch = (start + k + 1) % M32
        par = epoch % parities
        peer = 1 - who
        if phase == 0:
            b = [list(x) for x in bufs]
            b[peer][par] = (("v", who, k), epoch)
            return tuple(tuple(x) for x in b)
        val, e = bufs[who][par]
        if e != epoch:
            return None                                 # poll
        assert val == ("v", peer, k), f"rank {who} step {k}: read {val}"
        return shared

    init = tuple(tuple((None, start) for _ in range(parities)) for _ in range(2))
    return [n, n], init, step


def test_ll_allreduce_parity_buffers_all_interleavings_incl_epoch_wrap():
    for start in (0, M32 - 2):
        assert _explore(*_ll_model(steps=4, start=start)) >= 25


def test_ll_model_finds_the_overwrite_with_a_single_buffer():
    """One buffer instead of two parities: a rank that finished step k may push step k + 1 over a pair its
    peer has not read yet (the peer then polls for an epoch that is gone: stuck), which is why the kernel
    alternates buffers."""
    try:
        _explore(*_ll_model(steps=3, start=0, parities=1))
    except AssertionError:
        return
    raise AssertionError("the checker did not notice the single-buffer overwrite")
