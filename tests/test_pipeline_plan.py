"""The transport plan of the fused pipeline strategy, replayed on the CPU for many stage maps.

`FusedPipelineStrategy.train_step` posts the receive of the NEXT op before it enqueues the current one whenever
`can_post_early` allows it, into the stage engine's own buffers (activation slot ``m % slots``, gradient buffer ``m % 2``).
These tests walk every rank's op list exactly as `train_step` does and check, with a logical clock, that

* a transfer can never land in a buffer whose previous content still has a reader that has not finished,
* every flow (sender rank, receiver rank, direction, stage boundary) sees its sends and its receives in the same order,
* the slot count per stage is enough for the schedule (no micro-batch's activations are overwritten before its backward).
"""
import itertools

import pytest

from trustworthy_dl.parallel.pipeline import make_schedule
from trustworthy_dl.parallel.pipeline_fused import can_post_early, split_layers, stage_slots


def _replay(S, M, owner, kind):
    """owner[s] = rank of stage s.  Returns per-flow (sends, recvs) orders; raises AssertionError on a buffer hazard."""
    sched = make_schedule(S, M, kind)
    one_f_one_b = kind == "1f1b"
    ranks = sorted(set(owner))
    flows = {}
    for rank in ranks:
        my_ops = [op for op in sched if owner[op[1]] == rank]
        end_time = {}                       # op -> index after which it has finished (program order on the compute stream)
        posted = {}                         # op -> earliest time its receive may write its buffer
        def needs_recv(op):
            k, s, m = op
            return (k == "F" and s > 0 and owner[s - 1] != rank) or (k == "B" and s < S - 1 and owner[s + 1] != rank)
        def post(op, t):
            if needs_recv(op) and op not in posted:
                posted[op] = t
                k, s, m = op
                key = (owner[s - 1], rank, 0, s - 1) if k == "F" else (owner[s + 1], rank, 1, s)
                flows.setdefault(key, ([], []))[1].append(m)
        if my_ops:
            post(my_ops[0], -1)
        for i, op in enumerate(my_ops):
            k, s, m = op
            slots = stage_slots(s, S, M, one_f_one_b)
            nxt = my_ops[i + 1] if i + 1 < len(my_ops) else None
            early = nxt is not None and can_post_early(op, nxt, slots)
            if early:
                post(nxt, i - 1)            # may land any time after the previous op finished
            end_time[op] = i
            # sends of this op
            if k == "F" and s < S - 1 and owner[s + 1] != rank:
                flows.setdefault((rank, owner[s + 1], 0, s), ([], []))[0].append(m)
            if k == "B" and s > 0 and owner[s - 1] != rank:
                flows.setdefault((rank, owner[s - 1], 1, s - 1), ([], []))[0].append(m)
            if nxt is not None and not early:
                post(nxt, i)
        # hazards: the receive for (F, s, m) overwrites the slot of micro-batch m - slots, whose last reader is B(s, m - slots);
        # the receive for (B, s, m) overwrites the gradient buffer last read by B(s, m - 2)
        for op, t in posted.items():
            k, s, m = op
            slots = stage_slots(s, S, M, one_f_one_b)
            prev = ("B", s, m - slots) if k == "F" else ("B", s, m - 2)
            if prev[2] >= 0:
                assert prev in end_time and end_time[prev] <= t, f"rank {rank}: {op} may land at {t} before {prev} finished"
            if k == "F" and m - slots >= 0:
                assert end_time[("F", s, m - slots)] <= t
        # co-located neighbours hand activations / gradients over by a device copy out of the producer's buffers: the copy
        # must run before the producer re-uses that buffer (its next forward into the same slot, its next-but-one backward)
        pos = {op: i for i, op in enumerate(my_ops)}
        for (k, s, m) in my_ops:
            if k == "F" and s > 0 and owner[s - 1] == rank:
                again = ("F", s - 1, m + stage_slots(s - 1, S, M, one_f_one_b))
                assert pos[("F", s - 1, m)] < pos[(k, s, m)] and (again not in pos or pos[(k, s, m)] < pos[again]), (rank, k, s, m)
            if k == "B" and s < S - 1 and owner[s + 1] == rank:
                again = ("B", s + 1, m + 2)
                assert pos[("B", s + 1, m)] < pos[(k, s, m)] and (again not in pos or pos[(k, s, m)] < pos[again]), (rank, k, s, m)
        # slot sufficiency: between F(s, m) and B(s, m) no other forward of stage s uses the same slot
        for (k, s, m) in my_ops:
            if k != "F":
                continue
            slots = stage_slots(s, S, M, one_f_one_b)
            lo, hi = my_ops.index(("F", s, m)), my_ops.index(("B", s, m))
            for (k2, s2, m2) in my_ops[lo + 1:hi]:
                assert not (k2 == "F" and s2 == s and m2 % slots == m % slots), (rank, s, m, m2)
    return flows


def _owner_maps(S):
    yield list(range(S))                                            # one stage per rank
    for src in range(S):                                            # one reassignment: src's stage moves to dst
        for dst in range(S):
            if dst != src:
                o = list(range(S)); o[src] = dst
                yield o
    if S >= 4:
        yield [s % 2 for s in range(S)]                             # interleaved on two ranks
        o = list(range(S)); o[1] = 0; o[2] = 0                      # a rank holding three adjacent stages
        yield o


@pytest.mark.parametrize("S,M", [(2, 1), (2, 4), (3, 5), (4, 4), (4, 8), (5, 7), (8, 8), (8, 3)])
@pytest.mark.parametrize("kind", ["1f1b", "gpipe"])
def test_transport_plan_has_no_buffer_hazards_and_matched_flows(S, M, kind):
    for owner in _owner_maps(S):
        flows = _replay(S, M, owner, kind)
        for key, (sends, recvs) in flows.items():
            assert sends == recvs == sorted(sends), (owner, key, sends, recvs)      # one flow = one ordered stream of micro-batches
            assert len(sends) == M


def test_can_post_early_rules():
    assert can_post_early(("F", 1, 0), ("F", 1, 1), 3)             # different slot
    assert not can_post_early(("B", 1, 0), ("F", 1, 3), 3)         # the forward would land in the slot being back-propagated
    assert not can_post_early(("B", 2, 1), ("B", 2, 3), 2)         # same gradient buffer
    assert can_post_early(("B", 2, 1), ("B", 2, 2), 2)
    assert can_post_early(("B", 3, 0), ("F", 2, 1), 1)             # other stage on the same rank: buffers are per stage
    assert not can_post_early(("F", 3, 0), ("F", 3, 1), 1)


def test_split_layers_balances_head_and_embedding():
    assert split_layers(24, 4) == [(0, 6), (6, 13), (13, 20), (20, 24)]   # the last stage carries the LM head (~2.7 blocks)
    for L, S in itertools.product((4, 12, 24, 36), (1, 2, 3, 4, 8)):
        if S > L:
            continue
        r = split_layers(L, S)
        assert r[0][0] == 0 and r[-1][1] == L and all(a < b for a, b in r) and all(r[i][1] == r[i + 1][0] for i in range(S - 1))
