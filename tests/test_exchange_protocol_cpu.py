"""CPU: the two-buffer protocols of the multi-GPU gradient exchange (splatloc_b200/exchange.py,
GraphedViewStep exchange_mode "backward" and "next_step") under random interleavings of the ranks.

The NVLS kernels themselves need NVSwitch hardware (tests/test_gpu_exchange.py); what can be checked on
a CPU is the ordering argument the design rests on -- that with TWO buffers, a per-rank arrival counter
and one wait per step, every rank reads exactly the sum over all ranks of the previous step's rows, no
matter how far the ranks drift apart.  The model: a push becomes one delivery per destination GPU that
lands at an arbitrary later time, but before the pusher's arrive is visible (the arrive is a
release-ordered multimem.red, csrc/exchange.cu); a wait blocks on the local counters; zeroing and
reading act on the rank's own copy at once.  A scheduler advances randomly chosen ranks / deliveries.
Two deliberately broken orders show the model catches what the protocol prevents."""
import random

import numpy as np
import pytest

# what one step enqueues, in stream order (b = step % 2, o = 1 - b); graphs.py: GraphedViewStep._step
SPLIT_PHASE = ("wait_previous", "consume_o", "zero_o", "push_own_b", "arrive")          # push fused in the backward
NEXT_STEP = ("zero_b", "push_prev_o", "arrive", "wait_current", "consume_o")            # push beside the next forward
BROKEN_ZERO_AFTER_ARRIVE = ("wait_previous", "consume_o", "push_own_b", "arrive", "zero_o")
BROKEN_NO_WAIT = ("consume_o", "zero_o", "push_own_b", "arrive")


def rows_of(rank, step, width=4):
    return np.full(width, float((rank + 1) * 1000 + step + 1))


class Rank:
    def __init__(self, rank, world, order, n_steps):
        self.rank, self.world = rank, world
        self.buf = [np.zeros(4), np.zeros(4)]
        self.flags = [0] * world            # local copy of every rank's arrival counter
        self.waits = 0
        self.pending = 0                    # this rank's deliveries still in flight
        self.consumed = []                  # (step, copy of the buffer read)
        self.ops = [(s, op) for s in range(n_steps) for op in order]
        self.pc = 0

    def done(self):
        return self.pc == len(self.ops)


def run(order, world, n_steps, seed, flush):
    rng = random.Random(seed)
    ranks = [Rank(r, world, order, n_steps) for r in range(world)]
    if flush:                               # the sequence's tail: push the last step's rows, arrive, wait (flush_deferred)
        for rk in ranks:
            rk.ops += [(n_steps, "push_prev_o"), (n_steps, "arrive"), (n_steps, "wait_current"), (n_steps, "consume_o")]
    in_flight = []                          # (source rank, destination rank, buffer, values)
    while not all(rk.done() for rk in ranks) or in_flight:
        movable = [("deliver", i) for i in range(len(in_flight))]
        for rk in ranks:
            if rk.done():
                continue
            step, op = rk.ops[rk.pc]
            if op == "arrive" and rk.pending:
                continue                    # release ordering: the arrive is visible only after this rank's rows
            if op == "wait_previous" and min(rk.flags) < rk.waits:
                continue
            if op == "wait_current" and min(rk.flags) < rk.waits + 1:
                continue
            movable.append(("rank", rk.rank))
        assert movable, "deadlock"
        kind, i = rng.choice(movable)
        if kind == "deliver":
            src, dst, b, v = in_flight.pop(i)
            ranks[dst].buf[b] += v
            ranks[src].pending -= 1
            continue
        rk = ranks[i]
        step, op = rk.ops[rk.pc]
        rk.pc += 1
        b = step % 2
        o = 1 - b
        if op in ("wait_previous", "wait_current"):
            rk.waits += 1
        elif op == "consume_o":
            rk.consumed.append((step, rk.buf[o].copy()))
        elif op == "zero_o":
            rk.buf[o][:] = 0
        elif op == "zero_b":
            rk.buf[b][:] = 0
        elif op == "push_own_b":            # this step's rows into buffer b on every GPU
            for dst in range(world):
                in_flight.append((rk.rank, dst, b, rows_of(rk.rank, step)))
            rk.pending += world
        elif op == "push_prev_o":           # the previous step's rows into the previous step's buffer
            if step > 0:
                for dst in range(world):
                    in_flight.append((rk.rank, dst, o, rows_of(rk.rank, step - 1)))
                rk.pending += world
        elif op == "arrive":                # one multimem.red: every GPU's copy of flags[rank] (modelled as at once:
            for other in ranks:             # a late counter only delays a wait, it cannot break one)
                other.flags[rk.rank] += 1
        else:
            raise AssertionError(op)
    return ranks


def check(ranks, world, n_steps, flush):
    for rk in ranks:
        for step, got in rk.consumed:
            if step == 0:
                assert not got.any(), "nothing was pushed before the first step"
                continue
            want = sum(rows_of(r, step - 1) for r in range(world))
            assert np.array_equal(got, want), (rk.rank, step, got, want)
        assert len(rk.consumed) == n_steps + (1 if flush else 0)


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("order,flush", [(SPLIT_PHASE, False), (NEXT_STEP, True)], ids=["push_in_backward", "push_beside_next_forward"])
def test_two_buffers_suffice_under_any_interleaving(order, flush, world):
    for seed in range(60 if world < 8 else 15):
        ranks = run(order, world, n_steps=7, seed=seed, flush=flush)
        check(ranks, world, 7, flush)


@pytest.mark.parametrize("order", [BROKEN_ZERO_AFTER_ARRIVE, BROKEN_NO_WAIT], ids=["zero_after_arrive", "no_wait"])
def test_model_catches_broken_orders(order):
    failures = 0
    for seed in range(60):
        try:
            check(run(order, 3, n_steps=7, seed=seed, flush=False), 3, 7, False)
        except AssertionError:
            failures += 1
    assert failures > 0, "the model cannot tell a broken protocol from a correct one"
