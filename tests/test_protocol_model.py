"""Randomised-schedule model check of the two cross-GPU slot protocols ("flag-protocol stress test").

The kernels themselves are tested on GPUs; this file tests the PROTOCOL ARGUMENTS they rely on, by executing an abstract
model of W ranks under thousands of random interleavings and checking that no slot is ever overwritten between the moment
its reader starts and finishes reading it, and that every reader sees the sequence number it expects:

* mailbox protocol of the fused exchanges (ewdml/parallel/native_exchange.py): 2 parities, NO acknowledgements --
  sound because every rank both sends and receives in every exchange;
* window protocol of the collective engine (csrc/comm/collectives.cu): 2 parities + per-writer acknowledgement flags --
  needed because rooted operations (broadcast) have a writer that never waits for anybody.

Each protocol is also run in a deliberately broken variant (one parity / no acknowledgements) to show the model detects
the overwrite it is meant to exclude.
"""
import random

import pytest


class Violation(Exception):
    pass


class Rank:
    """A rank is a generator of micro-steps; `yield cond` blocks until cond() is true, `yield None` is a scheduling point."""

    def __init__(self, gen):
        self.gen, self.wait, self.done = gen, None, False

    def runnable(self):
        return not self.done and (self.wait is None or self.wait())

    def step(self):
        try:
            self.wait = next(self.gen)
        except StopIteration:
            self.done = True


def run(ranks, rng, max_steps=200000):
    for _ in range(max_steps):
        ready = [r for r in ranks if r.runnable()]
        if not ready:
            if all(r.done for r in ranks):
                return
            raise Violation("deadlock")
        rng.choice(ready).step()
    raise Violation("did not finish")


# ---------------------------------------------------------------------------------------------------------------------
# mailbox protocol: slot[owner][parity][src], flag[owner][parity][src]; exchange q: send(q) then recv(q) on every rank
# ---------------------------------------------------------------------------------------------------------------------
def mailbox_program(me, W, Q, slot, flag, parities):
    for q in range(1, Q + 1):
        par = q % parities
        for dst in range(W):                      # payload stores (scheduling point between peers), then the flags
            slot[dst][par][me] = q
            yield None
        for dst in range(W):
            flag[dst][par][me] = q
            yield None
        for src in range(W):                      # acquire every sender's flag
            yield (lambda s=src, p=par, qq=q: flag[me][p][s] >= qq)
        for src in range(W):                      # read: begin ... end, the slot must hold exchange q throughout
            if slot[me][par][src] != q:
                raise Violation(f"rank {me} exchange {q}: slot of {src} holds {slot[me][par][src]} at read start")
            yield None
            if slot[me][par][src] != q:
                raise Violation(f"rank {me} exchange {q}: slot of {src} overwritten during the read")


def run_mailbox(W, Q, parities, seed):
    rng = random.Random(seed)
    slot = [[[0] * W for _ in range(parities)] for _ in range(W)]
    flag = [[[0] * W for _ in range(parities)] for _ in range(W)]
    run([Rank(mailbox_program(r, W, Q, slot, flag, parities)) for r in range(W)], rng)


@pytest.mark.parametrize("W", [2, 3, 5])
def test_mailbox_two_parities_without_acks_is_sound(W):
    for seed in range(300):
        run_mailbox(W, 6, 2, seed)


def test_mailbox_single_parity_is_caught():
    with pytest.raises(Violation):
        for seed in range(300):
            run_mailbox(3, 6, 1, seed)


# ---------------------------------------------------------------------------------------------------------------------
# window protocol of the collective engine: an operation = (put_mask, src_mask) per rank; slot[owner][parity][writer]
# ---------------------------------------------------------------------------------------------------------------------
def engine_program(me, W, ops, slot, data, ack, use_acks):
    last_put = [[0, 0] for _ in range(W)]
    for q, op in enumerate(ops, start=1):
        par = q & 1
        puts, srcs = op(me, W)
        if use_acks:
            for p in puts:                        # the destination must have consumed my previous put of this parity
                need = last_put[p][par]
                if need:
                    yield (lambda pp=p, n=need: ack[me][pp] >= n)
        for p in puts:
            slot[p][par][me] = q
            last_put[p][par] = q
            yield None
        for p in puts:
            data[p][me] = q
            yield None
        for s in srcs:
            yield (lambda ss=s, qq=q: data[me][ss] >= qq)
        for s in srcs:
            if slot[me][par][s] != q:
                raise Violation(f"rank {me} op {q}: slot of {s} holds {slot[me][par][s]} at read start")
            yield None
            if slot[me][par][s] != q:
                raise Violation(f"rank {me} op {q}: slot of {s} overwritten during the read")
        for s in srcs:
            ack[s][me] = q
            yield None


def bcast(root):
    return lambda me, W: ([p for p in range(W) if p != root] if me == root else [], [] if me == root else [root])


def allgather(me, W):
    return list(range(W)), list(range(W))


def gather(root):
    return lambda me, W: ([root], list(range(W)) if me == root else [])


def scatter(root):
    return lambda me, W: (list(range(W)) if me == root else [], [root])


def run_engine(W, ops, use_acks, seed):
    rng = random.Random(seed)
    slot = [[[0] * W for _ in range(2)] for _ in range(W)]
    data = [[0] * W for _ in range(W)]
    ack = [[0] * W for _ in range(W)]
    run([Rank(engine_program(r, W, ops, slot, data, ack, use_acks)) for r in range(W)], rng)


@pytest.mark.parametrize("W", [2, 4])
def test_engine_window_protocol_with_acks_is_sound(W):
    rng = random.Random(7)
    makers = [lambda: bcast(rng.randrange(W)), lambda: allgather, lambda: gather(rng.randrange(W)), lambda: scatter(rng.randrange(W))]
    for seed in range(200):
        ops = [rng.choice(makers)() for _ in range(8)]
        run_engine(W, ops, True, seed)
    for seed in range(200):                       # the worst case: a root that broadcasts again and again without ever waiting
        run_engine(W, [bcast(0)] * 8, True, seed)


def test_engine_without_acks_lets_a_fast_root_overrun_a_slow_reader():
    with pytest.raises(Violation):
        for seed in range(300):
            run_engine(3, [bcast(0)] * 8, False, seed)


# ---------------------------------------------------------------------------------------------------------------------
# parameter-server mailboxes (K3): workers push into the server's slots, the server pulls into one slot per worker
# ---------------------------------------------------------------------------------------------------------------------
def ps_program(me, W, Q, push, pflag, pull, lflag, parities):
    for q in range(1, Q + 1):
        par = q % parities
        if me != 0:
            push[par][me] = q
            yield None
            pflag[par][me] = q
            yield (lambda p=par, qq=q: lflag[me][p] >= qq)
            if pull[me][par] != q:
                raise Violation(f"worker {me} step {q}: pull slot holds {pull[me][par]}")
            yield None
            if pull[me][par] != q:
                raise Violation(f"worker {me} step {q}: pull slot overwritten during the read")
        else:
            for w in range(1, W):
                yield (lambda ww=w, p=par, qq=q: pflag[p][ww] >= qq)
            for w in range(1, W):
                if push[par][w] != q:
                    raise Violation(f"server step {q}: push slot of {w} holds {push[par][w]}")
                yield None
                if push[par][w] != q:
                    raise Violation(f"server step {q}: push slot of {w} overwritten during the read")
            for w in range(1, W):
                pull[w][par] = q
                yield None
                lflag[w][par] = q
                yield None


def run_ps(W, Q, parities, seed):
    rng = random.Random(seed)
    push = [[0] * W for _ in range(parities)]
    pflag = [[0] * W for _ in range(parities)]
    pull = [[0] * parities for _ in range(W)]
    lflag = [[0] * parities for _ in range(W)]
    run([Rank(ps_program(r, W, Q, push, pflag, pull, lflag, parities)) for r in range(W)], rng)


@pytest.mark.parametrize("W", [2, 4])
def test_parameter_server_mailboxes_are_sound(W):
    for seed in range(300):
        run_ps(W, 6, 2, seed)
