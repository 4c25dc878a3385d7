"""CPU-only: an executable model of the LL small-message protocol (`ll_body`, src/coll/coll_experimental.cuh).

The device kernel could not be run when it was written, so its protocol — flag-in-data packets, two slot parities, the
"acknowledge every call to every peer, wait for n-2 before reusing a parity" rule, no barrier at either end — is
restated here as coroutines (one per rank and call, one `yield` per memory operation) and driven by a random scheduler
that interleaves ranks arbitrarily while keeping each rank's calls in stream order. Ranks may run as far ahead as the
protocol lets them; the scheduler is biased so that some rank is usually much slower than the others. Checked: no
deadlock, every consumer's result equals the rank-order sum, under rotating roots, reduce/allreduce mixes and message
sizes that shrink and grow (so slots keep stale packets of older, longer messages).

The multi-CTA instances of the kernel use a fixed partition of the message with one set of acknowledgement words per CTA,
so every CTA column runs exactly this protocol on its own share; one column is what is modelled.

The model has teeth: with the n-2 rule switched off it finds corrupted results (second test)."""
import random

import pytest


class Rank:
    def __init__(self, P, maxvec):
        # slots[parity][source][vector] = (data, flag); ack[source] = last LL call that source has finished
        self.slots = [[[(0, 0)] * maxvec for _ in range(P)] for _ in range(2)]
        self.ack = [0] * P
        self.results = {}  # seq -> list


def ll_call(ranks, me, seq, root, send, flow_control=True):
    """One rank's kernel for LL call `seq` as a coroutine; yields True after progress, False while spinning."""
    P = len(ranks)
    mine = ranks[me]
    parity = seq & 1
    consumer = root < 0 or root == me
    consumers = [d for d in range(P) if d != me and (root < 0 or d == root)]
    # 1. flow control: the slots of this parity were last used by call seq-2
    if flow_control and seq > 2:
        for d in consumers:
            while mine.ack[d] < seq - 2:
                yield False
    # 2. push my packets into every consumer's slot
    for d in consumers:
        for v, x in enumerate(send):
            ranks[d].slots[parity][me][v] = (x, seq)
            yield True
    # 3. consume: wait for every peer's packet, combine in rank order
    if consumer:
        out = []
        for v in range(len(send)):
            acc = 0
            for p in range(P):
                if p == me:
                    acc += send[v]
                    continue
                while mine.slots[parity][p][v][1] != seq:
                    yield False
                acc += mine.slots[parity][p][v][0]
                yield True
            out.append(acc)
        mine.results[seq] = out
    # 4. acknowledge the call to every peer
    for p in range(P):
        if p != me:
            ranks[p].ack[me] = seq
            yield True


def run_schedule(P, calls, seed, flow_control=True):
    """calls: list of (root, nvec). Returns (results per rank, expected per call) or raises on deadlock."""
    rng = random.Random(seed)
    maxvec = max(n for _, n in calls)
    ranks = [Rank(P, maxvec) for _ in range(P)]
    data = [[[rng.randrange(1 << 20) for _ in range(n)] for _ in range(P)] for _, n in calls]
    queues = [[ll_call(ranks, r, k + 1, root, data[k][r], flow_control) for k, (root, _) in enumerate(calls)] for r in range(P)]
    pos = [0] * P
    weights = [rng.choice([1, 1, 1, 20, 200]) for _ in range(P)]  # some ranks are usually much faster than others
    idle = 0
    while any(pos[r] < len(calls) for r in range(P)):
        live = [r for r in range(P) if pos[r] < len(calls)]
        r = rng.choices(live, weights=[weights[q] for q in live])[0]
        if rng.random() < 0.001:
            weights = [rng.choice([1, 1, 1, 20, 200]) for _ in range(P)]
        try:
            progressed = next(queues[r][pos[r]])
        except StopIteration:
            pos[r] += 1
            progressed = True
        idle = 0 if progressed else idle + 1
        if idle > 200000:
            raise AssertionError(f"deadlock: ranks at calls {pos}, acks {[rk.ack for rk in ranks]}")
    want = [[sum(data[k][r][v] for r in range(P)) for v in range(n)] for k, (_, n) in enumerate(calls)]
    return ranks, want


def check(P, calls, seed, flow_control=True):
    ranks, want = run_schedule(P, calls, seed, flow_control)
    bad = []
    for k, (root, _) in enumerate(calls):
        for r in range(P):
            if root < 0 or root == r:
                if ranks[r].results.get(k + 1) != want[k]:
                    bad.append((k + 1, r))
            else:
                assert k + 1 not in ranks[r].results
    return bad


def random_calls(rng, P, ncalls):
    calls = []
    for _ in range(ncalls):
        kind = rng.random()
        root = -1 if kind < 0.4 else rng.randrange(P)
        calls.append((root, rng.choice([1, 1, 2, 3, 5, 8])))
    return calls


@pytest.mark.parametrize("P", [2, 3, 4, 8])
def test_ll_protocol_model_is_correct_and_deadlock_free(P):
    for seed in range(12):
        rng = random.Random(1000 * P + seed)
        calls = random_calls(rng, P, 60)
        assert check(P, calls, seed) == [], (P, seed)
    # the patterns that stress the parity rule hardest: one fixed root with eager senders, and strict alternation
    assert check(P, [(0, 4)] * 80, 7) == []
    assert check(P, [(i % P, 1 + i % 3) for i in range(80)], 8) == []
    assert check(P, [(-1, 2), (P - 1, 5)] * 40, 9) == []


def test_the_model_catches_a_missing_flow_control_rule():
    """Without the n-2 acknowledgement rule a fast sender overwrites packets its consumer has not read yet."""
    found = False
    for seed in range(40):
        rng = random.Random(seed)
        calls = random_calls(rng, 4, 60)
        try:
            if check(4, calls, seed, flow_control=False):
                found = True
                break
        except AssertionError:  # a consumer waiting for a packet that was overwritten by a later call never wakes up
            found = True
            break
    assert found
