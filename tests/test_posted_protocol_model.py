"""CPU-only: an executable model of the posted reduce-to-root protocol (`posted_body`, src/coll/coll_posted.cuh).

One CTA column is modelled (the kernel's partition is fixed: chunk k belongs to CTA k mod grid on every rank in every call, so
every column runs exactly this protocol on its own chunks). Per rank: two staging slots in its OWN memory (call parity), at the
root one ready word per source, everywhere one acknowledgement word per acknowledging rank. A call is a coroutine with one
`yield` per memory operation; a random scheduler interleaves ranks arbitrarily, keeps each rank's calls in stream order and is
biased so that some rank — often the root — is much slower than the others, which is the situation the protocol exists for:
non-root ranks stage, signal and LEAVE, so they run ahead of the root by as much as the acknowledgements allow.

Checked: no deadlock; the root's result of every call is the sum of what every rank contributed to THAT call, with roots that
stay, alternate and rotate, and message lengths that shrink and grow (slots keep stale data of longer, older messages).

The model has teeth: without the wait for the acknowledgement of the call two back, and with the wait directed at the wrong rank
(the current root instead of the root of the call two back — the mistake a single ack word would make), it finds corrupted
results."""
import random

import pytest


class Rank:
    def __init__(self, P, maxvec):
        self.slot = [[0] * maxvec for _ in range(2)]  # my staged contribution, per call parity (read by the root through the switch)
        self.ready = [0] * P                          # at the root: ready[src] = last call src has staged
        self.ack = [0] * P                            # ack[r] = last call that root r has finished reading
        self.results = {}


def posted_call(ranks, me, seq, root, prev_root, send, wait_for="prev_root"):
    """One rank's kernel for posted call `seq`; yields True after progress, False while spinning."""
    P = len(ranks)
    mine = ranks[me]
    parity = seq & 1
    # 1. my slot of this parity was last read by the root of call seq-2
    waiter = {"prev_root": prev_root, "current_root": root, "nobody": -1}[wait_for]
    if seq > 2 and waiter >= 0 and waiter != me:
        while mine.ack[waiter] < seq - 2:
            yield False
    # 2. stage (the root stages too: the switch reads all P slots)
    for v, x in enumerate(send):
        mine.slot[parity][v] = x
        yield True
    if me != root:
        ranks[root].ready[me] = seq  # release
        yield True
        return
    # 3. the root: wait for every peer, reduce through the "switch" (reads every rank's slot), acknowledge
    for p in range(P):
        if p != me:
            while mine.ready[p] < seq:
                yield False
    out = []
    for v in range(len(send)):
        acc = 0
        for p in range(P):
            acc += ranks[p].slot[parity][v]
            yield True
        out.append(acc)
    mine.results[seq] = out
    for p in range(P):
        if p != me:
            ranks[p].ack[me] = seq
            yield True


def run_schedule(P, calls, seed, wait_for="prev_root", slow_root=False):
    rng = random.Random(seed)
    maxvec = max(n for _, n in calls)
    ranks = [Rank(P, maxvec) for _ in range(P)]
    data = [[[rng.randrange(1 << 20) for _ in range(n)] for _ in range(P)] for _, n in calls]
    roots = [root for root, _ in calls]
    queues = [[posted_call(ranks, r, k + 1, roots[k], roots[k - 2] if k >= 2 else -1, data[k][r], wait_for) for k in range(len(calls))] for r in range(P)]
    pos = [0] * P
    weights = [rng.choice([1, 1, 1, 20, 200]) for _ in range(P)]
    idle = 0
    while any(pos[r] < len(calls) for r in range(P)):
        live = [r for r in range(P) if pos[r] < len(calls)]
        w = [weights[q] for q in live]
        if slow_root:  # whoever is root of the call a rank is in gets little time: everybody else runs ahead
            w = [1 if roots[pos[q]] == q else 50 for q in live]
        r = rng.choices(live, weights=w)[0]
        if rng.random() < 0.001:
            weights = [rng.choice([1, 1, 1, 20, 200]) for _ in range(P)]
        try:
            progressed = next(queues[r][pos[r]])
        except StopIteration:
            pos[r] += 1
            progressed = True
        idle = 0 if progressed else idle + 1
        if idle > 200000:
            raise AssertionError(f"deadlock: ranks at calls {pos}")
    want = [[sum(data[k][r][v] for r in range(P)) for v in range(n)] for k, (_, n) in enumerate(calls)]
    return ranks, want


def check(P, calls, seed, **kw):
    ranks, want = run_schedule(P, calls, seed, **kw)
    return [(k + 1, root) for k, (root, _) in enumerate(calls) if ranks[root].results.get(k + 1) != want[k]]


def random_calls(rng, P, ncalls):
    calls, root = [], 0
    for _ in range(ncalls):
        if rng.random() < 0.35:
            root = rng.randrange(P)
        calls.append((root, rng.choice([1, 1, 2, 3, 5, 8])))
    return calls


@pytest.mark.parametrize("P", [2, 3, 4, 8])
def test_posted_protocol_model_is_correct_and_deadlock_free(P):
    n = 60 if P <= 4 else 36  # the cost of a schedule grows with P x calls; 36 calls still wrap the two parities 18 times
    for seed in range(6):
        rng = random.Random(2000 * P + seed)
        calls = random_calls(rng, P, n)
        assert check(P, calls, seed) == [], (P, seed)
        assert check(P, calls, seed, slow_root=True) == [], (P, seed)
    assert check(P, [(0, 4)] * n, 7, slow_root=True) == []                                   # one fixed root, eager peers
    assert check(P, [(i % P, 1 + i % 3) for i in range(n)], 8, slow_root=True) == []        # the root moves every call
    assert check(P, [((i // 2) % P, 2 + i % 2) for i in range(n)], 9, slow_root=True) == []  # ... every second call
    assert check(P, [(0, 3), (P - 1, 5)] * (n // 2), 10, slow_root=True) == []                      # two roots alternate: parity == root


def _finds_a_fault(**kw):
    for seed in range(60):
        rng = random.Random(seed)
        calls = random_calls(rng, 4, 60)
        try:
            if check(4, calls, seed, slow_root=True, **kw):
                return True
        except AssertionError:
            return True
    return False


def test_the_model_catches_missing_and_misdirected_flow_control():
    assert _finds_a_fault(wait_for="nobody")        # no n-2 rule: a rank restages a slot its root has not read yet
    assert _finds_a_fault(wait_for="current_root")  # waiting for the wrong rank's acknowledgement when the root has moved
