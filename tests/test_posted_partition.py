"""CPU-only: the index arithmetic of the posted reduce (`posted_body`, src/coll/coll_posted.cuh), restated.

The kernel's flow control rests on a FIXED partition: chunk k (kPostedChunkVecs 16-byte vectors) belongs to CTA k mod grid in every
call, whatever the message length, so that the bytes of a staging slot are only ever touched by one CTA index. Two loop nests walk
it: the vector nest (staging, the 16-byte multimem form, the pull path) and the unit nest of the 4- and 8-byte multimem forms, in
which unit m of a thread sits at j + m x kPostedChunkVecs inside the chunk so that a warp's instruction covers consecutive units.
Checked for both: every vector / unit of the message is visited exactly once, by the CTA that owns its chunk, for message
lengths around every boundary, several grids and CTA sizes (the loopback launch uses smaller ones)."""
import pytest

CH = 512  # kPostedChunkVecs
U = 4     # chunks in flight per thread


def vector_nest(nvec, G, threads):
    """-> {vector: cta} as visited by the staging / reduce loops."""
    seen = {}
    nchunks = (nvec + CH - 1) // CH
    for b in range(G):
        if b * CH >= nvec:  # has_work
            continue
        for k0 in range(b, nchunks, G * U):
            for tid in range(threads):
                for j in range(tid, CH, threads):
                    for u in range(U):
                        k = k0 + u * G
                        v = k * CH + j
                        if k < nchunks and v < nvec:
                            assert v not in seen, (v, b)
                            seen[v] = b
    return seen


def unit_nest(nbytes, G, threads, B):
    """-> {unit index: cta} as visited by the coalesced loop of the B-byte multimem forms."""
    seen = {}
    nvec = (nbytes + 15) // 16
    nchunks = (nvec + CH - 1) // CH
    M = 16 // B
    for b in range(G):
        if b * CH >= nvec:
            continue
        for k0 in range(b, nchunks, G * U):
            for tid in range(threads):
                for j in range(tid, CH, threads):
                    for u in range(U):
                        k = k0 + u * G
                        for m in range(M):
                            off = k * CH * 16 + (j + m * CH) * B
                            if k < nchunks and off < nbytes:
                                assert off % B == 0 and off // B not in seen
                                seen[off // B] = b
    return seen


SIZES_VEC = [1, 2, CH - 1, CH, CH + 1, 3 * CH + 7, 5 * CH]


@pytest.mark.parametrize("G,threads", [(4, 64), (3, 32), (7, 128)])
def test_vector_nest_visits_every_vector_once_in_its_owner(G, threads):
    for extra in (0, G * CH - 1, G * CH, G * CH + 1, G * U * CH, G * U * CH + CH + 3):
        for base in SIZES_VEC:
            nvec = base + extra
            seen = vector_nest(nvec, G, threads)
            assert len(seen) == nvec
            assert all(b == (v // CH) % G for v, b in seen.items())


@pytest.mark.parametrize("B", [4, 8])
@pytest.mark.parametrize("G,threads", [(4, 64), (3, 128)])
def test_unit_nest_visits_every_unit_once_in_its_owner_and_coalesced(G, threads, B):
    for nunits in (1, 3, 16 // B, 16 // B + 1, CH * 16 // B - 1, CH * 16 // B, CH * 16 // B + 1, (G * U + 1) * CH * 16 // B + 5, (2 * G + 1) * CH * 16 // B - 1):
        nbytes = nunits * B
        seen = unit_nest(nbytes, G, threads, B)
        assert len(seen) == nunits
        assert all(b == (unit * B // 16 // CH) % G for unit, b in seen.items())
    # coalescing: for fixed (chunk, m), consecutive threads read consecutive units
    k, m = 1, 1
    offs = [k * CH * 16 + (j + m * CH) * B for j in range(32)]
    assert all(b - a == B for a, b in zip(offs, offs[1:]))


def test_partition_does_not_depend_on_the_message_length():
    """What CTA b touches in a short message is a subset of what it touches in a long one: the slot bytes of a CTA index never
    change hands between calls, which is why one acknowledgement word per (root, CTA) is enough."""
    G, threads = 5, 64
    long = vector_nest(9 * G * CH + 17, G, threads)
    for nvec in (1, CH + 1, 2 * G * CH + 5):
        short = vector_nest(nvec, G, threads)
        assert all(long[v] == b for v, b in short.items())
