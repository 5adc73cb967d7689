"""Model check of the synchronisation protocol of csrc/peer.cu (CPU only; the kernel itself is tested on two GPUs by
tests/test_gpu_peer.py and exercised at 2 / 4 / 8 ranks by bench.py's known-answer `Exchange.verify`).

What is modelled, exactly as the kernel does it: W ranks, a fixed grid of G CTAs per launch, CTA b of a rank talks only
to CTA b of the other ranks; per CTA: epoch = epochs[b] + 1, publish its chunks into staging half (epoch & 1), raise
flags[peer][rank][b] = epoch, wait until flags[self][peer][b] - epoch >= 0 for every peer, read every rank's half,
epochs[b] = epoch.  Launches of one rank run in stream order (launch k + 1 starts after every CTA of launch k ended);
ranks drift freely.  A random scheduler interleaves single steps; every staged word carries (launch, writer) so a read
that sees anything but the current launch's data -- a stale half, or one a fast rank already overwrote -- is caught.
The control experiment (one staging half) must fail: that is what shows the check has teeth."""
import random

import pytest

G = 4                    # CTAs per launch (the kernel: 128); small so that schedules collide often
CHUNKS_PER_HALF = 8


def make_launches(rng, n_launches):
    """Each launch: the list of staging chunks it uses (identical on every rank, like the size-driven cuts of
    dls_peer_sgd_step).  Chunk c is owned by CTA c % G -- a fixed function of the staging offset."""
    return [sorted(rng.sample(range(CHUNKS_PER_HALF), rng.randint(1, CHUNKS_PER_HALF))) for _ in range(n_launches)]


def cta(rank, b, k, chunks, world, mem, halves):
    """One CTA of launch k as a generator; every `yield` is a point where another CTA (of any rank) may run."""
    mine = [c for c in chunks if c % G == b]
    if not mine:
        return
    epoch = mem["epochs"][rank][b] + 1
    half = epoch % halves
    for c in mine:                                           # phase 1: publish
        mem["stage"][rank][half][c] = (k, rank)
        yield
    for peer in range(world):                                # release: the flag follows the data
        if peer != rank:
            mem["flags"][peer][rank][b] = epoch
            yield
    for peer in range(world):                                # phase 3: acquire
        if peer != rank:
            while mem["flags"][rank][peer][b] - epoch < 0:
                yield "blocked"
    for c in mine:                                           # phase 4: read every rank's copy, in rank order
        for r in range(world):
            got = mem["stage"][r][half][c]
            assert got == (k, r), "rank %d CTA %d launch %d chunk %d: read %r from rank %d" % (rank, b, k, c, got, r)
            yield
    mem["epochs"][rank][b] = epoch


def run(world, launches, seed, halves=2):
    rng = random.Random(seed)
    mem = {"epochs": [[0] * G for _ in range(world)],
           "flags": [[[0] * G for _ in range(world)] for _ in range(world)],
           "stage": [[[None] * CHUNKS_PER_HALF for _ in range(halves)] for _ in range(world)]}
    position = [0] * world                                   # next launch of each rank
    live = [[] for _ in range(world)]                        # running CTAs of each rank's current launch
    steps = 0
    while True:
        for r in range(world):                               # stream order: the next launch starts when the previous one is done
            if not live[r] and position[r] < len(launches):
                k = position[r]
                position[r] += 1
                live[r] = [cta(r, b, k, launches[k], world, mem, halves) for b in range(G)]
        runnable = [(r, g) for r in range(world) for g in live[r]]
        if not runnable:
            return steps
        progressed = False
        rng.shuffle(runnable)
        # a rank may be arbitrarily slow: mostly advance one CTA, sometimes starve whole ranks for a while
        slow = set(rng.sample(range(world), rng.randint(0, world - 1))) if rng.random() < 0.3 else set()
        for r, g in runnable:
            if r in slow and rng.random() < 0.9:
                continue
            try:
                if next(g) != "blocked":
                    progressed = True
            except StopIteration:
                live[r].remove(g)
                progressed = True
            steps += 1
            if rng.random() < 0.5:
                break
        if not progressed and not slow:
            # everything runnable is blocked: only a deadlock if nobody can make progress in a full fair round
            if all(_blocked(g) for _, g in runnable):
                raise AssertionError("deadlock: every CTA is waiting for a flag")


def _blocked(g):
    try:
        return next(g) == "blocked"
    except StopIteration:
        return False


@pytest.mark.parametrize("world", [2, 4, 8])
def test_double_buffered_protocol_never_reads_stale_data_and_never_deadlocks(world):
    for seed in range(60):
        rng = random.Random(1000 + seed)
        launches = make_launches(rng, rng.randint(2, 9))
        run(world, launches, seed)


def test_a_single_staging_half_is_caught_by_the_model():
    """Control: with one half a fast rank overwrites data a slow peer has not read yet; some schedule must expose it."""
    caught = 0
    for seed in range(200):
        rng = random.Random(2000 + seed)
        launches = make_launches(rng, 6)
        try:
            run(4, launches, seed, halves=1)
        except AssertionError as exc:
            assert "read" in str(exc)
            caught += 1
    assert caught > 0
