"""CPU: an executable model of the cross-rank flag protocol of csrc/gradavg.cu and csrc/gradavg_view.cu.

Every (rank, block) of every launch is a coroutine that performs the same steps, in the same order,
as the CUDA kernels (waits on `done` / `ready` / `mid` words, bucket / staging / gather accesses,
epoch and `last_kind` bookkeeping by the last block).  A random scheduler interleaves them -- ranks
run ahead of each other by whole calls, like GPUs with launch skew -- and every load checks that the
cell it reads holds exactly the value the algorithm expects at that point (source rank, call number,
tile).  A stale or clobbered cell, a read before the write, or all coroutines blocked (deadlock)
fails the test.  It models the LOGIC of the protocol (which wait protects which buffer, epoch and
parity arithmetic, variant mixing), not the memory model: accesses are sequentially consistent.
compute-sanitizer is closed on this pool, so this plus the stress tests on real GPUs stand in for
a race detector.
"""
import random

import pytest

ONESHOT, TWOSHOT, ONESHOT_PUSH, TWOSHOT_PUSH, TWOSHOT_PIPE = 1, 2, 3, 4, 5
VIEW = 7            # ga_view_kernel (csrc/gradavg_view.cu): gradients live in the peer-mapped region, averaged in place
FLAG_ROWS = 16      # the view kernel indexes the flag rows by chunk, not by block


class Rank:
    def __init__(self, world, blocks):
        self.ready = [[0] * world for _ in range(max(blocks, FLAG_ROWS))]
        self.mid = [[0] * world for _ in range(max(blocks, FLAG_ROWS))]
        self.done = [0] * world
        self.epoch = 0
        self.last_kind = 0
        self.finished = 0
        self.send = {}          # bucket (pull) / staging (push): key -> value
        self.gather = {}
        self.grads = {}

    def store(self, layout, key, value):
        """A store into the bucket region.  The pull bucket, the two-shot staging slots and the two
        one-shot halves are different LAYOUTS of the same memory: a store in one layout destroys
        whatever another layout kept there (conservatively: all of it), except that the two one-shot
        halves do not overlap each other."""
        for k in [k for k in self.send if k[0] != layout and not (k[0] in ("h0", "h1") and layout in ("h0", "h1"))]:
            del self.send[k]
        self.send[(layout,) + key] = value


class World:
    def __init__(self, world, blocks, seed):
        self.W, self.B = world, blocks
        self.ranks = [Rank(world, blocks) for _ in range(world)]
        self.rng = random.Random(seed)

    # ---- one block of one launch, as a coroutine ------------------------------------------------
    def block(self, r, b, call, variant, T, launch):
        W, B, me = self.W, self.B, self.ranks[r]
        epoch, rounds, group, dbuf = launch["epoch"], launch["rounds"], launch["group"], launch["dbuf"]
        Ts = -(-T // W)

        def wait(words, target):
            for s in range(W):
                while words[s] - target < 0:
                    yield "blocked"
            yield "ok"

        def raise_flags(kind, value):
            for j in range(W):
                getattr(self.ranks[j], kind)[b][r] = value
            yield "ok"

        def expect(cell, want, what):
            assert cell == want, "%s: call %d rank %d block %d read %r, expected %r" % (what, call, r, b, cell, want)

        fold = ("out", call)
        mine = lambda t: ("in", call, r, t)

        if variant in (ONESHOT, TWOSHOT):                                   # ga_fused_kernel
            yield from wait(me.done, epoch - 1)
            tiles = [t for t in range(T) if t % B == b] if variant == ONESHOT else \
                [j * Ts + i for i in range(Ts) if i % B == b for j in range(W) if j * Ts + i < T]
            for t in tiles:                                                 # (1) pack
                expect(me.grads[t], mine(t), "pack")
                me.store("pull", (t,), mine(t))
                yield "ok"
            yield from raise_flags("ready", epoch)
            yield from wait(me.ready[b], epoch)
            if variant == ONESHOT:
                for t in tiles:                                             # (3) reduce everything
                    for s in range(W):
                        expect(self.ranks[s].send.get(("pull", t)), ("in", call, s, t), "one-shot pull")
                        yield "ok"
                    me.grads[t] = fold + (t,)
            else:
                own = [r * Ts + i for i in range(Ts) if i % B == b and r * Ts + i < T]
                for t in own:                                               # (3) reduce-scatter, in place
                    for s in range(W):
                        expect(self.ranks[s].send.get(("pull", t)), ("in", call, s, t), "two-shot pull reduce")
                        yield "ok"
                    me.store("pull", (t,), fold + (t,))
                    me.grads[t] = fold + (t,)
                yield from raise_flags("mid", epoch)
                yield from wait(me.mid[b], epoch)
                for i in range(Ts):                                         # (5) all-gather
                    if i % B != b:
                        continue
                    for jj in range(1, W):
                        j = (r + jj) % W
                        t = j * Ts + i
                        if t >= T:
                            continue
                        expect(self.ranks[j].send.get(("pull", t)), fold + (t,), "two-shot pull gather")
                        me.grads[t] = fold + (t,)
                        yield "ok"
            last_kind, last_epoch = 0, epoch

        elif variant == ONESHOT_PUSH:                                       # ga_push1_kernel
            half = (epoch & 1) if dbuf else 0
            if not (dbuf and me.last_kind == 1):
                yield from wait(me.done, epoch - 1)
            tiles = [t for t in range(b * group, min((b + 1) * group, T))]
            for t in tiles:
                expect(me.grads[t], mine(t), "scatter source")
                for jj in range(W):
                    j = (r + jj) % W
                    self.ranks[j].store("h%d" % half, (r, t), mine(t))
                    yield "ok"
            yield from raise_flags("ready", epoch)
            yield from wait(me.ready[b], epoch)
            for t in tiles:
                for s in range(W):
                    expect(me.send.get(("h%d" % half, s, t)), ("in", call, s, t), "one-shot push fold")
                    yield "ok"
                me.grads[t] = fold + (t,)
            last_kind, last_epoch = (1 if dbuf else 0), epoch

        else:                                                               # ga_push_kernel / ga_push_pipe_kernel
            yield from wait(me.done, epoch - 1)
            if variant == TWOSHOT_PUSH:
                rounds, slots_of = 1, (lambda q: [i for i in range(Ts) if i % B == b])
            else:
                slots_of = lambda q: [i for i in range((q * B + b) * group, (q * B + b + 1) * group) if i < Ts]

            def scatter(q):
                for i in slots_of(q):
                    for jj in range(W):
                        j = (r + jj) % W
                        t = j * Ts + i
                        if t >= T:
                            continue
                        expect(me.grads[t], mine(t), "scatter source")
                        self.ranks[j].store("stage", (r, i), mine(t))
                        yield "ok"

            def fold_round(q):
                for i in slots_of(q):
                    t = r * Ts + i
                    if t >= T:
                        continue
                    for s in range(W):
                        expect(me.send.get(("stage", s, i)), ("in", call, s, t), "two-shot push fold")
                        yield "ok"
                    me.grads[t] = fold + (t,)
                    for jj in range(1, W):
                        self.ranks[(r + jj) % W].gather[t] = fold + (t,)
                        yield "ok"

            def unpack(q):
                for i in slots_of(q):
                    for jj in range(1, W):
                        t = ((r + jj) % W) * Ts + i
                        if t >= T:
                            continue
                        expect(me.gather.get(t), fold + (t,), "unpack")
                        me.grads[t] = fold + (t,)
                        yield "ok"

            yield from scatter(0)
            yield from raise_flags("ready", epoch)
            for q in range(rounds):
                if q + 1 < rounds:
                    yield from scatter(q + 1)
                    yield from raise_flags("ready", epoch + q + 1)
                yield from wait(me.ready[b], epoch + q)
                yield from fold_round(q)
                yield from raise_flags("mid", epoch + q)
                if q > 0:
                    yield from wait(me.mid[b], epoch + q - 1)
                    yield from unpack(q - 1)
            yield from wait(me.mid[b], epoch + rounds - 1)
            yield from unpack(rounds - 1)
            last_kind, last_epoch = 0, epoch + rounds - 1

        # (6) last block of this rank: advance the epoch, tell everyone this rank is done
        me.finished += 1
        if me.finished == launch["blocks"]:
            me.finished = 0
            me.epoch = last_epoch
            me.last_kind = last_kind
            for j in range(W):
                self.ranks[j].done[r] = last_epoch
        yield "ok"


    # ---- ga_view_kernel: one coroutine per ROLE-WARP GROUP of a block -------------------------------
    # A scatter block is its data warps (stores, then a tick in shared memory) and its publisher (waits
    # for the tick, raises ready[c][r] everywhere); a fold block is its poller (waits for ready[c] from
    # every rank), its data warps (fold, store the result into EVERY rank's region) and its publisher
    # (mid[c][r]); the poller ends the block by waiting for mid[c] from every rank.  The data warps may
    # run ahead of the publisher, as in the kernel.  `grads` is the view region itself (in place).
    def view_parts(self, r, k, call, T, launch):
        W, me = self.W, self.ranks[r]
        epoch, C, nb, one = launch["epoch"], launch["chunk"], launch["nb"], launch["oneshot"]
        Ts = T if one else -(-T // W)
        NC = -(-Ts // C)
        chunks = list(range(k, NC, nb))
        fold = ("out", call)
        sh = {"issued_s": 0, "issued_f": 0, "ready": 0}          # the block's shared-memory words

        def expect(cell, want, what):
            assert cell == want, "%s: call %d rank %d block-pair %d read %r, expected %r" % (what, call, r, k, cell, want)

        def epilogue():
            me.finished += 1
            if me.finished == launch["blocks"]:
                me.finished = 0
                me.epoch = epoch
                me.last_kind = 0
                for j in range(W):
                    self.ranks[j].done[r] = epoch
            yield "ok"

        def slot(s, j):                      # one-shot: the slot rank j keeps for source s
            return s if s < j else s - 1

        def scatter_data():
            for s_ in range(W):
                while me.done[s_] - (epoch - 1) < 0:
                    yield "blocked"
            for q, c in enumerate(chunks):
                for i in range(c * C, min((c + 1) * C, Ts)):
                    for jj in range(1, W):
                        j = (r + jj) % W
                        t = i if one else j * Ts + i
                        if t >= T:
                            continue
                        expect(me.grads[t], ("in", call, r, t), "view scatter source")
                        self.ranks[j].store("stage", ((slot(r, j), t) if one else (r, i)), ("in", call, r, t))
                        yield "ok"
                sh["issued_s"] = q + 1
                yield "ok"
            yield from epilogue()

        def scatter_publisher():
            for q, c in enumerate(chunks):
                while sh["issued_s"] < q + 1:
                    yield "blocked"
                for j in range(W):
                    self.ranks[j].ready[c][r] = epoch
                yield "ok"
            yield from epilogue()

        def fold_poller():
            for q, c in enumerate(chunks):
                for s_ in range(W):
                    while me.ready[c][s_] - epoch < 0:
                        yield "blocked"
                sh["ready"] = q + 1
                yield "ok"
            if not one:
                for c in chunks:
                    for s_ in range(W):
                        while me.mid[c][s_] - epoch < 0:
                            yield "blocked"
            yield from epilogue()

        def fold_data():
            for q, c in enumerate(chunks):
                while sh["ready"] < q + 1:
                    yield "blocked"
                for i in range(c * C, min((c + 1) * C, Ts)):
                    t = i if one else r * Ts + i
                    if t >= T:
                        continue
                    for s_ in range(W):
                        if s_ == r:
                            expect(me.grads[t], ("in", call, r, t), "view fold, own tile")
                        else:
                            key = ("stage", slot(s_, r), t) if one else ("stage", s_, i)
                            expect(me.send.get(key), ("in", call, s_, t), "view fold, staged tile")
                        yield "ok"
                    for jj in range(W if not one else 1):
                        self.ranks[(r + jj) % W].grads[t] = fold + (t,)
                        yield "ok"
                sh["issued_f"] = q + 1
                yield "ok"
            yield from epilogue()

        def fold_publisher():
            for q, c in enumerate(chunks):
                while sh["issued_f"] < q + 1:
                    yield "blocked"
                if not one:
                    for j in range(W):
                        self.ranks[j].mid[c][r] = epoch
                yield "ok"
            yield from epilogue()

        return [scatter_data(), scatter_publisher(), fold_poller(), fold_data(), fold_publisher()]

    # ---- scheduler ----------------------------------------------------------------------------------
    def run(self, calls):
        """calls: list of (variant, tiles).  Each rank launches them in order; launch k+1 of a rank
        starts when all blocks of its launch k have finished (stream order)."""
        W, B = self.W, self.B
        next_call = [0] * W
        active = {r: [] for r in range(W)}
        steps = 0
        while True:
            for r in range(W):
                if not active[r] and next_call[r] < len(calls):
                    call = next_call[r]
                    variant, T = calls[call]
                    me = self.ranks[r]
                    for t in range(T):                                  # backward produced fresh gradients
                        me.grads[t] = ("in", call, r, t)
                    Ts = -(-T // W)
                    group = 2
                    if variant == TWOSHOT_PIPE:
                        blocks = min(B, -(-Ts // group))
                        rounds = -(-Ts // (blocks * group))
                    elif variant == ONESHOT_PUSH:
                        group = max(2, -(-T // B))
                        blocks, rounds = -(-T // group), 1
                    else:
                        blocks, rounds = B, 1
                    launch = {"epoch": me.epoch + 1, "rounds": rounds, "group": group, "blocks": blocks,
                              "dbuf": variant == ONESHOT_PUSH and T <= 6}      # "fits two halves" for small sets
                    if variant == VIEW:
                        one = T <= 2 and W > 1                                 # small sets take the one-shot form
                        chunk = 1 if T <= 6 else 2
                        NC = -(-(T if one else Ts) // chunk)
                        nb = min(B, NC)
                        launch.update(chunk=chunk, nb=nb, oneshot=one, blocks=5 * nb)
                        active[r] = [g for k in range(nb) for g in self.view_parts(r, k, call, T, launch)]
                    else:
                        active[r] = [self.block(r, b, call, variant, T, launch) for b in range(blocks)]
                    active[r] = [(g, call, T) for g in active[r]]
                    next_call[r] += 1
            runnable = [(r, k) for r in range(W) for k in range(len(active[r]))]
            if not runnable:
                break
            progressed = False
            self.rng.shuffle(runnable)
            # bias: occasionally let one rank run far ahead (launch skew)
            if self.rng.random() < 0.3:
                fav = self.rng.randrange(W)
                runnable.sort(key=lambda rk: rk[0] != fav)
            for r, k in runnable:
                g, call, T = active[r][k]
                try:
                    res = next(g)
                except StopIteration:
                    active[r][k] = None
                    progressed = True
                    break
                if res != "blocked":
                    progressed = True
                    break
            for r in range(W):
                if any(x is None for x in active[r]):
                    done_call, T = None, None
                    active[r] = [x for x in active[r] if x is not None]
                    if not active[r]:
                        call = next_call[r] - 1
                        T = calls[call][1]
                        for t in range(T):
                            assert self.ranks[r].grads[t] == ("out", call, t), \
                                "rank %d finished call %d with tile %d = %r" % (r, call, t, self.ranks[r].grads[t])
            steps += 1
            assert progressed, "deadlock: every block of every rank is waiting (calls %r)" % (calls,)
            assert steps < 2_000_000
        assert all(n == len(calls) for n in next_call)


@pytest.mark.parametrize("world", [2, 3, 4])
@pytest.mark.parametrize("blocks", [1, 3])
def test_random_interleavings_mixed_variants(world, blocks):
    variants = [ONESHOT, TWOSHOT, ONESHOT_PUSH, TWOSHOT_PUSH, TWOSHOT_PIPE, VIEW, VIEW]
    for seed in range(60):
        rng = random.Random(1000 * world + 10 * blocks + seed)
        calls = [(rng.choice(variants), rng.choice([1, 2, 5, 6, 7, 13])) for _ in range(8)]
        World(world, blocks, seed).run(calls)


def test_back_to_back_one_shot_push_shrinking_sets():
    """The case the fixed mid-bucket offset of the odd staging half exists for: consecutive
    double-buffered one-shot launches of different sizes, with no `done` wait between them."""
    for seed in range(20):
        World(3, 2, seed).run([(ONESHOT_PUSH, 6), (ONESHOT_PUSH, 2), (ONESHOT_PUSH, 5), (ONESHOT_PUSH, 1),
                               (TWOSHOT_PUSH, 6), (ONESHOT_PUSH, 3), (ONESHOT, 4), (ONESHOT_PUSH, 6)])


def test_model_catches_a_missing_wait():
    """Sanity of the checker itself: drop the `done` wait of the pull kernels (the protection against
    packing over a bucket a slow peer still reads) and some interleaving must be flagged."""
    class Broken(World):
        def block(self, r, b, call, variant, T, launch):
            self.ranks[r].done = [1 << 30] * self.W          # pretend everyone is always done
            return super().block(r, b, call, variant, T, launch)
    caught = 0
    for seed in range(40):
        try:
            Broken(3, 1, seed).run([(ONESHOT, 4)] * 6)
        except AssertionError:
            caught += 1
    assert caught > 0


def test_model_catches_single_buffer_without_wait():
    """One staging buffer and no `done` wait between one-shot launches is the hazard the two halves
    remove: the model must see a peer's fold reading the NEXT call's tiles."""
    class Broken(World):
        def block(self, r, b, call, variant, T, launch):
            launch = dict(launch, dbuf=False)
            self.ranks[r].done = [1 << 30] * self.W          # skip the wait although there is one buffer
            return super().block(r, b, call, variant, T, launch)
    caught = 0
    for seed in range(40):
        try:
            Broken(3, 1, seed).run([(ONESHOT_PUSH, 4)] * 6)
        except AssertionError:
            caught += 1
    assert caught > 0


def test_view_kernel_alone_and_back_to_back():
    """ga_view_kernel's chunk flags, role split and one-shot form, launch after launch with skew."""
    for world in (2, 3, 4):
        for blocks in (1, 2, 3):
            for seed in range(25):
                rng = random.Random(7 * world + blocks + 100 * seed)
                World(world, blocks, seed).run([(VIEW, rng.choice([1, 2, 3, 5, 7, 13, 16])) for _ in range(6)])


def test_model_catches_a_fold_that_does_not_wait_for_its_own_scatter():
    """One-shot form: the fold overwrites a tile in place, so it must not run before this rank's own scatter
    block has sent that tile -- the poller waits for all W ready words, its own included.  Drop that and the
    model must see a scatter reading an already averaged tile."""
    class Broken(World):
        def view_parts(self, r, k, call, T, launch):
            parts = super().view_parts(r, k, call, T, launch)
            me = self.ranks[r]
            for c in range(FLAG_ROWS):
                me.ready[c][r] = 1 << 30              # my own ready words look raised from the start
            return parts
    caught = 0
    for seed in range(60):
        try:
            Broken(2, 1, seed).run([(VIEW, 2)] * 4)
        except AssertionError:
            caught += 1
    assert caught > 0
