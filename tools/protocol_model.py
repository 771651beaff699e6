"""Randomised model checking of the hand-written mbarrier protocols (no GPU needed).

compute-sanitizer (racecheck / synccheck) is closed on this pool (profiles/r2_sanitizer.md), so the synchronisation
protocols of the four kernels that wait on mbarriers are restated here as small concurrent programs -- one generator
per warp role, same barriers, same counts, same parities, same order of operations as the CUDA source -- and run under
many random schedules.  Asynchronous agents (bulk copies, TMA loads, the tensor core's in-order MMA queue, DSMEM copies)
complete at random later times.  The simulator fails on
  * deadlock: nobody can run and somebody has not finished (a hang on the GPU),
  * a read that sees the wrong generation of a buffer (phase aliasing, a missing wait),
  * a write into a buffer that is still being read or whose previous contents were never consumed (a missing "empty"),
  * a work item processed twice or never, a counter not back at zero.

mbarrier semantics modelled (PTX ISA, "mbarrier"): a phase completes when the pending-arrival count AND the
transaction count are both zero; expect_tx raises the transaction count, complete_tx lowers it (it may go negative
when the data lands before the expectation is posted); try_wait.parity(P) is true iff the phase with parity P has
completed, i.e. the current phase's parity differs from P -- so a waiter that is two phases behind WOULD pass a wait it
should not, which is exactly the aliasing this tool looks for.

    python tools/protocol_model.py [--seeds N]       # every model, N random schedules each

Models: attn_fused (csrc/attn_fused.cu: ticket counter, producer / E / C warps, tile ring of 3, item ring of 4),
enc_exchange (csrc/kernels.cu enc_lstm_persistent_kernel: per-step DSMEM exchange between the CTAs of a cluster, barrier
re-armed by one thread), vocab_pair (csrc/vocab_tc.cu: two CTAs, two TMA producers, one MMA issuer, tcgen05.commit
multicast, double-buffered accumulator), dec_cell (csrc/dec_cell.cu: weight ring + two one-shot cluster exchanges).
"""
from __future__ import annotations

import argparse
import random
import sys


class ModelError(AssertionError):
    pass


class MBar:
    def __init__(self, name, count):
        self.name, self.count, self.pending, self.tx, self.phase = name, count, count, 0, 0

    def _check(self):
        if self.pending < 0:
            raise ModelError("%s: more arrivals than its count in one phase" % self.name)
        if self.pending == 0 and self.tx == 0:
            self.phase += 1
            self.pending = self.count

    def arrive(self):
        self.pending -= 1
        self._check()

    def expect_arrive(self, units):
        self.tx += units
        self.pending -= 1
        self._check()

    def complete_tx(self, units):
        self.tx -= units
        self._check()

    def passed(self, parity):
        return (self.phase & 1) != parity


class Buf:
    """a piece of shared memory with a generation tag; reads and writes are intervals"""

    def __init__(self, name, tag=None, track=True):
        self.name, self.tag, self.readers, self.writers, self.consumed, self.track = name, tag, 0, 0, True, track

    def begin_write(self, need_consumed=True):
        if self.readers:
            raise ModelError("%s: written while %d reader(s) are inside it" % (self.name, self.readers))
        if need_consumed and self.track and not self.consumed:
            raise ModelError("%s: overwritten before its contents (%r) were consumed" % (self.name, self.tag))
        self.writers += 1
        self.tag = None

    def end_write(self, tag, last=True):
        self.writers -= 1
        if last:
            self.tag, self.consumed = tag, False

    def begin_read(self, want):
        if self.writers or self.tag != want:
            raise ModelError("%s: read expecting %r, holds %r (%d writer(s) active)" % (self.name, want, self.tag, self.writers))
        self.readers += 1

    def end_read(self, want, consume=False):
        if self.writers or self.tag != want:
            raise ModelError("%s: changed under a reader (%r -> %r)" % (self.name, want, self.tag))
        self.readers -= 1
        if consume:
            self.consumed = True


class Sim:
    def __init__(self, seed):
        self.rng = random.Random(seed)
        # two kinds of schedule: uniformly random choices (even seeds), and random PRIORITIES with a few random change
        # points (odd seeds: an agent, or the asynchronous completions, can be starved for long stretches -- the
        # schedules that expose ordering bugs which uniform choices almost never produce)
        self.by_priority = bool(seed & 1)
        self.prio = {}
        self.agents = []          # [name, generator, blocked-on]
        self.inflight = []        # asynchronous completions (callables)
        self.named = {}           # named barrier id -> [arrived agents]
        self.steps = 0

    def spawn(self, name, gen):
        self.agents.append([name, gen, None])

    def later(self, fn):
        self.inflight.append(fn)

    def run(self, max_steps=2_000_000):
        while self.agents:
            runnable = []
            for a in self.agents:
                b = a[2]
                if b is None or (b[0] == "wait" and b[1].passed(b[2])) or (b[0] == "released"):
                    runnable.append(a)
            n = len(runnable) + len(self.inflight)
            if n == 0:
                raise ModelError("DEADLOCK: " + "; ".join("%s blocked on %s" % (a[0], self._what(a[2])) for a in self.agents))
            if self.by_priority:
                if self.prio and self.rng.random() < 0.02:         # a change point: reshuffle one priority
                    self.prio[self.rng.choice(sorted(self.prio))] = self.rng.random()
                best, k = -1.0, 0
                for i, a in enumerate(runnable):
                    pr = self.prio.setdefault(a[0], self.rng.random())
                    if pr > best:
                        best, k = pr, i
                if self.inflight:
                    pr = self.prio.setdefault("<async>", self.rng.random())
                    if pr > best or not runnable:
                        k = len(runnable) + self.rng.randrange(len(self.inflight))
            else:
                k = self.rng.randrange(n)
            if k >= len(runnable):
                self.inflight.pop(k - len(runnable))()
            else:
                a = runnable[k]
                a[2] = None
                try:
                    req = next(a[1])
                except StopIteration:
                    self.agents.remove(a)
                    continue
                if req is None:
                    pass
                elif req[0] == "wait":
                    a[2] = req
                elif req[0] == "barsync":                      # bar.sync id, n
                    lst = self.named.setdefault(req[1], [])
                    lst.append(a)
                    a[2] = ("barsync", req[1])
                    if len(lst) == req[2]:
                        for x in lst:
                            x[2] = ("released",)
                        self.named[req[1]] = []
            self.steps += 1
            if self.steps > max_steps:
                raise ModelError("step limit: livelock?")
        if self.inflight:
            while self.inflight:
                self.inflight.pop()()

    @staticmethod
    def _what(b):
        if b is None:
            return "nothing"
        if b[0] == "wait":
            return "%s parity %d (phase %d, pending %d, tx %d)" % (b[1].name, b[2], b[1].phase, b[1].pending, b[1].tx)
        return str(b[:2])


def wait(bar, parity):
    return ("wait", bar, parity)


# ------------------------------------------------------------------------------------------------
# attn_fused_kernel (csrc/attn_fused.cu)
# ------------------------------------------------------------------------------------------------
def model_attn_fused(seed, n_cta=3, n_art=4, nseg_max=3, nds=2, n_e=3, n_c=2, mutate=None):
    """n_cta persistent CTAs share the ticket counter; every CTA has a producer, n_e E agents (agent 0 = tid 0) and
    n_c C agents (agent 0 = ctid 0).  nds = 2 is the beam-4 configuration, 1 the beam-8 one."""
    sim = Sim(seed)
    rng = random.Random(seed ^ 0x5eed)
    nsg = [rng.randrange(0, nseg_max + 1) for _ in range(n_art)]          # segments of every article (0: finished)
    n_items = n_art * nseg_max
    ticket = [0]
    done_e, done_c = {}, {}
    IQ = 2 if mutate == "item_ring_of_2" else 4

    def cta(c):
        full = [MBar("cta%d.full%d" % (c, i), 1) for i in range(3)]
        empty = [MBar("cta%d.empty%d" % (c, i), 1) for i in range(3)]
        itembar = [MBar("cta%d.item%d" % (c, i), 1) for i in range(IQ)]
        dsfull = [MBar("cta%d.dsfull%d" % (c, i), 1) for i in range(nds)]
        dsfree = [MBar("cta%d.dsfree%d" % (c, i), 1) for i in range(nds)]
        efull, efree = MBar("cta%d.efull" % c, 1), MBar("cta%d.efree" % c, 1)
        items = [None] * IQ
        tiles = [Buf("cta%d.tile%d" % (c, i)) for i in range(3)]
        ds = [Buf("cta%d.ds%d" % (c, i)) for i in range(nds)]
        part = [Buf("cta%d.part.slice%d" % (c, e)) for e in range(n_e)]
        psm = Buf("cta%d.psm" % c)

        def draw():
            while True:
                t = ticket[0]
                ticket[0] += 1                                            # atomicAdd
                if t >= n_items:
                    if t == n_items + n_cta - 1 and mutate != "no_ticket_reset":
                        ticket[0] = 0
                    return -1
                art = t // nseg_max
                if t - art * nseg_max < nsg[art]:
                    return t

        def issue_copy(buf, tag, bar, pieces):
            buf.begin_write()
            for _ in range(pieces - 1):
                buf.writers += 1
            bar.expect_arrive(pieces)
            left = [pieces]

            def piece():
                left[0] -= 1
                buf.end_write(tag, last=left[0] == 0)
                bar.complete_tx(1)
            for _ in range(pieces):
                sim.later(piece)

        def producer():
            cur = draw()
            items[0] = cur
            itembar[0].arrive()
            yield
            nl = j = 0
            while cur >= 0:
                nxt = 0
                for states in (0, 1):
                    yield wait(empty[nl % 3], ((nl // 3) & 1) ^ 1)
                    issue_copy(tiles[nl % 3], ("H" if states else "F", cur), full[nl % 3], 1 + rng.randrange(3))
                    yield
                    if states == 0:
                        b = j % nds
                        if mutate != "no_dsfree_wait":
                            yield wait(dsfree[b], ((j // nds) & 1) ^ 1)
                        issue_copy(ds[b], cur, dsfull[b], 2)
                        yield
                        nxt = draw()
                        items[(j + 1) % IQ] = nxt
                        itembar[(j + 1) % IQ].arrive()
                        yield
                    nl += 1
                cur = nxt
                j += 1

        def e_agent(e):
            j = 0
            while True:
                yield wait(itembar[j % IQ], (j // IQ) & 1)
                it = items[j % IQ]
                if it < 0:
                    return
                db, b = j % nds, (2 * j) % 3
                yield wait(dsfull[db], (j // nds) & 1)
                yield wait(full[b], ((2 * j) // 3) & 1)
                ds[db].begin_read(it)
                tiles[b].begin_read(("F", it))
                yield                                                      # the E loop
                tiles[b].end_read(("F", it), consume=True)
                ds[db].end_read(it, consume=True)
                if j > 0 and mutate != "no_efree_wait":
                    yield wait(efree, (j - 1) & 1)
                part[e].begin_write()                                      # this warp's slice of the partial energies
                part[e].end_write(it)
                yield ("barsync", "cta%d.E" % c, n_e)
                if e == 0:
                    empty[b].arrive()
                    dsfree[db].arrive()
                    efull.arrive()
                    done_e[it] = done_e.get(it, 0) + 1
                yield
                j += 1

        def c_agent(k):
            j = 0
            while True:
                yield wait(itembar[j % IQ], (j // IQ) & 1)
                it = items[j % IQ]
                if it < 0:
                    return
                yield wait(efull, j & 1)
                for p in part:
                    p.begin_read(it)
                yield
                for p in part:
                    p.end_read(it, consume=(k == n_c - 1))
                if k == 0:
                    psm.begin_write(need_consumed=False)
                    psm.end_write(it)
                yield ("barsync", "cta%d.C" % c, n_c)
                if k == 0:
                    efree.arrive()
                b = (2 * j + 1) % 3
                yield wait(full[b], ((2 * j + 1) // 3) & 1)
                tiles[b].begin_read(("H", it))
                psm.begin_read(it)
                yield                                                      # the context loop
                psm.end_read(it)
                tiles[b].end_read(("H", it), consume=True)
                yield ("barsync", "cta%d.C" % c, n_c)
                if k == 0:
                    empty[b].arrive()
                    done_c[it] = done_c.get(it, 0) + 1
                yield
                j += 1

        sim.spawn("cta%d.producer" % c, producer())
        for e in range(n_e):
            sim.spawn("cta%d.E%d" % (c, e), e_agent(e))
        for k in range(n_c):
            sim.spawn("cta%d.C%d" % (c, k), c_agent(k))

    for c in range(n_cta):
        cta(c)
    sim.run()
    valid = {a * nseg_max + s for a in range(n_art) for s in range(nsg[a])}
    if set(done_e) != valid or set(done_c) != valid or any(v != 1 for v in done_e.values()) or any(v != 1 for v in done_c.values()):
        raise ModelError("items processed %r / %r, expected each of %r once" % (done_e, done_c, sorted(valid)))
    if ticket[0] != 0:
        raise ModelError("ticket counter left at %d" % ticket[0])
    return sim.steps


# ------------------------------------------------------------------------------------------------
# enc_lstm_persistent_kernel: the per-step exchange of h between the CTAs of a cluster (csrc/kernels.cu)
# ------------------------------------------------------------------------------------------------
def model_enc_exchange(seed, n_cta=4, grp_max=6, n_gate=2, n_other=2, mutate=None):
    """Per CTA: agent 0 = tid 0 (waits, re-arms the barrier, gate math, sends), n_gate - 1 more gate/sender agents, n_other
    threads that only do the mat-vec.  Every CTA sends its block of h(s) to every CTA (itself included) with one copy
    per destination that signals the destination's hbar[buffer]."""
    sim = Sim(seed)
    H = n_cta                                                     # one unit of bytes per source CTA
    hbar = [[MBar("cta%d.hbar%d" % (c, b), 1) for b in range(2)] for c in range(n_cta)]
    # hs[c][buf][src] = step whose h that block holds
    hs = [[[Buf("cta%d.hs%d.from%d" % (c, b, s), tag=-1) for s in range(n_cta)] for b in range(2)] for c in range(n_cta)]
    hstage = [[Buf("cta%d.hstage%d" % (c, b), tag=None) for b in range(2)] for c in range(n_cta)]
    n_thr = n_gate + n_other
    red = [[Buf("cta%d.red.t%d" % (c, t), track=False) for t in range(n_thr)] for c in range(n_cta)]
    for c in range(n_cta):
        for b in range(2):
            hbar[c][b].expect_arrive(H)                          # set-up: both barriers armed once (tid 0), then cluster.sync
            for s in range(n_cta):
                hs[c][b][s].consumed = True

    def thread(c, t):
        gate = t < n_gate
        for s in range(grp_max):
            cur = s & 1
            if s > 0:
                if not (mutate == "one_thread_skips_the_wait" and t == n_thr - 1):
                    yield wait(hbar[c][cur], ((s - 1) >> 1) & 1)
                if t == 0 and s + 1 < grp_max and mutate != "no_rearm":
                    hbar[c][cur].expect_arrive(H)
                    yield
            want = s - 1 if s > 0 else -1
            for src in range(n_cta):
                hs[c][cur][src].begin_read(want)
            yield                                                 # the mat-vec over hs[cur]
            for src in range(n_cta):
                hs[c][cur][src].end_read(want, consume=(t == n_thr - 1))
            # partial sums of step s overwrite those of step s-1
            red[c][t].begin_write()
            red[c][t].end_write(s)
            yield ("barsync", "cta%d.sync" % c, n_thr)           # __syncthreads
            if gate:
                for r in red[c]:
                    r.begin_read(s)
                yield
                for r in red[c]:
                    r.end_read(s)
                if s + 1 < grp_max and t == 0:
                    st = hstage[c][0 if mutate == "one_staging_buffer" else cur ^ 1]
                    st.begin_write(need_consumed=True)
                    st.end_write(s)
            if s + 1 < grp_max and gate:
                yield ("barsync", "cta%d.gates" % c, n_gate)     # bar.sync 1 among the gate warps
                if t == 0:                                        # (tid < 8: one copy per destination CTA)
                    left = [n_cta]
                    for dst in range(n_cta):
                        blk = hs[dst][cur ^ 1][c]
                        src_buf = hstage[c][0 if mutate == "one_staging_buffer" else cur ^ 1]

                        def land(blk=blk, bar=hbar[dst][cur ^ 1], s=s, src_buf=src_buf, left=left):
                            if src_buf.tag != s:
                                raise ModelError("%s rewritten (%r) before the copy of step %d read it" % (src_buf.name, src_buf.tag, s))
                            blk.begin_write()
                            blk.end_write(s)
                            left[0] -= 1
                            if left[0] == 0:
                                src_buf.consumed = True
                            bar.complete_tx(1)
                        sim.later(land)
                yield
        # final cluster.sync is a hardware barrier: nothing to model beyond "everybody gets here"

    for c in range(n_cta):
        for t in range(n_thr):
            sim.spawn("cta%d.t%d" % (c, t), thread(c, t))
    sim.run()
    return sim.steps


# ------------------------------------------------------------------------------------------------
# vocab_tc_kernel: one CTA pair (csrc/vocab_tc.cu)
# ------------------------------------------------------------------------------------------------
def model_vocab_pair(seed, n_tiles=5, KB=3, STAGES=4, n_epi=2, mutate=None):
    """CTA 0 = leader.  Barriers: full[s], a_full[kb], t_empty[a] live in the leader; empty[s], t_full[a] in each CTA
    (tcgen05.commit multicasts to both).  The tensor core executes the leader's instruction queue in order, at random
    times; a commit's arrivals happen when everything queued before it has executed."""
    sim = Sim(seed)
    full = [MBar("full%d" % s, 1) for s in range(STAGES)]
    a_full = [MBar("a_full%d" % k, 1) for k in range(KB)]
    t_empty = [MBar("t_empty%d" % a, 2 * n_epi) for a in range(2)]
    empty = [[MBar("cta%d.empty%d" % (c, s), 1) for s in range(STAGES)] for c in range(2)]
    t_full = [[MBar("cta%d.t_full%d" % (c, a), 1) for a in range(2)] for c in range(2)]
    stage = [[Buf("cta%d.stage%d" % (c, s)) for s in range(STAGES)] for c in range(2)]
    ablk = [[Buf("cta%d.A%d" % (c, k)) for k in range(KB)] for c in range(2)]
    acc = [[Buf("cta%d.acc%d" % (c, a)) for a in range(2)] for c in range(2)]
    queue = []                                                    # the tensor core's in-order queue
    tiles = list(range(n_tiles))
    drained = {}

    def tensor_core_step():
        op = queue.pop(0)
        if op[0] == "mma":
            _, s, kb, tile, a, first = op
            for c in range(2):
                stage[c][s].begin_read((tile, kb))
                ablk[c][kb].begin_read(kb)
                stage[c][s].end_read((tile, kb), consume=True)
                ablk[c][kb].end_read(kb)
                if first:
                    acc[c][a].begin_write()                       # accumulate = 0: overwrites -> must have been drained
                    acc[c][a].end_write((tile, 0))
                else:
                    if acc[c][a].tag != (tile, kb - 1) or acc[c][a].readers:
                        raise ModelError("accumulator %s holds %r at k-block %d of tile %d" % (acc[c][a].name, acc[c][a].tag, kb, tile))
                    acc[c][a].tag = (tile, kb)
        else:
            for b in op[1]:
                b.arrive()
        if queue:
            sim.later(tensor_core_step)

    def push(op):
        queue.append(op)
        if len(queue) == 1:
            sim.later(tensor_core_step)

    def producer(c):
        it = 0
        for tile in tiles:
            for kb in range(KB):
                s = it % STAGES
                if it < KB:
                    if c == 0:
                        a_full[kb].expect_arrive(2)
                    ablk[c][kb].begin_write()

                    def landA(kb=kb):
                        ablk[c][kb].end_write(kb)
                        a_full[kb].complete_tx(1)
                    sim.later(landA)
                    yield
                if not (mutate == "follower_skips_empty_wait" and c == 1):
                    yield wait(empty[c][s], ((it // STAGES) & 1) ^ 1)
                if c == 0:
                    full[s].expect_arrive(2)
                stage[c][s].begin_write()

                def landB(s=s, tile=tile, kb=kb):
                    stage[c][s].end_write((tile, kb))
                    full[s].complete_tx(1)
                sim.later(landB)
                yield
                it += 1

    def mma_issuer():
        it = 0
        for tcount, tile in enumerate(tiles):
            a = tcount & 1
            if mutate != "no_t_empty_wait":
                yield wait(t_empty[a], ((tcount >> 1) & 1) ^ 1)
            for kb in range(KB):
                s = it % STAGES
                if it < KB:
                    yield wait(a_full[kb], 0)
                yield wait(full[s], (it // STAGES) & 1)
                push(("mma", s, kb, tile, a, kb == 0))
                push(("commit", [empty[0][s]] if mutate == "commit_not_multicast" else [empty[0][s], empty[1][s]]))
                yield
                it += 1
            push(("commit", [t_full[0][a], t_full[1][a]]))
            yield

    def epilogue(c, w):
        for tcount, tile in enumerate(tiles):
            a = tcount & 1
            yield wait(t_full[c][a], (tcount >> 1) & 1)
            acc[c][a].begin_read((tile, KB - 1))
            yield
            acc[c][a].end_read((tile, KB - 1))
            drained[(c, a, tile)] = drained.get((c, a, tile), 0) + 1
            if drained[(c, a, tile)] == n_epi:
                acc[c][a].consumed = True
            t_empty[a].arrive()                                   # remote arrive on the leader's barrier
            yield

    for c in range(2):
        sim.spawn("cta%d.producer" % c, producer(c))
        for w in range(n_epi):
            sim.spawn("cta%d.epi%d" % (c, w), epilogue(c, w))
    sim.spawn("cta0.mma", mma_issuer())
    sim.run()
    if len(drained) != 2 * n_tiles or any(v != n_epi for v in drained.values()):
        raise ModelError("accumulators drained %r" % drained)
    return sim.steps


# ------------------------------------------------------------------------------------------------
# dec_cell_kernel: weight ring and the two one-shot exchanges inside a cluster (csrc/dec_cell.cu)
# ------------------------------------------------------------------------------------------------
def model_dec_cell(seed, n_cta=3, n_warp=3, CH=(2, 5, 3), STAGES=4, x_at=3, mutate=None):
    """Per CTA: a producer (inputs, then CH[0] + CH[1] + CH[2] weight chunks through a ring of STAGES) and n_warp compute
    warps.  x of every CTA goes to every CTA after stage 1 (x_bar), (c, h) after stage 2 (ch_bar); stage 2 starts with
    its h part and waits for x at chunk x_at of the stage."""
    sim = Sim(seed)
    n_chunks = sum(CH)
    full = [[MBar("cta%d.full%d" % (c, s), 1) for s in range(STAGES)] for c in range(n_cta)]
    empty = [[MBar("cta%d.empty%d" % (c, s), 1 if mutate == "empty_count_1" else n_warp) for s in range(STAGES)] for c in range(n_cta)]
    in_bar = [MBar("cta%d.in" % c, 1) for c in range(n_cta)]
    x_bar = [MBar("cta%d.x" % c, 1) for c in range(n_cta)]
    ch_bar = [MBar("cta%d.ch" % c, 1) for c in range(n_cta)]
    ring = [[Buf("cta%d.ring%d" % (c, s), track=False) for s in range(STAGES)] for c in range(n_cta)]
    xin = [[Buf("cta%d.x.from%d" % (c, s)) for s in range(n_cta)] for c in range(n_cta)]
    chin = [[Buf("cta%d.ch.from%d" % (c, s)) for s in range(n_cta)] for c in range(n_cta)]
    for c in range(n_cta):                                        # tid 0 arms both exchange barriers before the cluster barrier
        x_bar[c].expect_arrive(n_cta)
        ch_bar[c].expect_arrive(n_cta)

    def producer(c):
        in_bar[c].expect_arrive(1)
        sim.later(lambda: in_bar[c].complete_tx(1))
        yield
        for n in range(n_chunks):
            s = n % STAGES
            yield wait(empty[c][s], ((n // STAGES) & 1) ^ 1)
            full[c][s].expect_arrive(1)
            ring[c][s].begin_write()

            def land(s=s, n=n):
                ring[c][s].end_write(n)
                full[c][s].complete_tx(1)
            sim.later(land)
            yield

    def send(c, bufs, bars, tag):
        for dst in range(n_cta):
            def land(dst=dst):
                bufs[dst][c].begin_write()
                bufs[dst][c].end_write(tag)
                bars[dst].complete_tx(1)
            sim.later(land)

    def warp(c, w):
        n = 0

        def chunk():
            nonlocal n
            s = n % STAGES
            yield wait(full[c][s], (n // STAGES) & 1)
            ring[c][s].begin_read(n)
            yield
            ring[c][s].end_read(n)
            empty[c][s].arrive()
            n += 1
        yield wait(in_bar[c], 0)
        for _ in range(CH[0]):
            yield from chunk()
        yield ("barsync", "cta%d.compute" % c, n_warp)
        yield ("barsync", "cta%d.compute" % c, n_warp)
        if w == 0:
            send(c, xin, x_bar, "x")
        for i in range(CH[1]):
            if i == x_at:
                if mutate != "no_x_wait":
                    yield wait(x_bar[c], 0)
                for s in range(n_cta):
                    xin[c][s].begin_read("x")
                    xin[c][s].end_read("x")
            yield from chunk()
        yield ("barsync", "cta%d.compute" % c, n_warp)
        yield ("barsync", "cta%d.compute" % c, n_warp)
        if w == 0:
            send(c, chin, ch_bar, "ch")
        yield wait(ch_bar[c], 0)
        for s in range(n_cta):
            chin[c][s].begin_read("ch")
            chin[c][s].end_read("ch")
        for _ in range(CH[2]):
            yield from chunk()

    for c in range(n_cta):
        sim.spawn("cta%d.producer" % c, producer(c))
        for w in range(n_warp):
            sim.spawn("cta%d.w%d" % (c, w), warp(c, w))
    sim.run()
    return sim.steps


MODELS = {
    "attn_fused(beam 4)": lambda seed: model_attn_fused(seed, nds=2),
    "attn_fused(beam 8)": lambda seed: model_attn_fused(seed, nds=1),
    "attn_fused(1 CTA, many items)": lambda seed: model_attn_fused(seed, n_cta=1, n_art=5, nseg_max=4),
    "enc_exchange": model_enc_exchange,
    "enc_exchange(short)": lambda seed: model_enc_exchange(seed, grp_max=1 + seed % 4),
    "vocab_pair": model_vocab_pair,
    "vocab_pair(KB=8)": lambda seed: model_vocab_pair(seed, n_tiles=3, KB=8),
    "dec_cell": model_dec_cell,
}


# Deliberately broken variants: every one must be caught (tests/test_protocol_models.py) -- the checker's own check.
MUTANTS = {
    "attn_fused: item ring of 2 instead of 4": lambda seed: model_attn_fused(seed, mutate="item_ring_of_2"),
    "attn_fused (beam 8): producer does not wait for dsfree": lambda seed: model_attn_fused(seed, nds=1, mutate="no_dsfree_wait"),
    "attn_fused: E warps do not wait for efree": lambda seed: model_attn_fused(seed, mutate="no_efree_wait"),
    "attn_fused: ticket counter not reset": lambda seed: model_attn_fused(seed, mutate="no_ticket_reset"),
    "enc_exchange: barrier not re-armed": lambda seed: model_enc_exchange(seed, mutate="no_rearm"),
    "enc_exchange: one staging buffer": lambda seed: model_enc_exchange(seed, mutate="one_staging_buffer"),
    "enc_exchange: one thread skips the wait": lambda seed: model_enc_exchange(seed, mutate="one_thread_skips_the_wait"),
    "vocab_pair: MMA issuer does not wait for t_empty": lambda seed: model_vocab_pair(seed, mutate="no_t_empty_wait"),
    "vocab_pair: follower's producer skips its empty wait": lambda seed: model_vocab_pair(seed, mutate="follower_skips_empty_wait"),
    "vocab_pair: commit to the leader only": lambda seed: model_vocab_pair(seed, mutate="commit_not_multicast"),
    "dec_cell: empty barrier counts one warp": lambda seed: model_dec_cell(seed, mutate="empty_count_1"),
    "dec_cell: no wait for x": lambda seed: model_dec_cell(seed, mutate="no_x_wait"),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", type=int, default=300)
    args = ap.parse_args()
    for name, fn in MODELS.items():
        steps = 0
        for seed in range(args.seeds):
            steps += fn(seed)
        print("%-32s %5d schedules, %9d scheduling steps: no deadlock, no stale read, no early overwrite" % (name, args.seeds, steps))
    for name, fn in MUTANTS.items():
        caught = 0
        for seed in range(min(args.seeds, 100)):
            try:
                fn(seed)
            except ModelError:
                caught += 1
        print("mutant  %-55s caught in %3d of %d schedules" % (name, caught, min(args.seeds, 100)))


if __name__ == "__main__":
    sys.exit(main())
