"""Executable model of the shared-memory ring protocols of the tuned kernels (test infrastructure, CPU only).

The kernels stream planes through a TMA ring and pass per-plane results between warps through a second ring; what
keeps them correct is a handful of mbarrier / CTA-barrier rules (kernels_fused.cu smooth2u_tma, kernels_transfer.cu
resid_restrict_tma).  This module restates those rules as warp coroutines over abstract buffers and runs them under
random schedules: every read checks that the buffer holds the plane it expects (a refill or an overwrite that comes too
early shows up as a stale/destroyed read, a wait on the wrong parity as a deadlock).  It models the DESIGN, by hand —
it does not parse the CUDA source — so each protocol function cites the code it mirrors.
"""
import random


class Hazard(AssertionError):
    pass


class MBar:
    """mbarrier: `count` arrivals complete a phase; wait(parity) passes once the phase with that parity has completed"""

    def __init__(self, count):
        self.count, self.pending, self.phase = count, count, 0

    def arrive(self):
        self.pending -= 1
        if self.pending == 0:
            self.pending = self.count
            self.phase += 1

    def passed(self, parity):
        return (self.phase & 1) != (parity & 1)


class Sim:
    def __init__(self, nstage, nslot, nwarps, seed):
        self.rnd = random.Random(seed)
        self.stage = [None] * nstage          # plane held by a ring stage; ("loading", p) while a TMA is in flight
        self.full = [MBar(1) for _ in range(nstage)]
        self.slot = [[None] * nwarps for _ in range(nslot)]   # per warp: which plane's result it last stored there
        self.nwarps = nwarps
        self.inflight = []                    # (stage, plane, remaining delay)
        self.cta_waiting, self.cta_gen = set(), 0

    # ---- TMA ----
    def issue(self, stage, plane, arm=True):
        if isinstance(self.stage[stage], tuple):
            raise Hazard(f"stage {stage} re-armed while plane {self.stage[stage][1]} is still loading")
        self.stage[stage] = ("loading", plane)   # the old contents are gone from here on
        self.inflight.append([stage, plane, self.rnd.randint(0, 6)])

    def tick_tma(self):
        for it in list(self.inflight):
            it[2] -= 1
            if it[2] <= 0:
                self.inflight.remove(it)
                self.stage[it[0]] = it[1]
                self.full[it[0]].arrive()

    # ---- checked accesses ----
    def read_stage(self, stage, plane, who):
        if self.stage[stage] != plane:
            raise Hazard(f"{who}: stage {stage} holds {self.stage[stage]}, expected plane {plane}")

    def write_slot(self, slot, plane, warp):
        self.slot[slot][warp] = plane

    def read_slot(self, slot, plane, who):
        bad = [w for w in range(self.nwarps) if self.slot[slot][w] != plane]
        if bad:
            raise Hazard(f"{who}: slot {slot} expected plane {plane}, warps {bad} hold {[self.slot[slot][w] for w in bad]}")

    # ---- scheduler ----
    def run(self, warps, max_steps=2_000_000):
        """warps: generators yielding None (a scheduling point) or a callable condition to block on"""
        state = [(g, None) for g in warps]
        alive = list(range(len(state)))
        steps = 0
        while alive:
            steps += 1
            if steps > max_steps:
                raise Hazard("no progress (livelock)")
            self.tick_tma()
            runnable = [i for i in alive if state[i][1] is None or state[i][1]()]
            if not runnable:
                if self.inflight:
                    continue
                raise Hazard("deadlock: every warp is blocked and no load is in flight")
            i = self.rnd.choice(runnable)
            g = state[i][0]
            try:
                cond = next(g)
                state[i] = (g, cond)
            except StopIteration:
                alive.remove(i)

    def cta_barrier(self, warp):
        """bar.sync over all warps: returns a condition that holds once everybody has arrived at this generation"""
        gen = self.cta_gen
        self.cta_waiting.add(warp)
        if len(self.cta_waiting) == self.nwarps:
            self.cta_waiting.clear()
            self.cta_gen += 1
        return lambda: self.cta_gen > gen


def _wait(bar, parity):
    return lambda: bar.passed(parity)


# ---------------------------------------------------------------------------------
# smooth2u_tma (kernels_fused.cu): six-stage ring, two first-sweep slots, split-phase mbarrier, thread-0 issue
# ---------------------------------------------------------------------------------
def smooth2u(nzo, nwarps=4, seed=0, vz=False, ns=6, plain_arrive=True):
    sim = Sim(ns, 2, nwarps, seed)
    s1bar = MBar(nwarps)
    last, top = nzo + 2, nzo + 3          # first sweep on planes 1..last, v planes 0..top

    def issue(P):
        has_f = 1 <= P <= last
        if vz and not has_f:
            if plain_arrive:
                sim.full[P % ns].arrive() # plain arrival keeps the stage's parity in step (no load)
            return
        sim.issue(P % ns, P)

    def warp(w):
        if w == 0:
            for P in range(min(ns, top + 1)):
                issue(P)
        yield sim.cta_barrier(w)          # __syncthreads after the barrier init / first issues
        if not vz:
            yield _wait(sim.full[0], 0)
            sim.read_stage(0, 0, f"warp {w} prologue")
            yield _wait(sim.full[1], 0)
            sim.read_stage(1, 1, f"warp {w} prologue")
        for q in range(1, last + 1):
            par = (q // ns) & 1
            if vz:
                yield _wait(sim.full[q % ns], par)
            else:
                su = (q + 1) % ns
                yield _wait(sim.full[su], par ^ 1 if su == 0 else par)
                sim.read_stage(su, q + 1, f"warp {w} step {q} (U)")
            sim.read_stage(q % ns, q, f"warp {w} step {q}")
            yield None
            if q >= 2:
                yield _wait(s1bar, q & 1)
                if w == 0:
                    if q == 2 and ns <= top:
                        issue(ns)
                    if q + ns - 1 <= top:
                        issue(q + ns - 1)
            if q >= 3:
                sim.read_slot((q - 1) & 1, q - 1, f"warp {w} second sweep of plane {q - 1}")
                yield None
            sim.write_slot(q & 1, q, w)
            s1bar.arrive()
            yield None

    sim.run([warp(w) for w in range(nwarps)])
    return sim


# ---------------------------------------------------------------------------------
# resid_restrict_tma (kernels_transfer.cu): thread-0 issue after the per-coarse-plane barrier
# ---------------------------------------------------------------------------------
def resid_restrict(nzc, nwarps=4, seed=0, ns=6, nslot=5, refill_ahead=0):
    sim = Sim(ns, nslot, nwarps, seed)
    nvp = 2 * nzc + 3
    nxt = [0]

    def refill(j):
        while nxt[0] < nvp and nxt[0] <= 2 * j + 3 + ns + refill_ahead:
            sim.issue(nxt[0] % ns, nxt[0])
            nxt[0] += 1

    def warp(w):
        if w == 0:
            while nxt[0] < ns and nxt[0] < nvp:
                sim.issue(nxt[0] % ns, nxt[0])
                nxt[0] += 1
        yield sim.cta_barrier(w)
        yield _wait(sim.full[0], 0)
        sim.read_stage(0, 0, f"warp {w} prologue")
        yield _wait(sim.full[1 % ns], (1 // ns) & 1)
        sim.read_stage(1 % ns, 1, f"warp {w} prologue")

        def do_plane(t):
            q = t + 1
            yield _wait(sim.full[(q + 1) % ns], ((q + 1) // ns) & 1)
            sim.read_stage(q % ns, q, f"warp {w} residual plane {t}")
            sim.read_stage((q + 1) % ns, q + 1, f"warp {w} residual plane {t} (U)")
            yield None
            sim.write_slot(t % nslot, t, w)
            yield None

        def restrict(j):
            for t in (2 * j, 2 * j + 1, 2 * j + 2):
                sim.read_slot(t % nslot, t, f"warp {w} coarse plane {j}")
            yield None

        yield from do_plane(0)
        for j in range(nzc):
            yield from do_plane(2 * j + 1)
            yield from do_plane(2 * j + 2)
            yield sim.cta_barrier(w)
            if w == 0:
                refill(j)
            yield from restrict(j)

    sim.run([warp(w) for w in range(nwarps)])
    return sim


# ---------------------------------------------------------------------------------
# the retired producer-warp ring (full/empty mbarriers), as the single-sweep stencil_tma still uses it
# ---------------------------------------------------------------------------------
def producer_ring(nz, nwarps=4, seed=0, ns=4):
    sim = Sim(ns, 1, nwarps, seed)
    empty = [MBar(nwarps) for _ in range(ns)]
    nplanes = nz + 2

    def producer():
        for q in range(nplanes):
            if q >= ns:
                yield _wait(empty[q % ns], ((q // ns) - 1) & 1)
            sim.issue(q % ns, q)
            yield None

    def warp(w):
        yield _wait(sim.full[0], 0)
        sim.read_stage(0, 0, f"warp {w} prologue")
        yield _wait(sim.full[1 % ns], (1 // ns) & 1)
        sim.read_stage(1 % ns, 1, f"warp {w} prologue")
        for q in range(1, nz + 1):
            yield _wait(sim.full[(q + 1) % ns], ((q + 1) // ns) & 1)
            sim.read_stage(q % ns, q, f"warp {w} plane {q}")
            sim.read_stage((q + 1) % ns, q + 1, f"warp {w} plane {q} (U)")
            yield None
            if q == 1:
                empty[0].arrive()
            empty[q % ns].arrive()
            yield None

    sim.run([producer()] + [warp(w) for w in range(nwarps)])
    return sim


# ---------------------------------------------------------------------------------
# design study for the next round: prolongation+correction folded into the two-sweep post-smoother.
# Plane P lands uncorrected; at step q every warp adds the interpolated correction to ITS points of plane q+2 in place,
# before the arrival that ends the step.  The plane is read as U (own points) at step q+1 and as E/W/S/N (other warps'
# points) at step q+2 — behind the wait of step q+1, which proves every warp has finished step q.  No new barrier.
# ---------------------------------------------------------------------------------
def smooth2_prolong(nzo, nwarps=4, seed=0, ns=6, correct_ahead=2):
    sim = Sim(ns, 2, nwarps, seed)
    s1bar = MBar(nwarps)
    last, top = nzo + 2, nzo + 3
    corrected = [[None] * nwarps for _ in range(ns)]   # per stage and warp: plane whose correction that warp has stored

    def correct(w, P):
        if P > top:
            return
        sim.read_stage(P % ns, P, f"warp {w} correction of plane {P}")
        corrected[P % ns][w] = P

    def read_corrected(w, P, own, who):
        sim.read_stage(P % ns, P, who)
        ws = [w] if own else range(nwarps)
        bad = [x for x in ws if corrected[P % ns][x] != P]
        if bad:
            raise Hazard(f"{who}: plane {P} read before warps {bad} corrected it")

    def warp(w):
        if w == 0:
            for P in range(min(ns, top + 1)):
                sim.issue(P % ns, P)
        yield sim.cta_barrier(w)
        # prologue: planes 0 .. correct_ahead are corrected behind a CTA barrier of their own
        for P in range(0, correct_ahead + 1):
            yield _wait(sim.full[P % ns], (P // ns) & 1)
            correct(w, P)
        yield sim.cta_barrier(w)
        read_corrected(w, 0, True, f"warp {w} prologue")
        read_corrected(w, 1, True, f"warp {w} prologue")
        for q in range(1, last + 1):
            P = q + correct_ahead
            if P <= top and P > correct_ahead:
                yield _wait(sim.full[P % ns], (P // ns) & 1)
                correct(w, P)
                yield None
            read_corrected(w, q + 1, True, f"warp {w} step {q} (U)")
            read_corrected(w, q, False, f"warp {w} step {q} (E/W/S/N)")
            yield None
            if q >= 2:
                yield _wait(s1bar, q & 1)
                if w == 0:
                    if q == 2 and ns <= top:
                        sim.issue(0, ns)
                    if q + ns - 1 <= top:
                        sim.issue((q + ns - 1) % ns, q + ns - 1)
            if q >= 3:
                sim.read_slot((q - 1) & 1, q - 1, f"warp {w} second sweep of plane {q - 1}")
                yield None
            sim.write_slot(q & 1, q, w)
            s1bar.arrive()
            yield None

    sim.run([warp(w) for w in range(nwarps)])
    return sim
