"""Executable model of the peer-mapped ghost-plane handshake (kernels_comm.cu p2p_push / p2p_wait; test infrastructure,
CPU only).  Ranks are coroutines that run the SAME sequence of operators; between operators that read ghost planes there
is an exchange.  Each exchange e of a pair is: ack(e) to the peer, wait for the peer's ack(e), write the peer's ghost
buffer, signal(e); the consumer waits for signal(e).  The model checks, under random schedules, that
  * a consumer only ever reads ghost data of the exchange it expects (nothing stale, nothing from the future),
  * nobody writes a ghost buffer while its owner is still reading the previous contents,
  * nobody deadlocks;
and that dropping the ack (the tempting "the symmetric exchange orders it anyway" shortcut) is caught when two
consecutive exchanges refresh the SAME buffer (in-place operators: prolongation + correction after the residual +
restriction).  It models the design by hand; each step cites the kernel line it mirrors."""
import random


class Hazard(AssertionError):
    pass


class Rank:
    def __init__(self, r, world):
        self.r, self.world = r, world
        self.epoch = {p: 0 for p in self.peers()}    # exchanges started with peer p        (P2PCtl::epoch)
        self.ack_in = {p: 0 for p in self.peers()}   # written by p                           (P2PCtl::ack_in)
        self.sig_in = {p: 0 for p in self.peers()}   # written by p                           (P2PCtl::sig_in)
        self.ghost = {p: {} for p in self.peers()}   # buffer name -> exchange number of the data it holds (from peer p)
        self.reading = None                          # (buffer, expected) while a consumer kernel runs

    def peers(self):
        return [p for p in (self.r - 1, self.r + 1) if 0 <= p < self.world]


def run(world, schedule, seed, use_ack=True, steps_per_kernel=3):
    """schedule: list of buffer names; operator i first refreshes ghost buffer schedule[i] and then consumes it"""
    rnd = random.Random(seed)
    ranks = [Rank(r, world) for r in range(world)]

    def program(me):
        for buf in schedule:
            # ---- p2p_push ----
            e = {p: me.epoch[p] + 1 for p in me.peers()}
            if use_ack:
                for p in me.peers():                       # step 1: peer.ack_in = e (kernels_comm.cu:63)
                    ranks[p].ack_in[me.r] = e[p]
                yield
                while any(me.ack_in[p] < e[p] for p in me.peers()):   # step 2 (:64)
                    yield "blocked"
            for p in me.peers():                           # step 3: copy into the peer's ghost buffer (:68-94)
                peer = ranks[p]
                if peer.reading is not None and peer.reading[0] == buf:
                    raise Hazard(f"rank {me.r} overwrites ghost buffer {buf!r} of rank {p} while it is being read")
                peer.ghost[me.r][buf] = e[p]
                yield
            for p in me.peers():                           # step 4: sig_in = e at the peer, epoch = e here (:104-108)
                ranks[p].sig_in[me.r] = e[p]
                me.epoch[p] = e[p]
            yield
            # ---- p2p_wait ----
            while any(me.sig_in[p] < me.epoch[p] for p in me.peers()):
                yield "blocked"
            # ---- the consumer kernel: reads the ghost buffer over several scheduling steps ----
            me.reading = (buf, dict(me.epoch))
            for _ in range(steps_per_kernel):
                for p in me.peers():
                    if me.ghost[p].get(buf) != me.epoch[p]:
                        raise Hazard(f"rank {me.r} reads {buf!r} from rank {p}: holds exchange {me.ghost[p].get(buf)}, wants {me.epoch[p]}")
                yield
            me.reading = None
            yield

    progs = [program(r) for r in ranks]
    live = list(range(world))
    idle = 0
    while live:
        i = rnd.choice(live)
        try:
            st = next(progs[i])
        except StopIteration:
            live.remove(i)
            continue
        idle = idle + 1 if st == "blocked" else 0
        if idle > 2000 * world:
            raise Hazard("deadlock")
    return True
