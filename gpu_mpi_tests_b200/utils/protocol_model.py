"""Exhaustive interleaving check of the cross-GPU signalling protocols (a small explicit-state
model checker).

The reference never needs this: every exchange is bracketed by device-wide synchronisation
(mpi_stencil2d_gt.cc:202,254), so nothing overlaps.  The fused kernels here do overlap -- a rank
may be a whole launch ahead of its neighbour -- and `compute-sanitizer` cannot see across GPUs
(and is closed on this pool), so the protocol itself is checked: every rank is a little
sequential program of ATOMIC actions on shared words, and `explore()` walks every reachable
interleaving, failing on the first assertion a schedule can break.

Five models, matching csrc/stencil_fused.cu, csrc/collectives.cu and models/stencil2d.py (the last two,
`ll_allgather_model` and `host_feed_model`, are described at their definitions):

* `halo_model(n_ranks, steps, buffers)`: per step e a rank stores its boundary into each
  neighbour's inbox[e % buffers], publishes e in the neighbour's flag (release), waits until its
  own flags are >= e (acquire), reads its inboxes and finishes the step.  With `buffers=2` (what
  the kernels do) every read sees the value of step e; with `buffers=1` the checker finds the
  schedule in which a neighbour that is one launch ahead overwrites ghosts still to be read.
* `allgather_model(n_ranks, calls, entry_round)`: in-place allgather.  A rank may still be
  reading the previous result out of its buffer until it enters the next call, so peers may only
  store into it after it has signalled "entered" (round 0); round 1 signals "my stores are
  done".  Without the entry round the checker finds the overwritten read.

* `ll_allreduce_model(n_ranks, calls, buffers)`: the LL-cell all-reduce, whose cells are matched by
  EQUALITY with the epoch; two cell sets alternating by parity are enough, one is not (deadlock).
  With `gather_between="nowait"` a gather whose senders wait for nobody sits between two
  all-reduces and a fast sender laps a slow rank (deadlock again); `"barrier"` -- what
  gmt_gather_f64 does: it ends with an all-to-all round -- is safe.

Memory ordering is not modelled (actions are sequentially consistent); the mapping of "store
data, then publish" onto `fence.acq_rel.sys` + `st.release.sys` / `ld.acquire.sys` is argued in
DESIGN.md and exercised by the stress tests with injected delays.
"""
from __future__ import annotations

from collections import deque
from typing import Callable


class ProtocolViolation(AssertionError):
    def __init__(self, message: str, trace: list):
        super().__init__(message + "\nschedule: " + " ".join(f"r{r}:{a}" for r, a in trace))
        self.trace = trace


def explore(n_procs: int, init_shared: tuple, programs: list, max_states: int = 2_000_000) -> int:
    """programs[p] is a list of actions; an action is `(name, fn)` with
    fn(shared: list, p) -> None (done) | False (blocked: not enabled in this state); it may raise
    AssertionError.  Explores every interleaving (breadth first, states de-duplicated); returns
    the number of distinct states.  Deadlock (nobody enabled, somebody unfinished) is a violation."""
    start = (tuple([0] * n_procs), init_shared)
    seen = {start: None}
    queue = deque([start])
    while queue:
        state = queue.popleft()
        pcs, shared = state
        progressed = unfinished = False
        for p in range(n_procs):
            if pcs[p] >= len(programs[p]):
                continue
            unfinished = True
            name, fn = programs[p][pcs[p]]
            mem = list(shared)
            try:
                res = fn(mem, p)
            except AssertionError as exc:
                raise ProtocolViolation(str(exc), _trace(seen, state) + [(p, name)]) from None
            if res is False:
                continue
            progressed = True
            nxt = (pcs[:p] + (pcs[p] + 1,) + pcs[p + 1:], tuple(mem))
            if nxt not in seen:
                seen[nxt] = (state, p, name)
                if len(seen) > max_states:
                    raise RuntimeError("state space larger than max_states")
                queue.append(nxt)
        if unfinished and not progressed:
            raise ProtocolViolation("deadlock", _trace(seen, state))
    return len(seen)


def _trace(seen: dict, state) -> list:
    out = []
    while seen[state] is not None:
        state, p, name = seen[state]
        out.append((p, name))
    return out[::-1]


# ---------------------------------------------------------------------------- halo exchange
def halo_model(n_ranks: int = 2, steps: int = 3, buffers: int = 2):
    """Line of ranks, non-periodic (rank 0 has no low neighbour, the last no high neighbour).
    Shared words, per rank r and side s (0 low, 1 high): inbox[r][s][b] and flag[r][s]."""

    def idx_inbox(r, s, b):
        return (r * 2 + s) * buffers + b

    n_inbox = n_ranks * 2 * buffers

    def idx_flag(r, s):
        return n_inbox + r * 2 + s

    init = tuple([None] * n_inbox + [0] * (n_ranks * 2))
    programs = []
    for r in range(n_ranks):
        nbrs = [(0, r - 1)] if r > 0 else []
        if r < n_ranks - 1:
            nbrs.append((1, r + 1))
        prog = []
        for e in range(1, steps + 1):
            for s, nb in nbrs:   # my side s faces nb; it lands in nb's inbox of the opposite side
                def push(mem, p, nb=nb, s=s, e=e, r=r):
                    mem[idx_inbox(nb, 1 - s, e % buffers)] = (r, e)
                prog.append((f"push{e}->{nb}", push))

                def publish(mem, p, nb=nb, s=s, e=e):
                    assert mem[idx_flag(nb, 1 - s)] < e, "epochs must only grow"
                    mem[idx_flag(nb, 1 - s)] = e
                prog.append((f"flag{e}->{nb}", publish))
            for s, nb in nbrs:
                def wait(mem, p, s=s, e=e, r=r):
                    return None if mem[idx_flag(r, s)] >= e else False
                prog.append((f"wait{e}", wait))

                def read(mem, p, s=s, e=e, r=r, nb=nb):
                    got = mem[idx_inbox(r, s, e % buffers)]
                    assert got == (nb, e), f"rank {r} step {e}: ghosts hold {got}, expected {(nb, e)}"
                prog.append((f"read{e}", read))
        programs.append(prog)
    return n_ranks, init, programs


# ---------------------------------------------------------------------------- in-place allgather
def allgather_model(n_ranks: int = 2, calls: int = 2, entry_round: bool = True):
    """buf[r][slot] words; per call c every rank: (the USER fills its own slot -- a store the
    protocol does not control --) enters, stores its slot into every peer's buffer, signals
    completion, waits for everyone's completion, and then the USER reads the whole buffer."""
    W = n_ranks

    def idx_buf(r, slot):
        return r * W + slot

    def idx_sig(rnd, r, src):   # flag block of rank r: word [round][src]
        return W * W + (rnd * W + r) * W + src

    init = tuple([None] * (W * W) + [0] * (2 * W * W))
    programs = []
    for r in range(W):
        prog = []
        for c in range(1, calls + 1):
            def fill(mem, p, r=r, c=c):
                mem[idx_buf(r, r)] = (r, c)
            prog.append((f"fill{c}", fill))
            if entry_round:
                for q in range(W):
                    if q != r:
                        def entered(mem, p, q=q, r=r, c=c):
                            mem[idx_sig(0, q, r)] = c
                        prog.append((f"enter{c}->{q}", entered))
            for q in range(W):
                if q == r:
                    continue
                if entry_round:
                    def wait_entered(mem, p, q=q, r=r, c=c):
                        return None if mem[idx_sig(0, r, q)] >= c else False
                    prog.append((f"waitenter{c}<-{q}", wait_entered))

                def store(mem, p, q=q, r=r, c=c):
                    mem[idx_buf(q, r)] = (r, c)
                prog.append((f"store{c}->{q}", store))

                def done(mem, p, q=q, r=r, c=c):
                    mem[idx_sig(1, q, r)] = c
                prog.append((f"done{c}->{q}", done))
            for q in range(W):
                if q != r:
                    def wait_done(mem, p, q=q, r=r, c=c):
                        return None if mem[idx_sig(1, r, q)] >= c else False
                    prog.append((f"waitdone{c}<-{q}", wait_done))
            for slot in range(W):   # the user consumes the gathered array, one word at a time
                def use(mem, p, slot=slot, r=r, c=c):
                    got = mem[idx_buf(r, slot)]
                    assert got == (slot, c), f"rank {r} call {c}: slot {slot} holds {got}, expected {(slot, c)}"
                prog.append((f"use{c}[{slot}]", use))
        programs.append(prog)
    return W, init, programs


# ---------------------------------------------------------------------------- LL-cell all-reduce
def ll_allreduce_model(n_ranks: int = 2, calls: int = 3, buffers: int = 2, gather_between: str | None = None):
    """Every rank stores its value, tagged with the call number, into cell[c % buffers][its rank] of
    EVERY rank (one atomic store: data and epoch travel together), then takes the W cells of its own
    area one by one, each as soon as its tag EQUALS the call number (not >=: a newer tag would mean
    the value of this call is gone, and the wait would never end -- the checker reports that as a
    deadlock)."""
    W = n_ranks
    n_cells = W * buffers * W

    def idx(owner, b, src):
        return (owner * buffers + b) * W + src

    def idx_sig(r, src):   # round-1 flag words of rank r
        return n_cells + r * W + src

    # gather_between: every even call is a gather to rank 0 instead of an all-reduce (the epoch, hence
    # the parity of the NEXT all-reduce, still advances).  "nowait": senders signal the root and leave
    # (what an MPI_Gather lets them do); "barrier": everybody signals everybody and waits for all.
    init = tuple([None] * n_cells + [0] * (W * W))
    programs = []
    for r in range(W):
        prog = []
        for c in range(1, calls + 1):
            if gather_between and c % 2 == 0:
                targets = range(W) if gather_between == "barrier" else [0]
                for q in targets:
                    def sig(mem, p, q=q, r=r, c=c):
                        mem[idx_sig(q, r)] = c
                    prog.append((f"gsig{c}->{q}", sig))
                if gather_between == "barrier" or r == 0:
                    for q in range(W):
                        def gwait(mem, p, q=q, r=r, c=c):
                            return None if mem[idx_sig(r, q)] >= c else False
                        prog.append((f"gwait{c}<-{q}", gwait))
                continue
            for k in range(W):
                def push(mem, p, q=(r + k) % W, r=r, c=c):
                    mem[idx(q, c % buffers, r)] = (r, c)
                prog.append((f"push{c}->{(r + k) % W}", push))
            for src in range(W):
                def take_cell(mem, p, src=src, r=r, c=c):
                    got = mem[idx(r, c % buffers, src)]
                    if got is None or got[1] < c:
                        return False          # not there yet
                    if got[1] > c:
                        return False          # overwritten by a later call: this wait can never end
                    assert got == (src, c)
                prog.append((f"take{c}[{src}]", take_cell))
        programs.append(prog)
    return W, init, programs


# ---------------------------------------------------------------------------- LL-cell allgather (in place)
def ll_allgather_model(n_ranks: int = 2, calls: int = 3, buffers: int = 2):
    """In-place allgather of small slices through LL cells (gather_ll_kernel): per call the USER fills the
    rank's own slot of its buffer; the rank stores that value, tagged with the call number, into
    cell[c % buffers][its rank] of every PEER's staging area, takes the peers' cells from its own area
    (tag must EQUAL the call number) and copies each into its OWN user buffer; then the user reads the
    whole buffer.  Nobody ever writes another rank's user buffer, so there is no entry round -- the
    checker confirms that the user's reads never see anything but the current call's values, and that
    one cell set (buffers=1) deadlocks like the all-reduce."""
    W = n_ranks
    n_cells = W * buffers * W

    def idx_cell(owner, b, src):
        return (owner * buffers + b) * W + src

    def idx_buf(r, slot):
        return n_cells + r * W + slot

    init = tuple([None] * (n_cells + W * W))
    programs = []
    for r in range(W):
        prog = []
        for c in range(1, calls + 1):
            def fill(mem, p, r=r, c=c):
                mem[idx_buf(r, r)] = (r, c)
            prog.append((f"fill{c}", fill))
            for k in range(1, W):
                def push(mem, p, q=(r + k) % W, r=r, c=c):
                    mem[idx_cell(q, c % buffers, r)] = (r, c)
                prog.append((f"push{c}->{(r + k) % W}", push))
            for k in range(1, W):
                def take(mem, p, src=(r + k) % W, r=r, c=c):
                    got = mem[idx_cell(r, c % buffers, src)]
                    if got is None or got[1] != c:
                        return False          # not there yet -- or gone for good (reported as deadlock)
                    mem[idx_buf(r, src)] = got
                prog.append((f"take{c}[{(r + k) % W}]", take))
            for slot in range(W):
                def use(mem, p, slot=slot, r=r, c=c):
                    got = mem[idx_buf(r, slot)]
                    assert got == (slot, c), f"rank {r} call {c}: slot {slot} holds {got}, expected {(slot, c)}"
                prog.append((f"use{c}[{slot}]", use))
        programs.append(prog)
    return W, init, programs


# ---------------------------------------------------------------------------- host-fed sides, pipelined
def host_feed_model(steps: int = 4, buffers: int = 2, lookahead: int = 1):
    """The host as one more neighbour (Stencil2DWorkload.step_e2e_pipelined): process 0 is the host
    with its copy engine, process 1 the stream of fused kernels.  Per step e the host uploads the
    boundary values into inbox[e % buffers] and then publishes e in the flag word; the kernel of
    step e waits for flag >= e, reads the inbox (two separate reads, so that an upload can fall
    between them), and completes.  The host enqueues step e only after it has seen step
    e - 1 - lookahead complete (lookahead=1: the pipelined call, which waits for the PREVIOUS step's
    result after enqueuing the current one; lookahead=0: the synchronous call).  Two inboxes by
    epoch parity are safe with lookahead 1; with a single inbox the checker finds the upload of step
    e+1 landing between the two reads of step e."""
    INBOX, FLAG, DONE = 0, buffers, buffers + 1
    init = tuple([None] * buffers + [0, 0])
    host, kern = [], []
    for e in range(1, steps + 1):
        def may_enqueue(mem, p, e=e):
            return None if mem[DONE] >= e - 1 - lookahead else False
        host.append((f"enqueue{e}", may_enqueue))

        def upload(mem, p, e=e):
            mem[INBOX + e % buffers] = e
        host.append((f"upload{e}", upload))

        def publish(mem, p, e=e):
            mem[FLAG] = e
        host.append((f"flag{e}", publish))

        def wait(mem, p, e=e):
            return None if mem[FLAG] >= e else False
        kern.append((f"wait{e}", wait))
        for half in (0, 1):
            def read(mem, p, e=e, half=half):
                got = mem[INBOX + e % buffers]
                assert got == e, f"kernel of step {e} (read {half}) sees the boundary values of step {got}"
            kern.append((f"read{e}.{half}", read))

        def done(mem, p, e=e):
            mem[DONE] = e
        kern.append((f"done{e}", done))
    return 2, init, [host, kern]


def check(model: Callable, **kw) -> int:
    n, init, programs = model(**kw)
    return explore(n, init, programs)
