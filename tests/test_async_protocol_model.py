"""Executable model of the async arrival-ring protocol of csrc/ps_kernels.cu (push_async_kernel admission +
async_scan_kernel drain), explored under random interleavings.

The CUDA kernels are tested on hardware in test_ps_kernels.py; this model checks the PROTOCOL itself -- the part
that failed once in a way no single-GPU test showed (ranks blocked in a push could not serve their own ring: 601 ms
per step at 8 GPUs) -- for the properties the design relies on:
  * wait-free admission: a producer never blocks; it either takes a ticket or skips this owner for this step,
  * the ring never holds more than R tickets, so a slot is never overwritten before the owner consumed it, even when
    all W producers pass the admission check at the same instant (hence the margin of W slots),
  * arrivals are applied in ticket order and only when all producer CTAs of that ticket have signalled,
  * a skipped push keeps its gradient: the amount applied plus the amount still held equals the amount produced.
"""
from hypothesis import given, settings, strategies as st


class Owner:
    def __init__(self, R):
        self.R, self.tail, self.head = R, 0, 0
        self.done = [0] * R            # per-slot arrival counter (monotonic, G arrivals per use)
        self.hdr = [None] * R          # (src, amount) written by the producer's first CTA
        self.applied = []              # (ticket, src, amount) in application order


def run(W, R, G, steps, schedule, margin=None):
    """schedule: iterator of ints choosing which actor moves next.  Actors 0..W-1 are producers, W is the owner."""
    margin = W if margin is None else margin
    owner = Owner(R)
    held = [0] * W                     # gradient mass a producer still holds (kept when a push is skipped)
    produced = [0] * W
    pc = [("idle", None)] * W          # program counter of each producer
    step_of = [0] * W
    max_occ = 0
    overwritten = []

    def producer_move(r):
        state, data = pc[r]
        if state == "idle":
            if step_of[r] >= steps:
                return False
            step_of[r] += 1
            produced[r] += 1
            held[r] += 1                                        # backward pass adds this step's gradient
            pc[r] = ("check", (owner.tail, owner.head))         # ld.acquire tail, head (may be stale when used)
            return True
        if state == "check":
            tail, head = data
            if tail - head > R - margin:                        # no certain room: skip, keep the gradient
                pc[r] = ("idle", None)
            else:
                ticket = owner.tail                             # atom.add
                owner.tail += 1
                slot = ticket % R
                if ticket - owner.head >= R:
                    overwritten.append(ticket)                  # would clobber an unconsumed slot
                owner.hdr[slot] = (r, held[r], ticket)
                held[r] = 0                                     # slice is zeroed as it is sent
                pc[r] = ("signal", (slot, G))
            return True
        if state == "signal":                                   # the G CTAs signal one by one
            slot, left = data
            owner.done[slot] += 1
            pc[r] = ("signal", (slot, left - 1)) if left > 1 else ("idle", None)
            return True
        return False

    def owner_move():
        slot = owner.head % R
        need = (owner.head // R + 1) * G
        if owner.done[slot] < need:
            return False
        src, amount, ticket = owner.hdr[slot]
        owner.applied.append((ticket, src, amount))
        owner.head += 1
        return True

    idle_rounds = 0
    for choice in schedule:
        a = choice % (W + 1)
        moved = owner_move() if a == W else producer_move(a)
        max_occ = max(max_occ, owner.tail - owner.head)
        idle_rounds = 0 if moved else idle_rounds + 1
        if idle_rounds > 4 * (W + 1):
            break
    # drain: let everybody finish (no actor ever waits for another, so round-robin terminates)
    for _ in range(10 * (steps * W * (G + 2) + R)):
        progressed = any([producer_move(r) for r in range(W)] + [owner_move()])
        max_occ = max(max_occ, owner.tail - owner.head)
        if not progressed:
            break
    return owner, held, produced, max_occ, overwritten, pc


@settings(max_examples=150, deadline=None)
@given(W=st.integers(1, 8), extra=st.integers(0, 8), G=st.integers(1, 3), steps=st.integers(1, 6),
       schedule=st.lists(st.integers(0, 63), min_size=0, max_size=400))
def test_wait_free_ring_invariants(W, extra, G, steps, schedule):
    R = W + extra                                   # the admission margin needs R >= W
    owner, held, produced, max_occ, overwritten, pc = run(W, R, G, steps, schedule)
    assert not overwritten and max_occ <= R
    assert all(state == "idle" for state, _ in pc)                      # nobody is stuck: wait-free producers
    assert owner.head == owner.tail                                     # everything that got a ticket was applied
    assert [t for t, _, _ in owner.applied] == list(range(owner.tail))  # in ticket order
    applied_by = [0] * W
    for _, src, amount in owner.applied:
        applied_by[src] += amount
    assert [applied_by[r] + held[r] for r in range(W)] == produced       # skipped pushes keep their gradient


def test_all_producers_pass_the_check_at_the_same_instant():
    """Worst case for the margin: every producer reads (tail, head) before any of them takes a ticket."""
    W, R, G = 8, 8, 2
    sched = list(range(W)) + list(range(W)) + list(range(W))            # idle->check, then all take tickets
    owner, held, produced, max_occ, overwritten, _ = run(W, R, G, 1, sched)
    assert max_occ == R and not overwritten


def test_full_ring_skips_instead_of_blocking():
    W, R, G = 2, 3, 1
    # the owner never runs during the schedule: after R - W + 1 admitted tickets further pushes are skipped
    sched = [0, 0, 0, 1, 1, 1] * 4
    owner, held, produced, max_occ, overwritten, pc = run(W, R, G, 4, sched)
    assert not overwritten and max_occ <= R
    assert sum(amount for _, _, amount in owner.applied) + sum(held) == sum(produced)


def test_a_margin_smaller_than_the_number_of_producers_is_not_enough():
    """Why the check is `tail - head > R - W` and not `> R - 1`: two producers that read the counters before either
    takes its ticket both pass a margin-1 check when one slot is left, and the second overwrites a live slot."""
    W, R, G = 8, 8, 1
    sched = []
    for r in range(7):
        sched += [r, r, r]                        # seven producers: read, take ticket, signal
    sched += [7, 0, 7, 0]                         # producer 7 and producer 0 (next step) read, then both take tickets
    _, _, _, max_occ, overwritten, _ = run(W, R, G, 2, sched, margin=1)
    assert max_occ == R + 1 and overwritten == [8]
    _, _, _, max_occ, overwritten, _ = run(W, 2 * R, G, 2, sched)          # the real rule, with room for two rounds
    assert max_occ <= 2 * R and not overwritten
