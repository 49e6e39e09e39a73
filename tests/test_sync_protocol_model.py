"""Executable model of one SYNC round as the device path runs it (csrc/ps_kernels.cu: push_sync_kernel SEND / REDUCE,
pull_seg_kernel, wait_seg_kernel; dps_b200/api.py: Worker._pre / _send_group / _exchange; models/engine.py: the side
streams), explored under random interleavings of the ranks.

Nothing in a step waits on the host and no kernel carries a round number in its data: correctness rests on a chain of
happens-before edges between streams and arrival counters.  The kernels are tested on hardware (test_ps_kernels.py,
test_multi_gpu.py); this model checks the CHAIN, for every schedule hypothesis can find:

  * true barrier (reference: src/parameter_server/server.py:264-288 with SURVEY Q1 fixed): the forward AND the backward
    pass of step t on every rank read weights that contain exactly t updates of EVERY shard,
  * the owner's REDUCE of round t reads, for every source and layer group, the gradient of round t -- never a stale
    one, never one already overwritten by round t+1 (the staging slots are single-buffered),
  * a SEND never overwrites a staging slot whose previous content has not been reduced yet,
  * nobody deadlocks: every rank finishes all its steps.

Streams per rank and step t (K layer groups in forward order; the arena is sharded over all W owners):
  pull stream : P(t,0..K-1)   owner role: broadcast group k of the own shard to every worker, then signal SEGCOUNT[k];
                              starts after the main stream of step t-1 has ended (Worker._pre: side.wait_stream(cur))
  main stream : WAIT(t,k) FWD(t,k) for k = 0..K-1;  BWD(t,k) for k = K-1..0;  SEND(t,0);  REDUCE(t)
  wgrad stream: SEND(t,k) for k = K-1..1, each after BWD(t,k)   (engine.on_group_done);  joined before SEND(t,0)
"""
from hypothesis import given, settings, strategies as st
import pytest


class Violation(AssertionError):
    pass


def build_program(rank, K, steps, pull_waits_for_reduce=True):
    """Ops of one rank as (name, step, group, deps) in issue order; deps are indices of ops of the SAME rank that must
    have run before (stream order + cross-stream event edges).  Cross-RANK dependencies are the counters checked when an
    op runs."""
    ops, idx = [], {}

    def add(name, t, k, deps):
        ops.append((name, t, k, [d for d in deps if d is not None]))
        idx[(name, t, k)] = len(ops) - 1
        return len(ops) - 1

    main_prev = None          # last op of the main stream
    for t in range(steps):
        pull_prev = main_prev if pull_waits_for_reduce else None
        pulls = []
        for k in range(K):
            pull_prev = add("P", t, k, [pull_prev] + ([pulls[-1]] if pulls else []))
            pulls.append(pull_prev)
        m = main_prev
        for k in range(K):
            m = add("WAIT", t, k, [m])
            m = add("FWD", t, k, [m])
        w = None              # wgrad stream
        for k in range(K - 1, -1, -1):
            m = add("BWD", t, k, [m])
            if k >= 1:
                w = add("SEND", t, k, [m, w])      # behind BWD(t,k), hence behind the whole forward pass
        m = add("SEND", t, 0, [m, w])
        m = add("REDUCE", t, None, [m, pulls[-1]])      # main joins the pull stream at the end of the step
        main_prev = m
    return ops


class World:
    def __init__(self, W, K, steps, **variant):
        self.W, self.K, self.steps = W, K, steps
        self.prog = [build_program(r, K, steps, **variant) for r in range(W)]
        self.done = [[False] * len(p) for p in self.prog]
        self.version = [0] * W                                    # owner o's master holds `version[o]` updates
        self.weights = [[[None] * W for _ in range(K)] for _ in range(W)]     # [worker][group][owner] -> version
        self.segcount = [[0] * K for _ in range(W)]               # [worker][group]
        self.stage = [[[None] * K for _ in range(W)] for _ in range(W)]       # [owner][source][group] -> round
        self.arrive = [[0] * W for _ in range(W)]                 # [owner][source]
        self.reduced = [0] * W                                    # rounds reduced by owner o

    def ready(self, r, i):
        if self.done[r][i]:
            return False
        name, t, k, deps = self.prog[r][i]
        if not all(self.done[r][d] for d in deps):
            return False
        if name == "WAIT":                                        # wait_seg_kernel: every owner delivered group k
            return self.segcount[r][k] >= (t + 1) * self.W
        if name == "REDUCE":                                      # every source signalled all K groups of this round
            return all(self.arrive[r][s] >= (t + 1) * self.K for s in range(self.W))
        return True

    def run(self, r, i):
        name, t, k, _ = self.prog[r][i]
        W, K = self.W, self.K
        if name == "P":
            for w in range(W):
                self.weights[w][k][r] = self.version[r]
                self.segcount[w][k] += 1
        elif name in ("FWD", "BWD"):
            got = self.weights[r][k]
            if any(v != t for v in got):
                raise Violation(f"{name} step {t} group {k} on rank {r} read shard versions {got}, wanted all {t}")
        elif name == "SEND":
            for o in range(W):
                if self.stage[o][r][k] is not None and self.reduced[o] < t:
                    raise Violation(f"SEND round {t} group {k} of source {r} overwrites owner {o}'s slot before its "
                                    f"round {self.stage[o][r][k]} was reduced")
                self.stage[o][r][k] = t
                self.arrive[o][r] += 1
        elif name == "REDUCE":
            for s in range(W):
                for g in range(K):
                    if self.stage[r][s][g] != t:
                        raise Violation(f"REDUCE round {t} on owner {r} read round {self.stage[r][s][g]} of source {s} "
                                        f"group {g}")
            self.version[r] += 1
            self.reduced[r] = t + 1
        self.done[r][i] = True

    def explore(self, schedule):
        """schedule: ints choosing (rank, which of its ready ops) -- any op whose stream predecessors are done may go."""
        it = iter(schedule)
        while True:
            ready = [(r, i) for r in range(self.W) for i in range(len(self.prog[r])) if self.ready(r, i)]
            if not ready:
                break
            c = next(it, 0)
            self.run(*ready[c % len(ready)])
        return all(all(d) for d in self.done)


@settings(max_examples=200, deadline=None)
@given(W=st.integers(1, 4), K=st.integers(1, 4), steps=st.integers(1, 4),
       schedule=st.lists(st.integers(0, 1 << 16), min_size=0, max_size=600))
def test_sync_round_invariants_under_any_interleaving(W, K, steps, schedule):
    world = World(W, K, steps)
    finished = world.explore(schedule)
    assert finished, "deadlock: some rank could not finish its steps"
    assert world.version == [steps] * W and world.reduced == [steps] * W


def test_a_fast_rank_cannot_run_ahead_of_the_barrier():
    """Adversarial schedule: always advance the lowest rank as far as it can go.  It must stall at the first WAIT /
    REDUCE that needs the others, and the round invariants still hold when they catch up."""
    world = World(4, 3, 3)
    while True:
        ready = [(r, i) for r in range(4) for i in range(len(world.prog[r])) if world.ready(r, i)]
        if not ready:
            break
        world.run(*min(ready))
    assert all(all(d) for d in world.done)


def _violates(**variant):
    hit = 0
    for seed in range(300):
        import random
        rnd = random.Random(seed)
        world = World(3, 3, 3, **variant)
        try:
            world.explore(rnd.randrange(1 << 16) for _ in range(5000))
        except Violation:
            hit += 1
    return hit


def test_the_model_has_teeth_pull_must_follow_the_reduce():
    """Without the edge 'pull of step t+1 starts after the main stream of step t' (Worker._pre) an owner can broadcast
    its shard BEFORE reducing into it: some forward pass then reads a stale shard.  The model must see that."""
    assert _violates(pull_waits_for_reduce=False) > 0


def test_the_model_has_teeth_send_must_follow_all_forward_waits():
    """The per-group SENDs of the backward pass are safe only because they come after the forward pass, which waited
    for EVERY group of EVERY owner's pull, and each owner pulls after its own reduce.  Issue a SEND without that edge
    (as if it could overtake the forward pass) and it can overwrite a slot that an owner has not reduced yet."""
    with pytest.raises(Violation):
        # direct construction: source 0 of a 2-rank world sends round 1 before owner 1 has reduced round 0
        world = World(2, 2, 2)
        world.stage[1][0][1] = 0
        world.reduced[1] = 0
        world.prog[0] = [("SEND", 1, 1, [])]
        world.done[0] = [False]
        world.run(0, 0)
