"""Push / Pull / Async kernels (csrc/ps_kernels.cu) on ONE GPU with W emulated ranks driven by a single
launch (rank_lo=0, rank_hi=W) -- kernels that wait on each other are never run as separate launches
on one device.  Oracle: plain fp32 PyTorch restatement of the reference numerics
(reference: src/parameter_server/server.py:126-186; fp16 wire: src/workers/worker.py:264-268)."""
import numpy as np
import pytest
import torch

from dps_b200.config import Config
from dps_b200.models.resnet import build_model
from dps_b200.store.shardmap import build_layout

pytestmark = pytest.mark.gpu


def _setup(W, **cfg_kw):
    from dps_b200.parallel.ps_device import DevicePS
    model = build_model(100, seed=0)
    layout = build_layout(model, W)
    cfg = Config(workers=W, **cfg_kw)
    ps = DevicePS(layout, cfg, rank=0, world=W, device_index=0, emulate=True, timeout_s=5.0)
    flat = torch.zeros(layout.total)
    layout.flatten_into(flat, dict(model.named_parameters()))
    ps.load_master(flat)
    return model, layout, ps, flat.cuda()


def test_layout_views_roundtrip_cpu_side():
    model = build_model(100, seed=0)
    layout = build_layout(model, 8)
    flat = torch.zeros(layout.total)
    layout.flatten_into(flat, dict(model.named_parameters()))
    back = layout.unflatten(flat)
    for n, p in model.named_parameters():
        assert torch.equal(back[n], p.detach()), n
    sh = layout.to_shards(flat)
    assert sh.shape == (8, layout.shard_elems) and torch.equal(layout.from_shards(sh), flat)


@pytest.mark.parametrize("W,momentum,wd", [(1, 0.0, 0.0), (4, 0.0, 0.0), (8, 0.9, 5e-4), (3, 0.9, 0.0)])
def test_sync_push_pull_matches_reference_numerics(W, momentum, wd):
    model, layout, ps, flat0 = _setup(W, mode="sync", lr=0.1, momentum=momentum, weight_decay=wd)
    p = flat0.clone()
    m = torch.zeros_like(p)
    for step in range(3):
        gs = [torch.randn(layout.total, device="cuda", generator=torch.Generator("cuda").manual_seed(100 * step + r)) * 0.05
              for r in range(W)]
        for r in range(W):
            ps.grad[r].copy_(gs[r])
        ps.sync_round()
        torch.cuda.synchronize()
        g = torch.stack([x.half().float() for x in gs]).sum(0) / W          # fp16 wire, fp32 mean
        g = g + wd * p
        if momentum:
            m = momentum * m + g
            g = m
        p = p - 0.1 * g
        got = ps.gather_master()
        assert torch.allclose(got, p, rtol=1e-5, atol=1e-6), float((got - p).abs().max())
        for r in range(W):
            assert torch.equal(ps.wbf16[r], got.to(torch.bfloat16))               # pull: cast + broadcast
            assert torch.equal(ps.small[r], got[: layout.n_small])
            assert float(ps.grad[r].abs().max()) == 0.0                           # consumed gradients are zeroed
    assert ps.error_bits(0) == 0
    assert all(ps.version(r) == 3 for r in range(W))
    assert ps.flag_words(0)[64:64 + W].tolist() == [3] * W                       # heartbeats
    ps.close()


def test_sync_uncompressed_wire_is_exact():
    W = 2
    model, layout, ps, flat0 = _setup(W, mode="sync", lr=0.5, compress="none")
    gs = [torch.randn(layout.total, device="cuda") for _ in range(W)]
    for r in range(W):
        ps.grad[r].copy_(gs[r])
    ps.sync_round()
    want = flat0 - 0.5 * ((gs[0] + gs[1]) / 2)
    assert torch.allclose(ps.gather_master(), want, rtol=1e-6, atol=1e-7)
    ps.close()


def test_sync_matches_reference_servicer_by_name(ref_modules):
    """Same fp16 gradients through the reference servicer (numpy) and through the fused kernels."""
    W = 2
    model, layout, ps, flat0 = _setup(W, mode="sync", lr=0.1)
    ref = ref_modules["server"].ParameterServerServicer(mode="sync", total_workers=W, learning_rate=0.1)
    ref.parameters = {k: v.detach().numpy().copy() for k, v in model.state_dict().items()}
    names = [n for n, _ in model.named_parameters()]
    for r in range(W):
        torch.manual_seed(r)
        gdict = {n: (torch.randn_like(p) * 0.01) for n, p in model.named_parameters()}
        gflat = torch.zeros(layout.total)
        layout.flatten_into(gflat, gdict)
        ps.grad[r].copy_(gflat.cuda())
        g16 = {n: g.numpy().astype(np.float16) for n, g in gdict.items()}
        ref._handle_sync_gradients(r, ref.decompress_gradients(g16), 0)
    ps.sync_round()
    got = layout.unflatten(ps.gather_master().cpu())
    for n in names:
        np.testing.assert_allclose(got[n].numpy(), ref.parameters[n], rtol=1e-5, atol=1e-6, err_msg=n)
    ps.close()


def _cpu_async_sim(events, bound, alpha=0.1, floor=0.1):
    """events: list of (src, version_seen) in ticket order -> list of (accepted, weight, staleness)."""
    v, out = 0, []
    for src, seen in events:
        s = v - seen
        if s > bound:
            out.append((False, 0.0, s))
        else:
            out.append((True, max(floor, 1.0 / (1.0 + alpha * s)), s))
            v += 1
    return out, v


def test_async_ring_staleness_and_rejection():
    W, bound = 3, 2
    model, layout, ps, flat0 = _setup(W, mode="async", lr=0.1, staleness_bound=bound)
    p = flat0.clone()
    # schedule: all pull at v0; ranks push in order 0,1,2 (staleness 0,1,2); rank 0 pushes twice more
    # without re-pulling (staleness 3 -> rejected, then still 3 -> rejected); rank 1 re-pulls and pushes (0)
    script = [("pull", 0), ("pull", 1), ("pull", 2), ("push", 0), ("push", 1), ("push", 2), ("serve", None),
              ("push", 0), ("push", 0), ("serve", None), ("pull", 1), ("push", 1), ("serve", None)]
    seen = {r: 0 for r in range(W)}
    events, grads = [], []
    ver = 0
    k = 0
    for op, r in script:
        if op == "pull":
            ps.pull_async(r)
            torch.cuda.synchronize()
            seen[r] = ps.version(0)
        elif op == "push":
            g = torch.randn(layout.total, device="cuda", generator=torch.Generator("cuda").manual_seed(k)) * 0.05
            k += 1
            ps.grad[r].copy_(g)
            ps.push_async(r)
            events.append((r, seen[r]))
            grads.append(g.half().float())
        else:
            for o in range(W):
                ps.serve_async(o)
            torch.cuda.synchronize()
    sim, v_final = _cpu_async_sim(events, bound)
    for (acc, w, s), g in zip(sim, grads):
        if acc:
            p = p - 0.1 * (w * g)
    got = ps.gather_master()
    assert torch.allclose(got, p, rtol=1e-5, atol=1e-6), float((got - p).abs().max())
    st = ps.async_stats(0)
    assert st["async_updates"] == sum(a for a, _, _ in sim) and st["rejected_updates"] == sum(not a for a, _, _ in sim)
    assert st["rejected_updates"] == 2 and st["version"] == v_final
    assert st["running_max_staleness"] == max(s for _, _, s in sim)
    # every shard saw the same arrival order here, so all versions agree
    assert all(ps.version(o) == v_final for o in range(W))
    # published bf16 shard == cast of master; a fresh pull delivers it
    ps.pull_async(2)
    torch.cuda.synchronize()
    assert torch.equal(ps.wbf16[2], got.to(torch.bfloat16))
    assert ps.error_bits(0) == 0
    ps.close()


def test_barrier_timeout_drops_dead_rank_and_flags_it():
    """SURVEY Q3 / §5.3 fault injection: rank 2 of 3 never pushes.  Owners do not hang: after the bounded
    wait they average over the ranks present, record the missing rank and raise the sticky error bit."""
    from dps_b200.ops import _lib as L
    from dps_b200.parallel.watchdog import decode_error_bits, missing_ranks
    W = 3
    model, layout, ps, flat0 = _setup(W, mode="sync", lr=0.1)
    ps.timeout_ns = int(0.2e9)
    ps.tables.rank_hi = 2                                     # ranks 0 and 1 are alive; rank 2 is dead
    gs = [torch.randn(layout.total, device="cuda", generator=torch.Generator("cuda").manual_seed(r)) * 0.05
          for r in range(2)]
    for r in range(2):
        ps.grad[r].copy_(gs[r])
    ps.push_sync()
    torch.cuda.synchronize()
    want = flat0 - 0.1 * (gs[0].half().float() + gs[1].half().float()) / 2
    got = ps.gather_master()
    sh_want, sh_got = layout.to_shards(want), layout.to_shards(got)
    for o in range(2):
        assert torch.allclose(sh_got[o], sh_want[o], rtol=1e-5, atol=1e-6)
        assert decode_error_bits(ps.error_bits(o)) == ["barrier_timeout"]
        assert missing_ranks(int(ps.flag_words(o)[L.F_MISSING]) & 0xFFFFFFFF, W) == [2]
    assert torch.equal(sh_got[2], layout.to_shards(flat0)[2])          # the dead rank's shard is untouched
    ps.close()


def test_heartbeat_watchdog_logic():
    from dps_b200.parallel.watchdog import HeartbeatWatchdog
    hb, now = [0, 0, 0], [0.0]
    wd = HeartbeatWatchdog(lambda: hb, 3, timeout_s=5.0, clock=lambda: now[0])
    assert wd.poll() == []
    for t in range(1, 8):
        now[0] = float(t)
        hb[0] += 1
        hb[1] += 1                                            # rank 2 never advances
        stale = wd.poll()
    assert stale == [2]


def test_checkpoint_resume_through_job(tmp_path):
    """W16: save from the sharded store, load into the reference-shaped torch module, resume a new Job."""
    from dps_b200 import Config
    from dps_b200.api import Job
    from dps_b200.models.resnet import ResNet18
    cfg = Config(mode="sync", workers=1, batch_size=16, lr=0.05, momentum=0.9)
    job = Job(cfg, rank=0, world=1, device_index=0)
    g = torch.Generator().manual_seed(0)
    x, y = torch.randn(16, 3, 32, 32, generator=g), torch.randint(0, 100, (16,), generator=g)
    for _ in range(3):
        job.worker.train_step(x, y).item()
    path = str(tmp_path / "ck.pt")
    job.server.save_checkpoint(path)
    sd = job.server.state_dict()
    mom = job.ps.gather_momentum().clone()
    job.close()
    m = ResNet18(100)
    m.load_state_dict(torch.load(path, weights_only=True), strict=True)        # reference-compatible file
    assert torch.equal(m.fc.weight, sd["fc.weight"])
    job2 = Job(cfg.replace(resume=path), rank=0, world=1, device_index=0)
    assert job2.server.global_step == 3
    sd2 = job2.server.state_dict()
    for k in ("conv1.weight", "layer3.0.shortcut.0.weight", "fc.bias", "bn1.running_var"):
        assert torch.allclose(sd2[k], sd[k]), k
    assert torch.allclose(job2.ps.gather_momentum(), mom)
    l = job2.worker.train_step(x, y).item()
    assert l == l and job2.server.global_step == 4
    job2.close()


def test_async_push_is_wait_free_when_the_ring_is_full():
    """A push never blocks on another rank: with no room in the owner's ring it is skipped, counted, and the
    gradient slice is kept (not zeroed) so it is sent with the next push."""
    W = 2
    model, layout, ps, flat0 = _setup(W, mode="async", lr=0.1, async_ring_slots=1)      # R = 2: room for W - ... 
    g = torch.randn(layout.total, device="cuda") * 0.01
    ps.grad[0].copy_(g)
    ps.push_async(0)                        # admitted on both shards (rings empty)
    torch.cuda.synchronize()
    assert float(ps.grad[0].abs().max()) == 0.0
    ps.grad[0].copy_(g)
    ps.push_async(0)                        # nobody served: no room (margin of W slots) -> skipped, not blocked
    torch.cuda.synchronize()
    assert ps.async_stats(0)["dropped_pushes"] == W
    assert torch.equal(ps.grad[0], g)       # kept for the next push
    for o in range(W):
        ps.serve_async(o)
    ps.push_async(0)                        # room again
    torch.cuda.synchronize()
    assert float(ps.grad[0].abs().max()) == 0.0 and ps.error_bits(0) == 0
    for o in range(W):
        ps.serve_async(o)
    torch.cuda.synchronize()
    want = flat0 - 0.1 * g.half().float() - 0.1 * g.half().float()      # two accepted pushes, staleness 0 then 1->w
    st = ps.async_stats(0)
    assert st["async_updates"] == 2
    ps.close()


def test_segmented_pull_delivers_layer_groups_in_order():
    """Pull overlapped with forward: five layer groups, per-group delivery counters, same bytes as the plain pull."""
    from dps_b200.ops import _lib as L
    W = 4
    model, layout, ps, flat0 = _setup(W, mode="sync", lr=0.1)
    bounds = [0] + [layout.entries[n].offset for n in ("layer2.0.conv1.weight", "layer3.0.conv1.weight",
                                                        "layer4.0.conv1.weight", "fc.weight")] + [layout.total]
    ps.set_segments(bounds)
    assert ps.nseg == 5
    new = torch.randn(layout.total, device="cuda")
    sh = layout.to_shards(new)
    for r in range(W):
        ps.master[r].copy_(sh[r])
        ps.wbf16[r].zero_()
    ps.pull_segments()
    for r in range(W):
        for k in range(ps.nseg):
            ps.wait_segment(k, rank=r)
    torch.cuda.synchronize()
    for r in range(W):
        assert torch.equal(ps.wbf16[r], new.to(torch.bfloat16))
        assert torch.equal(ps.small[r], new[: layout.n_small])
        counts = ps.flag_words(r)[L.F_SEGCOUNT:L.F_SEGCOUNT + 5].tolist()
        assert counts == [W * ps.ctas] * 5
    assert ps.error_bits(0) == 0
    ps.close()


@pytest.mark.parametrize("W,momentum", [(1, 0.0), (4, 0.9), (8, 0.0)])
def test_segmented_push_equals_monolithic_push(W, momentum):
    """Push overlapped with backward: SEND per layer group (last group first, each with its own CTA count), then one
    REDUCE over the whole shard -- same master, same zeroed gradients as the single launch."""
    model, layout, ps, flat0 = _setup(W, mode="sync", lr=0.1, momentum=momentum)
    bounds = [0] + [layout.entries[n].offset for n in ("layer2.0.conv1.weight", "layer3.0.conv1.weight",
                                                        "layer4.0.conv1.weight", "fc.weight")] + [layout.total]
    ps.set_segments(bounds)
    p, m = flat0.clone(), torch.zeros_like(flat0)
    for step in range(3):
        gs = [torch.randn(layout.total, device="cuda", generator=torch.Generator("cuda").manual_seed(10 * step + r)) * 0.05
              for r in range(W)]
        for r in range(W):
            ps.grad[r].copy_(gs[r])
        ctas = [ps.send_ctas(k) for k in range(ps.nseg)]
        for k in range(ps.nseg - 1, -1, -1):
            ps.push_segment(k, ps.SEND, ctas=ctas[k])
        ps.push_segment(-1, ps.REDUCE, arrivals_per_step=sum(ctas))
        ps.pull()
        ps.step += 1
        torch.cuda.synchronize()
        g = torch.stack([x.half().float() for x in gs]).sum(0) / W
        if momentum:
            m = momentum * m + g
            g = m
        p = p - 0.1 * g
        got = ps.gather_master()
        assert torch.allclose(got, p, rtol=1e-5, atol=1e-6), float((got - p).abs().max())
        for r in range(W):
            assert float(ps.grad[r].abs().max()) == 0.0
            assert torch.equal(ps.wbf16[r], got.to(torch.bfloat16))
    assert ps.error_bits(0) == 0 and all(ps.version(r) == 3 for r in range(W))
    ps.close()


def test_missed_barrier_removes_the_source_from_the_membership_mask():
    """A source that misses one barrier is not waited for again (bounded step time afterwards) and its stale staging
    slot is never read again; the sticky error bit stays up for the host watchdog."""
    import time
    from dps_b200.ops import _lib as L
    W = 2
    model, layout, ps, flat0 = _setup(W, mode="sync", lr=0.1)
    ps.timeout_ns = int(0.25e9)
    gen = lambda s: torch.randn(layout.total, device="cuda", generator=torch.Generator("cuda").manual_seed(s)) * 0.05
    # round 0: both sources arrive
    g0 = [gen(0), gen(1)]
    for r in range(W):
        ps.grad[r].copy_(g0[r])
    ps.push_sync(); ps.step += 1
    torch.cuda.synchronize()
    p = flat0 - 0.1 * (g0[0].half().float() + g0[1].half().float()) / 2
    # round 1: owner 0 never sees source 1's arrival (as if rank 1 had died before this push)
    g1 = [gen(2), gen(3)]
    for r in range(W):
        ps.grad[r].copy_(g1[r])
    ps.flags[0][L.F_ARRIVE + 1] -= ps.ctas
    torch.cuda.synchronize()
    t0 = time.time(); ps.push_sync(); ps.step += 1; torch.cuda.synchronize(); t_timeout = time.time() - t0
    assert t_timeout >= 0.2
    assert int(ps.flag_words(0)[L.F_ALIVE]) == 0b01 and int(ps.flag_words(1)[L.F_ALIVE]) == 0b11
    assert ps.error_bits(0) & L.ERR_BARRIER_TIMEOUT and ps.error_bits(1) == 0
    want = layout.to_shards(p)
    both = (g1[0].half().float() + g1[1].half().float()) / 2
    want0 = want[0] - 0.1 * layout.to_shards(g1[0].half().float())[0]          # owner 0: mean over the present source
    want1 = want[1] - 0.1 * layout.to_shards(both)[1]
    got = layout.to_shards(ps.gather_master())
    assert torch.allclose(got[0], want0, rtol=1e-5, atol=1e-6) and torch.allclose(got[1], want1, rtol=1e-5, atol=1e-6)
    # round 2: owner 0 no longer waits for source 1 -> no second timeout
    g2 = [gen(4), gen(5)]
    for r in range(W):
        ps.grad[r].copy_(g2[r])
    t0 = time.time(); ps.push_sync(); ps.step += 1; torch.cuda.synchronize(); t_after = time.time() - t0
    assert t_after < 0.1, t_after
    got2 = layout.to_shards(ps.gather_master())
    assert torch.allclose(got2[0], want0 - 0.1 * layout.to_shards(g2[0].half().float())[0], rtol=1e-5, atol=1e-6)
    ps.close()
