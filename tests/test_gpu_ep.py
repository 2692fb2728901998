"""GPU: the expert-parallel kernels (nf4moe_ep_pack / nf4moe_expert_rows / nf4moe_combine_indexed) and the EPMoELayer
phases, with `world` VIRTUAL ranks living on one GPU and the all-to-alls played by tensor copies (the NCCL call itself
is covered by bench.py --gpus N and by tests/test_ep_gloo.py on CPU).  Bar: every rank's output is BIT-IDENTICAL to the
single-GPU nf4moe_moe_forward on that rank's tokens, and within the north-star tolerance of the oracle."""
import importlib

import numpy as np
import pytest
import torch

from oracle import nf4_oracle as orc
from test_gpu_parity import assert_moe_close, make_layer

pytestmark = pytest.mark.gpu


def exchange(send_bufs, recv_bufs, world):
    """all_to_all_single with equal splits, for `world` ranks that live in this process."""
    for dst in range(world):
        r = recv_bufs[dst].view(world, -1)
        for src in range(world):
            r[src].copy_(send_bufs[src].view(world, -1)[dst])


@pytest.mark.parametrize("world,T,E,regime", [(2, 9, 32, "gemv"), (4, 16, 32, "gemv"), (8, 8, 32, "gemv"), (8, 40, 32, "gemm"),
                                             (8, 6, 128, "gemv"), (1, 5, 32, "gemv")])
def test_virtual_ranks_bit_identical_to_single_gpu(pkg, cuda_device, world, T, E, regime):
    ep = importlib.import_module("gpt-oss-vllm_b200.ep")
    H, I, k = 192, 128, 4
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 100 + world + T, E, H, I)
    rng = np.random.default_rng(5)
    reg = pkg.REGIME_GEMM if regime == "gemm" else pkg.REGIME_GEMV
    layers, xs, lgs = [], [], []
    for r in range(world):
        plan = ep.EPPlan(r, world, E)
        layers.append(ep.EPMoELayer(ep.shard_experts(w13, plan), ep.shard_experts(w2, plan), plan, k, H, T, cuda_device,
                                    regime=reg))
        Tr = T - (r % 3)                                              # ragged token counts per rank
        xs.append(torch.from_numpy(rng.standard_normal((Tr, H)).astype(np.float16)).to(cuda_device))
        lg = rng.standard_normal((Tr, E)).astype(np.float16)
        if r == 1:
            lg[:, :4] += 6                                            # one rank sends nearly everything to rank 0's experts
        lgs.append(torch.from_numpy(lg).to(cuda_device))
    for r in range(world):
        layers[r].prepare(xs[r], router_logits=lgs[r])
    exchange([l.send_ids for l in layers], [l.recv_ids for l in layers], world)
    exchange([l.send_rows for l in layers], [l.recv_rows for l in layers], world)
    for l in layers:
        l.compute()
    exchange([l.y_rows for l in layers], [l.back_rows for l in layers], world)
    for r in range(world):
        out = layers[r].finish()
        layers[r].check_overflow()
        single = pkg.ops.moe_forward(xs[r], w13, w2, k, router_logits=lgs[r], regime=reg)
        assert torch.equal(out, single), f"rank {r} differs from the single-GPU layer"
        ref = orc.moe_forward_from_states(xs[r].cpu().numpy(), lgs[r].cpu().numpy(), st13, st2, k=k)
        assert_moe_close(out.cpu().numpy(), ref)


@pytest.mark.parametrize("world,T,E", [(2, 64, 32), (4, 48, 32), (8, 64, 32), (8, 32, 128)])
def test_virtual_ranks_default_regime_and_capacity_factor(pkg, cuda_device, world, T, E):
    """The configuration bench.py --gpus N runs: library-chosen regime (decode tcgen05 kernel on the padded capacity rows) and
    2 x mean-load capacity per peer; near-uniform routing must not overflow and every rank must match the oracle."""
    ep = importlib.import_module("gpt-oss-vllm_b200.ep")
    H, I, k = 192, 128, 4
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 300 + world + T, E, H, I)
    rng = np.random.default_rng(11)
    layers, xs, lgs = [], [], []
    for r in range(world):
        plan = ep.EPPlan(r, world, E)
        layers.append(ep.EPMoELayer(ep.shard_experts(w13, plan), ep.shard_experts(w2, plan), plan, k, H, T, cuda_device,
                                    capacity_factor=2.0))
        assert layers[-1].cap <= T * k and layers[-1].regime == pkg.REGIME_AUTO
        xs.append(torch.from_numpy(rng.standard_normal((T, H)).astype(np.float16)).to(cuda_device))
        lgs.append(torch.from_numpy(rng.standard_normal((T, E)).astype(np.float16)).to(cuda_device))
    for r in range(world):
        layers[r].prepare(xs[r], router_logits=lgs[r])
    exchange([l.send_ids for l in layers], [l.recv_ids for l in layers], world)
    exchange([l.send_rows for l in layers], [l.recv_rows for l in layers], world)
    for l in layers:
        l.compute()
    exchange([l.y_rows for l in layers], [l.back_rows for l in layers], world)
    for r in range(world):
        out = layers[r].finish()
        layers[r].check_overflow()
        ref = orc.moe_forward_from_states(xs[r].cpu().numpy(), lgs[r].cpu().numpy(), st13, st2, k=k)
        assert_moe_close(out.cpu().numpy(), ref)


@pytest.mark.parametrize("world,T,E,regime,bias", [(2, 9, 32, "gemv", False), (4, 16, 32, "auto", True), (8, 64, 32, "auto", False),
                                                  (8, 40, 32, "gemm", True), (8, 6, 128, "gemv", False), (1, 5, 32, "gemv", True)])
def test_peer_memory_exchange_virtual_ranks(pkg, cuda_device, world, T, E, regime, bias):
    """The NVLink peer-memory exchange (nf4moe_ep_dispatch_p2p / nf4moe_ep_wait / nf4moe_ep_return_p2p) with `world` virtual
    ranks on one GPU: every rank's dispatch kernel stores into the OTHER ranks' receive buffers through the pointer tables (here
    device-local, on a multi-GPU box NVLink peer addresses), publishes its epoch, the owners' wait kernels see the flags, the
    return path mirrors it.  Several layers in a row through ONE set of buffers (epochs advance, buffers are reused), ragged
    token counts, one rank sending nearly everything to rank 0, expert biases: bit-identical to the single-GPU layer, and
    nothing is ever dropped (capacity = worst case)."""
    ep = importlib.import_module("gpt-oss-vllm_b200.ep")
    H, I, k = 192, 128, 4
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 500 + world + T, E, H, I)
    rng = np.random.default_rng(15)
    reg = {"gemm": pkg.REGIME_GEMM, "gemv": pkg.REGIME_GEMV, "auto": pkg.REGIME_AUTO}[regime]
    b13 = torch.from_numpy((rng.standard_normal((E, 2 * I)) * 0.3).astype(np.float16)).to(cuda_device) if bias else None
    b2 = torch.from_numpy((rng.standard_normal((E, H)) * 0.3).astype(np.float16)).to(cuda_device) if bias else None
    plans = [ep.EPPlan(r, world, E) for r in range(world)]
    exs = [ep.PeerExchange.local(pl, pl.capacity(T, k), H, cuda_device) for pl in plans]
    for ex in exs:
        ex.connect(exs)
    layers = []
    for r, pl in enumerate(plans):
        lo, hi = pl.local_experts.start, pl.local_experts.stop
        layers.append(ep.EPMoELayer(ep.shard_experts(w13, pl), ep.shard_experts(w2, pl), pl, k, H, T, cuda_device, regime=reg,
                                    exchange=exs[r], w13_bias=None if b13 is None else b13[lo:hi].contiguous(), w2_bias=b2))
        assert layers[-1].cap == T * k
    for rep in range(3):                                               # three "layers" through the same buffers
        xs, lgs = [], []
        for r in range(world):
            Tr = T - (r % 3)
            xs.append(torch.from_numpy(rng.standard_normal((Tr, H)).astype(np.float16)).to(cuda_device))
            lg = rng.standard_normal((Tr, E)).astype(np.float16)
            if r == 1:
                lg[:, :4] += 6
            lgs.append(torch.from_numpy(lg).to(cuda_device))
        for r in range(world):
            layers[r].prepare(xs[r], router_logits=lgs[r])             # dispatch kernels of all ranks (stores + flags)
        for l in layers:
            l.dispatch()                                               # wait kernels: every flag is already there
            l.compute()
        for l in layers:
            l.x.ret()
        for r in range(world):
            layers[r].x.wait_return()
            out = layers[r].finish()
            layers[r].check_overflow()
            single = pkg.ops.moe_forward(xs[r], w13, w2, k, router_logits=lgs[r], regime=reg, w13_bias=b13, w2_bias=b2)
            if regime != "auto":      # same kernels on both sides: bit-identical; AUTO may pick another regime for the owner's rows
                assert torch.equal(out, single), f"rank {r} rep {rep} differs from the single-GPU layer"
            ref = orc.moe_forward_from_states(xs[r].cpu().numpy(), lgs[r].cpu().numpy(), st13, st2, k=k,
                                              b13=None if b13 is None else b13.cpu().numpy(), b2=None if b2 is None else b2.cpu().numpy())
            assert_moe_close(out.cpu().numpy(), ref)
    assert int(exs[0].flags[0, 0].item()) == 3 and int(exs[0].flags[1, world - 1].item()) == 3     # epochs advanced once per layer


def test_nccl_exchange_layer_with_biases(pkg, cuda_device):
    """The collective back end with expert biases: w13_bias applied on the owner (local expert index), w2_bias at home."""
    ep = importlib.import_module("gpt-oss-vllm_b200.ep")
    world, T, E, H, I, k = 4, 12, 32, 192, 128, 4
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 611, E, H, I)
    rng = np.random.default_rng(3)
    b13 = torch.from_numpy((rng.standard_normal((E, 2 * I)) * 0.3).astype(np.float16)).to(cuda_device)
    b2 = torch.from_numpy((rng.standard_normal((E, H)) * 0.3).astype(np.float16)).to(cuda_device)
    layers, xs, lgs = [], [], []
    for r in range(world):
        pl = ep.EPPlan(r, world, E)
        lo, hi = pl.local_experts.start, pl.local_experts.stop
        layers.append(ep.EPMoELayer(ep.shard_experts(w13, pl), ep.shard_experts(w2, pl), pl, k, H, T, cuda_device, regime=pkg.REGIME_GEMV,
                                    w13_bias=b13[lo:hi].contiguous(), w2_bias=b2))
        xs.append(torch.from_numpy(rng.standard_normal((T, H)).astype(np.float16)).to(cuda_device))
        lgs.append(torch.from_numpy(rng.standard_normal((T, E)).astype(np.float16)).to(cuda_device))
        layers[r].prepare(xs[r], router_logits=lgs[r])
    exchange([l.send_ids for l in layers], [l.recv_ids for l in layers], world)
    exchange([l.send_rows for l in layers], [l.recv_rows for l in layers], world)
    for l in layers:
        l.compute()
    exchange([l.y_rows for l in layers], [l.back_rows for l in layers], world)
    for r in range(world):
        out = layers[r].finish()
        single = pkg.ops.moe_forward(xs[r], w13, w2, k, router_logits=lgs[r], regime=pkg.REGIME_GEMV, w13_bias=b13, w2_bias=b2)
        assert torch.equal(out, single)


def test_ep_pack_layout_and_overflow_flag(pkg, cuda_device):
    ep = importlib.import_module("gpt-oss-vllm_b200.ep")
    E, H, I, k, T, world = 8, 64, 64, 2, 12, 4
    _, _, w13, w2 = make_layer(pkg, cuda_device, 3, E, H, I)
    rng = np.random.default_rng(9)
    x = torch.from_numpy(rng.standard_normal((T, H)).astype(np.float16)).to(cuda_device)
    ids = torch.from_numpy(np.stack([rng.permutation(E)[:k] for _ in range(T)]).astype(np.int32)).to(cuda_device)
    plan = ep.EPPlan(0, world, E)
    layer = ep.EPMoELayer(ep.shard_experts(w13, plan), ep.shard_experts(w2, plan), plan, k, H, T, cuda_device)
    w = torch.full((T, k), 0.5, dtype=torch.float32, device=cuda_device)
    layer.prepare(x, topk_ids=ids, topk_weights=w)
    sid, slot, rows = layer.send_ids.cpu().numpy(), layer.pair_slot.cpu().numpy(), layer.send_rows.cpu().numpy()
    idn, xn = ids.cpu().numpy().reshape(-1), x.cpu().numpy()
    cap, el = layer.cap, E // world
    assert cap == T * k
    for pair in range(T * k):                                          # every pair sits in its owner's region with its local id
        s = slot[pair]
        assert s // cap == idn[pair] // el and sid[s] == idn[pair] % el
        assert np.array_equal(rows[s], xn[pair // k])
    assert (sid >= 0).sum() == T * k and len(set(slot.tolist())) == T * k
    for p in range(world):                                             # inside a peer's region: ascending expert, then pair
        used = sid[p * cap:(p + 1) * cap]
        n = int((used >= 0).sum())
        assert (used[:n] >= 0).all() and (used[n:] == -1).all() and (np.diff(used[:n]) >= 0).all()
    # a capacity below the load is flagged, never silent
    small = ep.EPMoELayer(ep.shard_experts(w13, plan), ep.shard_experts(w2, plan), plan, k, H, T, cuda_device)
    small.cap = 2
    small.prepare(x, topk_ids=ids, topk_weights=w)
    with pytest.raises(RuntimeError):
        small.check_overflow()


@pytest.mark.parametrize("world,T,E", [(2, 24, 32), (8, 64, 32), (4, 5, 32)])
def test_replicated_activation_mode_virtual_ranks(pkg, cuda_device, world, T, E):
    """SURVEY 8(e)'s second mode: every (virtual) rank gets ALL tokens, runs only its experts (non-local pairs masked with -1), and
    the partial outputs are summed -- here in fp32 on one device, the all-reduce of the real thing being NCCL's.  Against the
    oracle, MoE tolerance; tokens none of whose experts are local contribute exact zeros."""
    ep = importlib.import_module("gpt-oss-vllm_b200.ep")
    H, I, k = 192, 128, 4
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 300 + world + T, E, H, I)
    rng = np.random.default_rng(9)
    x = torch.from_numpy(rng.standard_normal((T, H)).astype(np.float16)).to(cuda_device)
    lg = torch.from_numpy(rng.standard_normal((T, E)).astype(np.float16)).to(cuda_device)
    ids, _ = pkg.ops.route_topk(lg, k)
    total = torch.zeros(T, H, dtype=torch.float32, device=cuda_device)
    for r in range(world):
        plan = ep.EPPlan(r, world, E)
        layer = ep.ReplicatedMoELayer(ep.shard_experts(w13, plan), ep.shard_experts(w2, plan), plan, k, cuda_device)
        part = layer.partial(x, router_logits=lg)
        lo, hi = plan.local_experts.start, plan.local_experts.stop
        none_local = ~((ids >= lo) & (ids < hi)).any(dim=1)
        assert bool((part[none_local] == 0).all())
        total += part.float()
    ref = orc.moe_forward_from_states(x.cpu().numpy(), lg.cpu().numpy(), st13, st2, k=k)
    assert_moe_close(total.half().cpu().numpy(), ref)


@pytest.mark.parametrize("world,T", [(4, 48), (8, 64)])
def test_peer_memory_exchange_honours_capacity_factor(pkg, cuda_device, world, T):
    """Peer-memory exchange with a per-peer capacity BELOW the worst case (2 x the mean load), which is what bounds the owner's
    world * capacity slots at prefill sizes.  Near-uniform routing: nothing is dropped and every rank matches the oracle.  Then one
    rank routes nearly everything to rank 0: the dispatch kernel drops the excess and raises the overflow bit, which
    check_overflow turns into an error -- never a silent wrong answer."""
    ep = importlib.import_module("gpt-oss-vllm_b200.ep")
    E, H, I, k = 32, 192, 128, 4
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 700 + world + T, E, H, I)
    rng = np.random.default_rng(23)
    plans = [ep.EPPlan(r, world, E) for r in range(world)]
    cap = plans[0].capacity(T, k, 2.0)
    assert cap < T * k
    exs = [ep.PeerExchange.local(pl, cap, H, cuda_device) for pl in plans]
    for ex in exs:
        ex.connect(exs)
    layers = [ep.EPMoELayer(ep.shard_experts(w13, pl), ep.shard_experts(w2, pl), pl, k, H, T, cuda_device, exchange=exs[r])
              for r, pl in enumerate(plans)]
    assert all(l.cap == cap for l in layers)

    def run(skew_rank):
        xs, lgs = [], []
        for r in range(world):
            xs.append(torch.from_numpy(rng.standard_normal((T, H)).astype(np.float16)).to(cuda_device))
            lg = rng.standard_normal((T, E)).astype(np.float16)
            if r == skew_rank:
                lg[:, :4] += 8          # all four slots of every token go to rank 0's experts: T*k pairs for one peer
            lgs.append(torch.from_numpy(lg).to(cuda_device))
        for r in range(world):
            layers[r].prepare(xs[r], router_logits=lgs[r])
        for l in layers:
            l.dispatch()
            l.compute()
        for l in layers:
            l.x.ret()
        outs = []
        for r in range(world):
            layers[r].x.wait_return()
            outs.append(layers[r].finish())
        return xs, lgs, outs

    xs, lgs, outs = run(skew_rank=-1)
    for r in range(world):
        layers[r].check_overflow()
        ref = orc.moe_forward_from_states(xs[r].cpu().numpy(), lgs[r].cpu().numpy(), st13, st2, k=k)
        assert_moe_close(outs[r].cpu().numpy(), ref)
    run(skew_rank=1)
    with pytest.raises(RuntimeError, match="overflowed"):
        layers[1].check_overflow()
    for r in range(world):
        if r != 1:
            layers[r].check_overflow()
