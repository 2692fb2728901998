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
