"""CPU, world_size 2 over gloo: the expert-parallel choreography of gpt-oss-vllm_b200/ep.py (plan, fixed-capacity send
buffers, three all-to-alls, slot maps) with the device steps played by the oracle (tests/ep_oracle_ops.py).  The result on
every rank must equal the single-process oracle layer on that rank's tokens (to the last fp16 bit or two: CPU matmul blocking)."""
import importlib
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
E, H, I, K, T = 8, 128, 64, 4, 11


def _inputs():
    from oracle import nf4_oracle as orc
    rng = np.random.default_rng(77)
    st13 = [orc.quantize_nf4((rng.standard_normal((2 * I, H)) * 0.05).astype(np.float16)) for _ in range(E)]
    st2 = [orc.quantize_nf4((rng.standard_normal((H, I)) * 0.05).astype(np.float16)) for _ in range(E)]
    d13 = np.stack([orc.dequant_state(s) for s in st13])
    d2 = np.stack([orc.dequant_state(s) for s in st2])
    return rng, d13, d2


def _worker(rank, world, port, skew, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ep_oracle_ops import OracleEPOps
        from oracle import nf4_oracle as orc
        ep = importlib.import_module("gpt-oss-vllm_b200.ep")
        rng, d13, d2 = _inputs()
        xs = [rng.standard_normal((T - r, H)).astype(np.float16) for r in range(world)]     # ragged: ranks differ in T
        lgs = [rng.standard_normal((T - r, E)).astype(np.float16) for r in range(world)]
        if skew:                                                                          # everything to rank 0's experts
            for lg in lgs:
                lg[:, : E // world] += 8
        plan = ep.EPPlan(rank, world, E)
        lo, hi = plan.local_experts.start, plan.local_experts.stop
        layer = ep.EPMoELayer(None, None, plan, K, H, T, "cpu", device_ops=OracleEPOps(d13[lo:hi], d2[lo:hi]))
        out = layer.forward(torch.from_numpy(xs[rank]), router_logits=torch.from_numpy(lgs[rank]))
        layer.check_overflow()
        ref = orc.moe_forward_ref(xs[rank], lgs[rank], d13, d2, k=K)
        got = out.numpy().astype(np.float32)
        # torch's CPU matmul may block rows differently for different batch sizes (the owner multiplies rows of several
        # ranks at once), so allow last-bit fp16 differences; a choreography error would be O(1)
        q.put((rank, float(np.abs(got - ref.astype(np.float32)).max()), float(np.mean(got == ref.astype(np.float32))), layer.cap,
               int((layer.send_ids >= 0).sum())))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("skew", [False, True])
def test_ep_world2_matches_single_process_oracle(skew):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + (7 if skew else 0)
    procs = [ctx.Process(target=_worker, args=(r, world, port, skew, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, max_abs, frac_equal, cap, used in res:
        assert max_abs <= 1e-3 and frac_equal >= 0.98, f"rank {rank} differs from the single-process oracle: {max_abs} {frac_equal}"
        assert cap == T * K and used == (T - rank) * K


def _worker_replicated(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ep_oracle_ops import OracleEPOps
        from oracle import nf4_oracle as orc
        ep = importlib.import_module("gpt-oss-vllm_b200.ep")
        rng, d13, d2 = _inputs()
        x = rng.standard_normal((T, H)).astype(np.float16)            # the SAME tokens on every rank (replicated activations)
        lg = rng.standard_normal((T, E)).astype(np.float16)
        plan = ep.EPPlan(rank, world, E)
        lo, hi = plan.local_experts.start, plan.local_experts.stop
        layer = ep.ReplicatedMoELayer(None, None, plan, K, "cpu", device_ops=OracleEPOps(d13[lo:hi], d2[lo:hi]))
        part = layer.partial(torch.from_numpy(x), router_logits=torch.from_numpy(lg)).clone()
        out = layer.forward(torch.from_numpy(x), router_logits=torch.from_numpy(lg))
        ref = orc.moe_forward_ref(x, lg, d13, d2, k=K).astype(np.float32)
        got = out.numpy().astype(np.float32)
        ids, _ = orc.route_topk(lg, K)
        has_local = ((ids >= lo) & (ids < hi)).any(axis=1)
        q.put((rank, float(np.abs(got - ref).max()), float(np.linalg.norm(got - ref) / np.linalg.norm(ref)),
               bool((part.numpy()[~has_local] == 0).all())))
    finally:
        dist.destroy_process_group()


def test_replicated_activation_mode_world2():
    """SURVEY 8(e)'s second mode over gloo: every rank holds all tokens, computes its own experts' share, one all-reduce."""
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + 13
    procs = [ctx.Process(target=_worker_replicated, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, max_abs, rel, zero_rows_ok in res:
        assert max_abs <= 1e-2 and rel <= 2e-3, f"rank {rank}: {max_abs} {rel}"       # fp16 sum of two partials: tolerance, not bits
        assert zero_rows_ok, "tokens with no local expert must contribute exactly zero"


def test_plan_partition_and_capacity():
    ep = importlib.import_module("gpt-oss-vllm_b200.ep")
    p = ep.EPPlan(3, 8, 32)
    assert list(p.local_experts) == [12, 13, 14, 15] and p.owner(12) == 3 and p.owner(31) == 7 and p.experts_per_rank == 4
    assert p.capacity(32, 4) == 128                       # worst case: cannot overflow
    assert p.capacity(32, 4, factor=2.0) == 32            # 2 x the mean load of 16, multiple of 16
    assert ep.EPPlan(0, 8, 128).experts_per_rank == 16    # the 120b shape
    with pytest.raises(ValueError):
        ep.EPPlan(0, 3, 32)
    with pytest.raises(RuntimeError):                     # no CPU fallback unless a device_ops object is injected
        ep.EPMoELayer(None, None, p, 4, 64, 8, "cpu")


def test_owner_slot_limit_is_checked_before_any_kernel_runs():
    """Every owner buckets world * capacity receive slots, and nf4moe_bucket takes 204,800 pairs: at the worst-case capacity
    8 ranks stop at 6,400 tokens per rank (an 8,192-token prefill chunk used to fail inside the first forward with EINVAL).  The
    plan refuses at construction and says what to do; a capacity factor brings the same chunk well inside the limit."""
    ep = importlib.import_module("gpt-oss-vllm_b200.ep")
    p = ep.EPPlan(0, 8, 32)
    assert p.check_owner_slots(p.capacity(6400, 4)) == 204800
    with pytest.raises(ValueError, match="capacity_factor"):
        p.check_owner_slots(p.capacity(8192, 4), 2880)
    cap = p.capacity(8192, 4, factor=2.0)
    assert cap == 8192 and p.check_owner_slots(cap, 2880) == 65536
    with pytest.raises(ValueError, match="receive slots"):          # make_exchange checks before it touches any device
        ep.make_exchange(p, 4, 2880, 8192, "cpu")
    with pytest.raises(ValueError, match="receive slots"):          # ... and so does the layer, whoever built the exchange
        class _Ex:
            cap, kind = 8192 * 4, "p2p"
        ep.EPMoELayer(None, None, p, 4, 2880, 8192, "cpu", device_ops=object(), exchange=_Ex())
