"""The fused sm_100a step vs the fp32 autograd oracle, same weights, same batch."""
import pytest
import torch

from fedrec_b200 import ops
from fedrec_b200.config import FedRecConfig
from fedrec_b200.data import make_synthetic_mind
from fedrec_b200.engine import TrainStep
from fedrec_b200.models import UserModel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _models(**kw):
    ops.require_native()
    data = make_synthetic_mind(num_news=600, num_train=8, num_valid=2, title_len=30, seed=5)
    cfg = FedRecConfig(title_len=30, dropout=0.0, batch_size=6, **kw)
    torch.manual_seed(0)
    a = UserModel(None, data.news_index, DEV, config=cfg)
    torch.manual_seed(0)
    b = UserModel(None, data.news_index, DEV, config=cfg)
    b.load_state_dict(a.state_dict())
    return a, b


@pytest.mark.parametrize("kw", [{}, {"mask_padding": True}, {"dedup": False}, {"cuda_graph": False}],
                         ids=["default", "mask_padding", "no_dedup", "no_graph"])
def test_fused_step_matches_oracle(kw):
    a, b = _models(**kw)
    g = torch.Generator(device=DEV).manual_seed(1)
    cand = torch.randint(1, 600, (6, 5), device=DEV, generator=g, dtype=torch.int32)
    his = torch.randint(0, 600, (6, 50), device=DEV, generator=g, dtype=torch.int32)
    fused, oracle = TrainStep(a, impl="fused"), TrainStep(b, impl="torch")
    la, lb = fused.run(cand, his), oracle.run(cand, his)
    assert float(la) == pytest.approx(float(lb), abs=5e-3)
    ga, gb = a.flat.grad, b.flat.grad
    assert float(gb.norm()) > 0
    rel = float((ga - gb).norm() / gb.norm())
    assert rel < 3e-2, rel
    # per tensor: within 3% of its own norm. The two additive-attention fc1 layers are the exception: their gradients are
    # small because the per-position factors sum to ~0 (the soft-max Jacobian), so what is left is dominated by the bf16
    # rounding of much larger terms; they must stay within 1% of the whole gradient's norm instead.
    report = []
    for name in a.flat.names:
        x, y = a.flat.view(name, grad=True), b.flat.view(name, grad=True)
        err, ny = float((x - y).norm()), float(y.norm())
        report.append((name, round(err / max(ny, 1e-30), 4)))
        if ny < 1e-4 * float(gb.norm()):
            # true value ~0 (a soft-max is invariant to a constant shift of its logits: the att_fc2 biases, and with few
            # valid positions the att_fc2 weights): only an absolute bound means anything
            assert err < 1e-3 * float(gb.norm()), (name, err, ny, report)
        else:
            if "additive_attention.att_fc1" in name:
                assert err < 0.1 * ny or err < 1e-2 * float(gb.norm()), (name, err, ny, float(gb.norm()), report)
            else:
                assert err < 0.03 * ny, (name, err, ny, report)
    sa, sb = fused.scores(cand, his), oracle.scores(cand, his)
    torch.testing.assert_close(sa, sb, atol=1e-2, rtol=1e-2)


def test_few_steps_track_oracle_loss():
    a, b = _models()
    a.cfg.lr = b.cfg.lr = 1e-3
    fused, oracle = TrainStep(a, impl="fused"), TrainStep(b, impl="torch")
    g = torch.Generator(device=DEV).manual_seed(2)
    for _ in range(5):
        cand = torch.randint(1, 600, (6, 5), device=DEV, generator=g, dtype=torch.int32)
        his = torch.randint(0, 600, (6, 50), device=DEV, generator=g, dtype=torch.int32)
        la, lb = fused.run(cand, his), oracle.run(cand, his)
        a.adam_step(); b.adam_step()
        assert float(la) == pytest.approx(float(lb), abs=1e-2)


def test_trunk_cache_is_exact_across_steps():
    """With the opt-in trunk cache, hidden states of news seen before are gathered from HBM instead of recomputed:
    losses and gradients must be bit-for-bit those of the uncached run (the trunk is frozen, the kernels deterministic
    up to the order of the atomics in the gradient accumulation)."""
    a, b = _models(trunk_cache=True)
    c, d = _models()
    c.load_state_dict(a.state_dict())
    cached, plain = TrainStep(a, impl="fused"), TrainStep(c, impl="fused")
    g = torch.Generator(device=DEV).manual_seed(3)
    for it in range(4):
        cand = torch.randint(1, 80, (6, 5), device=DEV, generator=g, dtype=torch.int32)      # small id range: many hits
        his = torch.randint(0, 80, (6, 50), device=DEV, generator=g, dtype=torch.int32)
        la, lc = cached.run(cand, his), plain.run(cand, his)
        # identical hidden states; the only run-to-run difference is the order of the atomic gradient accumulations
        assert float(la) == pytest.approx(float(lc), abs=1e-4), it
        rel = float((a.flat.grad - c.flat.grad).norm() / c.flat.grad.norm())
        assert rel < (1e-4 if it == 0 else 2e-2), (it, rel)      # later steps: two separately trained models drift slightly
        a.adam_step(); c.adam_step()
    valid = a._fused._cache["valid"]
    assert 0 < int(valid.sum()) <= 80
    # replacing the trunk weights invalidates the cache
    a.on_weights_changed()
    assert int(a._fused._cache["valid"].sum()) == 0


def test_fused_step_with_upstream_toy_shapes():
    """Title length 50 (the shipped toy shard) uses the 64-row attention slots; 3 users per batch."""
    ops.require_native()
    data = make_synthetic_mind(num_news=300, num_train=8, num_valid=2, title_len=50, seed=9)
    cfg = FedRecConfig(title_len=50, dropout=0.0, batch_size=3)
    torch.manual_seed(0)
    a = UserModel(None, data.news_index, DEV, config=cfg)
    torch.manual_seed(0)
    b = UserModel(None, data.news_index, DEV, config=cfg)
    b.load_state_dict(a.state_dict())
    g = torch.Generator(device=DEV).manual_seed(4)
    cand = torch.randint(1, 300, (3, 5), device=DEV, generator=g, dtype=torch.int32)
    his = torch.randint(0, 300, (3, 50), device=DEV, generator=g, dtype=torch.int32)
    la, lb = TrainStep(a, impl="fused").run(cand, his), TrainStep(b, impl="torch").run(cand, his)
    assert float(la) == pytest.approx(float(lb), abs=5e-3)
    rel = float((a.flat.grad - b.flat.grad).norm() / b.flat.grad.norm())
    assert rel < 5e-2, rel


def test_captured_graph_follows_trunk_weight_changes():
    """load_state_dict / full-state FedAvg replace the frozen trunk's weights: the packed bf16 copies are refilled in
    place, so the already captured CUDA graph computes with the new weights."""
    a, b = _models()
    g = torch.Generator(device=DEV).manual_seed(7)
    cand = torch.randint(1, 600, (6, 5), device=DEV, generator=g, dtype=torch.int32)
    his = torch.randint(0, 600, (6, 50), device=DEV, generator=g, dtype=torch.int32)
    fused, oracle = TrainStep(a, impl="fused"), TrainStep(b, impl="torch")
    l0 = float(fused.run(cand, his))                                    # captures the graph
    sd = {k: v.clone() for k, v in a.state_dict().items()}
    for k in sd:
        if "DistillBert" in k and k.endswith("lin1.weight"):
            sd[k] = sd[k] * 0.5
        if "DistillBert" in k and "word_embeddings" in k:
            sd[k] = sd[k] * 2.0
    a.load_state_dict(sd)
    b.load_state_dict(sd)
    la, lb = fused.run(cand, his), oracle.run(cand, his)
    assert float(la) == pytest.approx(float(lb), abs=5e-3)
    assert abs(float(la) - l0) > 1e-6                                   # the change was visible at all
    rel = float((a.flat.grad - b.flat.grad).norm() / b.flat.grad.norm())
    assert rel < 5e-2, rel


@pytest.mark.parametrize("graph", [True, False], ids=["graph", "eager"])
def test_pipelined_optimizer_equals_step_then_aggregate_then_adam(graph):
    """attach_optimizer moves clip + Adam of step i to the start of step i+1 (beside the frozen trunk). Same arithmetic
    in the same order as running them back to back: weights and losses must agree to fp32 rounding of the atomics."""
    from fedrec_b200.parallel import GradAggregator
    from fedrec_b200.parallel.symm import SymmetricHeap
    a, b = _models(cuda_graph=graph)
    a.cfg.lr = b.cfg.lr = 1e-3
    pipe, seq = TrainStep(a, impl="fused"), TrainStep(b, impl="fused")
    agg_a = GradAggregator(a.flat, 1, 0, "local", clip_norm=0.05, noise_multiplier=0.0, seed=3, symm=SymmetricHeap.single(a.device))
    agg_b = GradAggregator(b.flat, 1, 0, "local", clip_norm=0.05, noise_multiplier=0.0, seed=3, symm=SymmetricHeap.single(b.device))
    assert pipe.attach_optimizer(agg_a, 1e-3, 0.9, 0.999, 1e-8) and pipe.pipelined
    g = torch.Generator(device=DEV).manual_seed(11)
    for it in range(5):
        cand = torch.randint(1, 600, (6, 5), device=DEV, generator=g, dtype=torch.int32)
        his = torch.randint(0, 600, (6, 50), device=DEV, generator=g, dtype=torch.int32)
        la = float(pipe.run(cand, his))
        lb = float(seq.run(cand, his))
        agg_b.allreduce_mean_and_adam_(1e-3, 0.9, 0.999, 1e-8)
        assert la == pytest.approx(lb, abs=2e-4), it
    assert a.flat.step == 4 and b.flat.step == 5           # the pipelined model still owes the last update
    pipe.flush()
    pipe.flush()                                           # idempotent
    torch.cuda.synchronize()
    assert a.flat.step == 5
    assert agg_a.symm.pad_words(73, 1)[0] == 5             # device-side step counter agrees
    # The first moment is linear in the gradients: it must agree to the rounding of the atomics in the gradient
    # accumulation. The weights themselves are compared in bulk: Adam divides by sqrt(v), so a coordinate whose gradient
    # is rounding noise (the cancellation-prone text-head fc1) moves by +-lr whichever way that rounding went.
    rel_m = float((a.flat.exp_avg - b.flat.exp_avg).norm() / b.flat.exp_avg.norm())
    assert rel_m < 5e-3, rel_m
    rel_v = float((a.flat.exp_avg_sq - b.flat.exp_avg_sq).norm() / b.flat.exp_avg_sq.norm())
    assert rel_v < 5e-3, rel_v
    diff = (a.flat.param - b.flat.param).abs()
    assert float(diff.max()) < 5e-3 and float((diff > 2e-5).float().mean()) < 5e-2, (float(diff.max()), float((diff > 2e-5).float().mean()))
    rel_p = float((a.flat.param - b.flat.param).norm() / b.flat.param.norm())
    assert rel_p < 1e-3, rel_p
    # inference and checkpoints see the flushed weights
    cand = torch.randint(1, 600, (6, 5), device=DEV, generator=g, dtype=torch.int32)
    his = torch.randint(0, 600, (6, 50), device=DEV, generator=g, dtype=torch.int32)
    pipe.run(cand, his); seq.run(cand, his); agg_b.allreduce_mean_and_adam_(1e-3, 0.9, 0.999, 1e-8)
    torch.testing.assert_close(pipe.scores(cand, his), seq.scores(cand, his), atol=1e-4, rtol=1e-3)
    assert a.flat.step == 6


def test_run_packed_takes_pinned_host_ids():
    a, b = _models()
    fa, fb = TrainStep(a, impl="fused"), TrainStep(b, impl="fused")
    g = torch.Generator(device=DEV).manual_seed(12)
    cand = torch.randint(1, 600, (6, 5), device=DEV, generator=g, dtype=torch.int32)
    his = torch.randint(0, 600, (6, 50), device=DEV, generator=g, dtype=torch.int32)
    ids = torch.cat([his.reshape(-1), cand.reshape(-1)]).cpu().pin_memory()
    la = float(fa.run_packed(ids, 6, 5, 50))
    lb = float(fb.run(cand, his))
    assert la == pytest.approx(lb, abs=1e-5)
    rel = float((a.flat.grad - b.flat.grad).norm() / b.flat.grad.norm())
    assert rel < 1e-4, rel


@pytest.mark.parametrize("graph", [True, False], ids=["graph", "eager"])
def test_run_host_pulls_ids_from_pinned_memory_and_pushes_the_loss_back(graph):
    """run_host / loss_of: the step's first kernel reads its ids from a host-mapped ring, its last kernel writes
    {loss, sequence number} into host memory. Same losses and gradients as run() on device-resident ids, for more steps
    than the ring has slots (the host is held back instead of overwriting a slot that has not been consumed)."""
    a, b = _models(cuda_graph=graph)
    fa, fb = TrainStep(a, impl="fused"), TrainStep(b, impl="fused")
    fa.enable_host_io(6, 5, 50, slots=3)
    g = torch.Generator(device=DEV).manual_seed(13)
    tickets, want = [], []
    for it in range(7):
        cand = torch.randint(1, 600, (6, 5), device=DEV, generator=g, dtype=torch.int32)
        his = torch.randint(0, 600, (6, 50), device=DEV, generator=g, dtype=torch.int32)
        ids = torch.cat([his.reshape(-1), cand.reshape(-1)]).cpu().numpy()
        tickets.append(fa.run_host(ids))
        want.append(float(fb.run(cand, his)))
        rel = float((a.flat.grad - b.flat.grad).norm() / b.flat.grad.norm())
        assert rel < 1e-4, (it, rel)
        if it >= 1:
            assert fa.loss_of(tickets[it - 1]) == pytest.approx(want[it - 1], abs=1e-5)
    assert tickets == list(range(1, 8))
    assert fa.loss_of(tickets[-1]) == pytest.approx(want[-1], abs=1e-5)
