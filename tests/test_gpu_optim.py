"""Fused AdamW + EMA step (fit_b200.optim) against torch.optim.AdamW + the reference's EMA update on CPU."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _mk(sizes, seed):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(n, generator=g) * 0.05 for n in sizes]


@pytest.mark.parametrize("wd,every", [(0.0, 1), (0.01, 2)])
def test_adamw_ema_matches_torch(wd, every):
    from fit_b200.optim import FusedAdamWEMA
    from oracle.optim_oracle import AdamWEMAOracle
    sizes = [4, 7, 1152, 16384, 16385, 50001, 3 * 16384 + 12]   # ragged: below / at / across the 16384-element chunk size
    init = _mk(sizes, 0)
    params = [torch.nn.Parameter(p.clone().cuda()) for p in init]
    opt = FusedAdamWEMA(params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=wd, ema_decay=0.99, every_n_steps=every)
    ora = AdamWEMAOracle(init, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=wd, ema_decay=0.99, every_n_steps=every)
    for step in range(4):
        grads = _mk(sizes, 100 + step)
        for p, g in zip(params, grads):
            p.grad = g.clone().cuda()
        v0 = params[0]._version
        opt.step()
        ora.step(grads)
        assert params[0]._version > v0   # weight caches keyed by the version counter must notice the update
        for i, (p, q) in enumerate(zip(params, ora.params)):
            m, v = ora.state(i)
            # fp32 with a different contraction of the same formula: a few ulp
            torch.testing.assert_close(p.detach().cpu(), q.detach(), rtol=2e-6, atol=1e-8)
            torch.testing.assert_close(opt.exp_avg[i].cpu(), m, rtol=2e-6, atol=1e-9)
            torch.testing.assert_close(opt.exp_avg_sq[i].cpu(), v, rtol=2e-6, atol=1e-12)
            torch.testing.assert_close(opt.ema_params[i].cpu(), ora.ema[i], rtol=2e-6, atol=1e-8)


def test_no_ema_and_errors():
    from fit_b200.optim import FusedAdamWEMA
    p = torch.nn.Parameter(torch.ones(100, device="cuda"))
    opt = FusedAdamWEMA([p], ema_decay=None)
    with pytest.raises(RuntimeError):
        opt.step()                       # no gradient
    p.grad = torch.ones_like(p)
    opt.step()
    assert opt.ema_params == () and float(p[0]) < 1.0
    with pytest.raises(RuntimeError):
        FusedAdamWEMA([torch.nn.Parameter(torch.ones(4))])   # CPU parameter: no fallback
    with pytest.raises(ValueError):
        FusedAdamWEMA([p], ema_decay=1.5)


def test_swap_ema_weights_roundtrip():
    from fit_b200.optim import FusedAdamWEMA
    p = torch.nn.Parameter(torch.arange(10, dtype=torch.float32, device="cuda"))
    opt = FusedAdamWEMA([p], ema_decay=0.5)
    p.grad = torch.ones_like(p)
    opt.step()
    before, ema = p.detach().clone(), opt.ema_params[0].clone()
    with opt.swap_ema_weights_ctx():
        assert torch.equal(p.detach(), ema)
    assert torch.equal(p.detach(), before) and torch.equal(opt.ema_params[0], ema)


@pytest.mark.parametrize("number_of_tokens", [32, 96, 256])
def test_masked_subsample_matches_reference_collate(number_of_tokens):
    """masked_FiT subsample / collate on device (row f4) vs the CPU restatement, same RNG stream: bit-exact."""
    from fit_b200.masked import subsample_batch
    from oracle.subsample_oracle import collate
    g = torch.Generator().manual_seed(7)
    sizes = [256, 200, 96, 32, 31, 1, 160, 97]      # over-long, exact, shorter than every choice, ragged
    lat = [torch.randn(n, 16, generator=g) for n in sizes]
    pos = [torch.randn(n, 72, generator=g) for n in sizes]
    msk = [torch.ones(n, dtype=torch.bool) for n in sizes]
    msk[1][5] = False                                # mask bytes travel with their tokens
    want = collate(lat, pos, msk, number_of_tokens, generator=torch.Generator().manual_seed(11))
    got = subsample_batch(lat, pos, msk, number_of_tokens, "cuda", generator=torch.Generator().manual_seed(11))
    for w, o in zip(want, got):
        assert o.shape == w.shape and torch.equal(o.cpu(), w)


@pytest.mark.parametrize("layout", ["own", "lightning", "adamw"])
def test_optimizer_resume_matches_uninterrupted_run(layout):
    """save -> load -> step == an uninterrupted run (moments, step count for the bias correction, EMA copy), for the optimizer's own
    state_dict, the reference EMA optimizer's layout written by checkpoint.to_lightning_checkpoint (FiT/ema.py:322-347) and a bare
    torch.optim.AdamW state (per-parameter 'step' only)."""
    from fit_b200.checkpoint import to_lightning_checkpoint
    from fit_b200.optim import FusedAdamWEMA
    g = torch.Generator().manual_seed(3)
    shapes = [(40, 24), (17,), (3, 5, 7)]
    grads = [[torch.randn(s, generator=g).cuda() for s in shapes] for _ in range(5)]

    def make():
        gg = torch.Generator().manual_seed(9)
        return [torch.nn.Parameter(torch.randn(s, generator=gg).cuda()) for s in shapes]

    def run(opt, params, steps):
        for k in steps:
            for p, gr in zip(params, grads[k]):
                p.grad = gr.clone()
            opt.step()

    pa = make(); oa = FusedAdamWEMA(pa, lr=1e-2, ema_decay=0.9)
    run(oa, pa, range(5))
    pb = make(); ob = FusedAdamWEMA(pb, lr=1e-2, ema_decay=0.9)
    run(ob, pb, range(3))
    if layout == "own":
        state = ob.state_dict()
    elif layout == "lightning":
        mod = torch.nn.Module()
        for i, p in enumerate(pb):
            mod.register_parameter(f"p{i}", p)
        state = to_lightning_checkpoint(mod, "FiT-S/8", optimizer=ob, global_step=3)["optimizer_states"][0]
    else:
        state = {"state": {i: {"step": torch.tensor(3.0), "exp_avg": m.clone(), "exp_avg_sq": v.clone()}
                           for i, (m, v) in enumerate(zip(ob.exp_avg, ob.exp_avg_sq))},
                 "param_groups": [{"lr": 1e-2, "betas": (0.9, 0.999), "eps": 1e-8, "weight_decay": 0.0, "params": [0, 1, 2]}]}
    pc = [torch.nn.Parameter(p.detach().clone()) for p in pb]
    oc = FusedAdamWEMA(pc, lr=1e-2, ema_decay=0.9)
    if layout == "adamw":           # a bare AdamW state carries no EMA: continue from the interrupted run's copy
        for e, src in zip(oc.ema_params, ob.ema_params):
            e.copy_(src)
    oc.load_state_dict(state)
    assert oc.step_count == 3
    run(oc, pc, range(3, 5))
    for a, c in zip(pa, pc):
        assert torch.equal(a.detach(), c.detach())
    for a, c in zip(oa.ema_params, oc.ema_params):
        assert torch.equal(a, c)


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
@pytest.mark.parametrize("mlp_ratio", [4.0, 3.7])   # 3.7: hidden width 355 -> padded to 368 (the FiT-L case: 2730 -> 2736)
def test_optimizer_keeps_engine_weight_buffers_current(precision, mlp_ratio):
    """attach_optimizer: the step kernel writes the bf16 / packed copies of the updated parameters into the training engine's weight buffers
    (fitb200_adam_tensor.shadow).  After a step they are bit-identical to a fresh re-pack of the new parameters, the engine does not refill
    them, and a training step through them matches a model rebuilt from the same parameters."""
    from fit_b200.engine_train import TrainWeights
    from fit_b200.models.fit import FiT
    from fit_b200.optim import FusedAdamWEMA
    from fit_b200.train import _train_engine, attach_optimizer, masked_mse_loss
    from oracle import fit_oracle as fo
    from oracle import synth
    cfg = dict(patch_size=2, in_channels=4, hidden_size=144, depth=2, num_heads=2, mlp_ratio=mlp_ratio, class_dropout_prob=0.1, num_classes=10,
               learn_sigma=False)
    sd = fo.random_state_dict(cfg, seed=3)
    m = FiT(**cfg, precision=precision)
    m.load_state_dict(sd)
    m = m.cuda().train()
    m.y_embedder.dropout_prob = 0.0
    x, t, y, pos, mask, target = synth.mixed_batch(cfg, [(16, 16), (8, 32)], 64, seed=5, rope_dim=72)
    opt = attach_optimizer(m, FusedAdamWEMA(m.parameters(), lr=1e-2, ema_decay=0.99))
    eng = _train_engine(m)
    refills = []
    orig = TrainWeights.refresh
    TrainWeights.refresh = lambda self, model: (refills.append(1), orig(self, model))[1]
    try:
        for _ in range(2):
            for p in m.parameters():
                p.grad = None
            loss = masked_mse_loss(m(x.cuda(), t.cuda(), y.cuda(), pos.cuda(), mask.cuda(), train=True), target.cuda(), mask.cuda())
            loss.backward()
            opt.step()
        W = eng.weights()
        assert not refills, "the engine re-packed its weights although the optimizer keeps them current"
    finally:
        TrainWeights.refresh = orig
    fresh = TrainWeights(m, precision)
    assert torch.equal(W.ada_w, fresh.ada_w) and torch.equal(W.ada_b, fresh.ada_b) and torch.equal(W.fin_w, fresh.fin_w)
    for a, b in zip(W.blocks, fresh.blocks):
        for k in ("qkv_w", "proj_w", "fc1_w", "fc1_b", "fc2_w", "qkv_b", "proj_b", "fc2_b"):
            assert torch.equal(a[k], b[k]), k
    # a parameter change behind the optimizer's back (load_state_dict) still refills
    m.load_state_dict({k: v.cuda() for k, v in sd.items()})
    W2 = eng.weights()
    assert W2 is W and torch.equal(W2.blocks[0]["qkv_w"].float(), m.blocks[0].attn.qkv.weight.detach().to(W2.wdt).float())


def test_capturable_optimizer_matches_host_scalar_path():
    """capturable=True (step number and hyper-parameters in device memory, bias corrections formed in the kernel) against the host-scalar
    path over several steps, incl. the every_n_steps EMA rule and an lr change through sync_device_state()."""
    from fit_b200.optim import FusedAdamWEMA
    sizes = [5, 1152, 16385, 40000]
    init = _mk(sizes, 1)
    pa = [torch.nn.Parameter(p.clone().cuda()) for p in init]
    pb = [torch.nn.Parameter(p.clone().cuda()) for p in init]
    a = FusedAdamWEMA(pa, lr=1e-3, weight_decay=0.01, ema_decay=0.99, every_n_steps=2)
    b = FusedAdamWEMA(pb, lr=1e-3, weight_decay=0.01, ema_decay=0.99, every_n_steps=2)
    b.capturable = True
    for step in range(5):
        if step == 3:
            a.lr = b.lr = 5e-4
            b.sync_device_state()
        grads = _mk(sizes, 200 + step)
        for p, q, g in zip(pa, pb, grads):
            p.grad, q.grad = g.clone().cuda(), g.clone().cuda()
        a.step(); b.step()
        for i in range(len(sizes)):
            torch.testing.assert_close(pb[i].detach(), pa[i].detach(), rtol=1e-6, atol=1e-9)
            torch.testing.assert_close(b.ema_params[i], a.ema_params[i], rtol=1e-6, atol=1e-9)
    assert a.step_count == b.step_count == 5 and int(b._step_dev) == 5


def test_captured_train_step_matches_eager_steps():
    """CapturedTrainStep (forward + loss + backward + AdamW/EMA as one replayed CUDA graph) against the same steps issued eagerly: same
    parameters and EMA after 4 steps on changing batches (atomically accumulated bias / gate gradients make the two runs differ by fp32
    summation order only), ONE capture, and the parameter version counters advance so that the sampling engine repacks."""
    from fit_b200.models.fit import FiT
    from fit_b200.optim import FusedAdamWEMA
    from fit_b200.train import CapturedTrainStep, attach_optimizer, masked_mse_loss
    from oracle import fit_oracle as fo
    from oracle import synth
    cfg = dict(patch_size=2, in_channels=4, hidden_size=144, depth=2, num_heads=2, mlp_ratio=4.0, class_dropout_prob=0.1, num_classes=10,
               learn_sigma=False)
    sd = fo.random_state_dict(cfg, seed=3)
    models = []
    for _ in range(2):
        m = FiT(**cfg, precision="bf16")
        m.load_state_dict(sd)
        m = m.cuda().train()
        m.y_embedder.dropout_prob = 0.0
        models.append(m)
    eager, graphed = models
    oe = attach_optimizer(eager, FusedAdamWEMA(eager.parameters(), lr=1e-5, ema_decay=0.99))
    og = FusedAdamWEMA(graphed.parameters(), lr=1e-5, ema_decay=0.99)
    step = CapturedTrainStep(graphed, og)
    losses = []
    for k in range(4):
        x, t, y, pos, mask, target = [b.cuda() for b in synth.mixed_batch(cfg, [(16, 16), (8, 32)], 64, seed=20 + k, rope_dim=72)]
        for p in eager.parameters():
            p.grad = None
        le = masked_mse_loss(eager(x, t, y, pos, mask, train=True), target, mask)
        le.backward()
        oe.step()
        v0 = next(graphed.parameters())._version
        lg = step(x, t, y, pos, mask, target)
        losses.append((float(le), float(lg)))
        assert next(graphed.parameters())._version > v0
    assert step.captures == 1 and og.step_count == 4
    for le, lg in losses:
        assert abs(le - lg) < 2e-3 * abs(le)
    # Adam divides by sqrt(v): an element whose gradient is rounding noise (e.g. the key bias, whose gradient is analytically zero) moves by
    # +-lr per step in either run, so parameters are compared against that bound and the FIRST MOMENTS (linear in the gradients) tightly
    lr, steps = og.lr, 4
    for i, ((n, p), q) in enumerate(zip(eager.named_parameters(), graphed.parameters())):
        assert float((q.detach() - p.detach()).abs().max()) <= 2.0 * steps * lr * 1.01, n
        me, mg = oe.exp_avg[i].double(), og.exp_avg[i].double()
        den = float(me.norm()) + 1e-7 * me.numel() ** 0.5
        assert float((mg - me).norm()) / den < 5e-2, n
    for e1, e2 in zip(oe.ema_params, og.ema_params):
        assert float((e2 - e1).abs().max()) <= 2.0 * steps * lr * 1.01
