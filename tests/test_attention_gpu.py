"""GPU parity: kernel (2b) attention step vs oracle/attention_oracle.py (restating asr.py:410-462).

Tolerance (north_star: attention outputs within 1e-4 relative in fp32):
  |a-b| <= 1e-4*|b| + 1e-6 for alpha (values in [0,1]) and 1e-4*|b| + 1e-5*max|ctx| for the context,
checked against the oracle evaluated in float64.
"""
import importlib

import pytest
import torch

from oracle import attention_oracle as ao

pytestmark = pytest.mark.gpu


def _attn():
    return importlib.import_module("end-to-end-asr-pytorch_b200._attn")


def make_weights(A, D, dec, C=10, Kw=100, seed=0):
    g = torch.Generator().manual_seed(seed)

    def n(*s, sd=1.0):
        return torch.randn(*s, generator=g) * sd

    return dict(phi=n(A, dec, sd=dec ** -0.5), psi_w=n(A, D, sd=D ** -0.5), psi_b=n(A, sd=0.1),
                loc_conv=n(C, 1, 2 * Kw + 1, sd=(2 * Kw + 1) ** -0.5), loc_proj=n(A, C, sd=C ** -0.5),
                gen_w=n(1, A, sd=A ** -0.5), gen_b=n(1, sd=0.1))


def make_batch(B, T, D, dec, seed=1, nstates=4):
    g = torch.Generator().manual_seed(seed)
    h = torch.randn(B, T, D, generator=g)
    lens = sorted([int(v) for v in torch.randint(max(1, T // 2), T + 1, (B,), generator=g)], reverse=True)
    lens[0] = T
    for b, n in enumerate(lens):
        h[b, n:] = 0
    s = [torch.randn(B, dec, generator=g) for _ in range(nstates)]
    return h, lens, s


def worst(a, b, rtol, atol):
    err = (a.double().cpu() - b.double()).abs()
    tol = rtol * b.double().abs() + atol
    return float((err / tol).max())


SHAPES = [  # B, T, D, A, dec
    (4, 50, 64, 32, 48),
    (16, 97, 512, 256, 256),     # timit-shaped (M1)
    (32, 400, 640, 300, 320),    # libri-shaped (M2)
    (3, 750, 640, 300, 320),     # longest utterance after 4x subsampling
]


@pytest.mark.parametrize("B,T,D,A,dec", SHAPES)
@pytest.mark.parametrize("cluster", [0, 1, 2])
def test_loc_steps_match_oracle(cuda, B, T, D, A, dec, cluster):
    at = _attn()
    w = make_weights(A, D, dec)
    h, lens, states = make_batch(B, T, D, dec)
    keys64 = ao.project_keys(h.double(), w["psi_w"].double(), w["psi_b"].double())
    w64 = {k: v.double() for k, v in w.items()}
    prev64 = ao.uniform_alignment(lens, T, torch.float64)
    wd = {k: v.to(cuda) for k, v in w.items()}
    hd, kd = h.to(cuda), keys64.float().to(cuda)
    ld = at.lens_i32(lens, cuda)
    prev = None
    for step, s in enumerate(states):
        a64, c64 = ao.loc_step(s.double(), h.double(), keys64.float().double(), lens, prev64, w64)
        att, ctx, q = at.step_forward("loc", s.to(cuda), wd["phi"], kd, hd, ld, prev, wd["loc_conv"], wd["loc_proj"],
                                      wd["gen_w"], wd["gen_b"], 2.0, cluster=cluster, want_q=True)
        torch.cuda.synchronize()
        wa = worst(att, a64, 1e-4, 1e-6)
        wc = worst(ctx, c64, 1e-4, 1e-5 * float(c64.abs().max()))
        wq = worst(q, torch.tanh(s.double() @ w64["phi"].t()), 1e-4, 1e-6)
        assert wa <= 1.0 and wc <= 1.0 and wq <= 1.0, "step %d: alpha %.2f ctx %.2f q %.2f (err/tol)" % (step, wa, wc, wq)
        for b, n in enumerate(lens):
            assert float(att[b, n:].abs().sum()) == 0.0
        prev = att                   # feed our own alignment back, as Attention.prev_att does (asr.py:456)
        prev64 = att.double().cpu()  # and keep the oracle on the same trajectory


@pytest.mark.parametrize("B,T,D,A,dec", [(4, 50, 64, 32, 48), (8, 400, 640, 300, 320)])
def test_loc_chain_of_20_steps_against_free_running_oracle(cuda, B, T, D, A, dec):
    """Error growth through prev_att (asr.py:456): kernel and float64 oracle each feed back THEIR OWN alignment for
    20 steps -- nothing re-synchronises them -- and the last step's alpha / ctx still meet the 1e-4 bar."""
    at = _attn()
    w = make_weights(A, D, dec)
    h, lens, states = make_batch(B, T, D, dec, seed=21, nstates=20)
    keys64 = ao.project_keys(h.double(), w["psi_w"].double(), w["psi_b"].double())
    w64 = {k: v.double() for k, v in w.items()}
    wd = {k: v.to(cuda) for k, v in w.items()}
    hd, kd = h.to(cuda), keys64.float().to(cuda)
    ld = at.lens_i32(lens, cuda)
    prev, prev64 = None, ao.uniform_alignment(lens, T, torch.float64)
    worst_a = worst_c = 0.0
    for s in states:
        a64, c64 = ao.loc_step(s.double(), h.double(), keys64.float().double(), lens, prev64, w64)
        att, ctx, _ = at.step_forward("loc", s.to(cuda), wd["phi"], kd, hd, ld, prev, wd["loc_conv"], wd["loc_proj"],
                                      wd["gen_w"], wd["gen_b"], 2.0)
        worst_a = max(worst_a, worst(att, a64, 1e-4, 1e-6))
        worst_c = max(worst_c, worst(ctx, c64, 1e-4, 1e-5 * float(c64.abs().max())))
        prev, prev64 = att, a64                      # two independent trajectories
    assert worst_a <= 1.0 and worst_c <= 1.0, "over 20 free-running steps: alpha %.2f ctx %.2f (err/tol)" % (worst_a, worst_c)


@pytest.mark.parametrize("B,T,D,A,dec", SHAPES[:3])
def test_dot_steps_match_oracle(cuda, B, T, D, A, dec):
    at = _attn()
    w = make_weights(A, D, dec)
    h, lens, states = make_batch(B, T, D, dec, seed=5)
    keys64 = ao.project_keys(h.double(), w["psi_w"].double(), w["psi_b"].double())
    ld = at.lens_i32(lens, cuda)
    for s in states[:2]:
        a64, c64 = ao.dot_step(s.double(), h.double(), keys64.float().double(), lens, {"phi": w["phi"].double()})
        att, ctx, _ = at.step_forward("dot", s.to(cuda), w["phi"].to(cuda), keys64.float().to(cuda), h.to(cuda), ld,
                                      None, None, None, None, None, 2.0)
        wa = worst(att, a64, 1e-4, 1e-6)
        wc = worst(ctx, c64, 1e-4, 1e-5 * float(c64.abs().max()))
        assert wa <= 1.0 and wc <= 1.0, "alpha %.2f ctx %.2f (err/tol)" % (wa, wc)


def test_shared_memory_rows_for_beam_decode(cuda):
    """20 hypotheses attending over one utterance (batch stride 0) == the same utterance repeated."""
    at = _attn()
    B, T, D, A, dec = 20, 400, 640, 300, 320
    w = make_weights(A, D, dec)
    g = torch.Generator().manual_seed(3)
    h1 = torch.randn(1, T, D, generator=g)
    s = torch.randn(B, dec, generator=g)
    prev = torch.softmax(torch.randn(B, T, generator=g), -1)
    keys1 = ao.project_keys(h1, w["psi_w"], w["psi_b"])
    a_ref, c_ref = ao.loc_step(s.double(), h1.double().expand(B, T, D), keys1.double().expand(B, T, A), [T] * B,
                               prev.double(), {k: v.double() for k, v in w.items()})
    wd = {k: v.to(cuda) for k, v in w.items()}
    att, ctx, _ = at.step_forward("loc", s.to(cuda), wd["phi"], keys1.to(cuda), h1.to(cuda), at.lens_i32([T] * B, cuda),
                                  prev.to(cuda), wd["loc_conv"], wd["loc_proj"], wd["gen_w"], wd["gen_b"], 2.0,
                                  shared_memory=True)
    wa = worst(att, a_ref, 1e-4, 1e-6)
    wc = worst(ctx, c_ref, 1e-4, 1e-5 * float(c_ref.abs().max()))
    assert wa <= 1.0 and wc <= 1.0, "alpha %.2f ctx %.2f (err/tol)" % (wa, wc)


def rel(a, b):
    """max |a-b| / max|b| : gradients are compared relative to the tensor's own scale."""
    return float((a.double().cpu() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("B,T,D,A,dec", SHAPES)
@pytest.mark.parametrize("mode", ["loc", "dot"])
@pytest.mark.parametrize("cluster", [0, 1])
def test_step_backward_matches_autograd_of_oracle(cuda, B, T, D, A, dec, mode, cluster):
    """Kernel backward vs float64 autograd through the oracle step; every gradient within 1e-4 of its
    tensor's max magnitude."""
    at = _attn()
    w = make_weights(A, D, dec)
    h, lens, states = make_batch(B, T, D, dec, seed=11, nstates=1)
    g = torch.Generator().manual_seed(99)
    keys = ao.project_keys(h, w["psi_w"], w["psi_b"])
    prev = torch.softmax(torch.randn(B, T, generator=g), -1) * ao.pad_mask(lens, T).logical_not()
    prev = prev / prev.sum(-1, keepdim=True)
    Ga, Gc = torch.randn(B, T, generator=g), torch.randn(B, D, generator=g)
    # ---- oracle, float64 autograd
    leaf = {k: v.double().clone().requires_grad_(True) for k, v in w.items() if k not in ("psi_w", "psi_b")}
    s64 = states[0].double().clone().requires_grad_(True)
    k64 = keys.double().clone().requires_grad_(True)
    p64 = prev.double().clone().requires_grad_(True)
    if mode == "loc":
        a64, c64 = ao.loc_step(s64, h.double(), k64, lens, p64, leaf)
    else:
        a64, c64 = ao.dot_step(s64, h.double(), k64, lens, leaf)
    mask = ao.pad_mask(lens, T).logical_not().double()
    ((a64 * Ga.double() * mask).sum() + (c64 * Gc.double()).sum()).backward()
    # ---- kernel
    wd = {k: v.to(cuda) for k, v in w.items()}
    ld = at.lens_i32(lens, cuda)
    loc = mode == "loc"
    att, ctx, q = at.step_forward(mode, states[0].to(cuda), wd["phi"], keys.to(cuda), h.to(cuda), ld,
                                  prev.to(cuda) if loc else None, wd["loc_conv"] if loc else None,
                                  wd["loc_proj"] if loc else None, wd["gen_w"] if loc else None,
                                  wd["gen_b"] if loc else None, 2.0, cluster=cluster, want_q=True)
    C, Kw = (10, 100) if loc else (0, 0)
    o_wl, o_wc, o_wg, o_bg, stride = at.pgrad_layout(A, C, Kw) if loc else (0, 0, 0, 0, 1)
    slots = at.num_slots(B, T, D, A, dec, mode, C, Kw, cluster)
    pgrad = torch.zeros(slots, stride, device=cuda) if loc else None
    d_keys = torch.zeros(B, T, A, device=cuda)
    d_prev, d_zq, d_s = at.step_backward(mode, wd["phi"], keys.to(cuda), h.to(cuda), ld, prev.to(cuda) if loc else None,
                                         wd["loc_conv"] if loc else None, wd["loc_proj"] if loc else None,
                                         wd["gen_w"] if loc else None, 2.0, q, att, Gc.to(cuda), Ga.to(cuda),
                                         d_keys, pgrad, dec, cluster=cluster)
    torch.cuda.synchronize()
    errs = {"d_s": rel(d_s, s64.grad), "d_keys": rel(d_keys, k64.grad),
            "d_phi": rel(d_zq.t() @ states[0].to(cuda), leaf["phi"].grad)}
    if loc:
        pg = pgrad.sum(0)
        errs["d_prev"] = rel(d_prev, p64.grad * mask)
        errs["d_loc_proj"] = rel(pg[o_wl:o_wc].view(A, C), leaf["loc_proj"].grad)
        errs["d_loc_conv"] = rel(pg[o_wc:o_wg].view(C, 1, 2 * Kw + 1), leaf["loc_conv"].grad)
        errs["d_gen_w"] = rel(pg[o_wg:o_bg].view(1, A), leaf["gen_w"].grad)
        # d/d(gen_energy.bias) of a softmax is exactly 0 (shift invariance): absolute check
        errs["d_gen_b"] = float(pg[o_bg].abs().cpu()) / float(Ga.abs().max()) * 1e-1
    bad = {k: v for k, v in errs.items() if not v <= 1e-4}
    assert not bad, "gradient mismatch (err / max|ref|): %s" % errs
