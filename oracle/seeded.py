"""TEST INFRASTRUCTURE ONLY: deterministic parameter values and gradient sub-samples shared by tools/make_golden.py
(which applies them to the UNMODIFIED reference model) and the GPU tests (which apply them to the B200 mirror).

The mid-size fixtures (tests/golden/midsize_*.pt) hold what the reference computed for a model of the shipped layer
widths (attention 300, decoder 320, encoder 2x320 bidirectional): 3.6 M parameters.  Storing the weights and every
gradient would put tens of MB per fixture into the history, so the weights are a pure function of (parameter name,
shape, seed) that both sides evaluate -- parameter names are the same on both sides, state_dicts interchange -- and
gradients of tensors above `FULL_LIMIT` elements are stored as a fixed pseudo-random sample of `SAMPLE` entries plus
their float64 sum and Euclidean norm over ALL entries.
"""
import zlib

import torch

FULL_LIMIT = 1 << 16
SAMPLE = 1 << 14


def _gen(name, seed):
    return torch.Generator().manual_seed((zlib.crc32(name.encode()) ^ (seed * 0x9E3779B1)) & 0x7FFFFFFF)


def reseed_parameters(model, seed):
    """p <- N(0,1)/sqrt(fan_in) for matrices / conv kernels, 0.1*N(0,1) for vectors; in place, by parameter name."""
    with torch.no_grad():
        for name, p in model.named_parameters():
            g = _gen(name, seed)
            v = torch.randn(p.shape, generator=g, dtype=torch.float32)
            if p.dim() >= 2:
                fan_in = 1
                for d in p.shape[1:]:
                    fan_in *= int(d)
                v = v / float(fan_in) ** 0.5
            else:
                v = v * 0.1
            p.copy_(v.to(p.device, p.dtype))


def sample_index(name, numel, seed):
    return torch.randperm(numel, generator=_gen("idx/" + name, seed))[:SAMPLE]


def summarise_gradient(name, grad, seed):
    """-> what the fixture stores for one parameter gradient."""
    g = grad.detach().cpu().reshape(-1)
    g64 = g.double()
    out = {"numel": g.numel(), "sum": float(g64.sum()), "norm": float(g64.norm()), "absmax": float(g.abs().max())}
    if g.numel() <= FULL_LIMIT:
        out["full"] = g.clone().reshape(grad.shape)
    else:
        out["sample"] = g[sample_index(name, g.numel(), seed)].clone()
    return out


def compare_gradient(name, grad, ref, seed, rel=1e-4):
    """max |ours - reference| over the stored entries <= rel * the reference tensor's largest entry (most entries
    are ~0), and the whole-tensor sum / norm agree likewise.  Returns None or a failure message."""
    g = grad.detach().cpu().reshape(-1)
    tol = rel * ref["absmax"] + 1e-7
    if "full" in ref:
        err = float((g - ref["full"].reshape(-1)).abs().max())
    else:
        err = float((g[sample_index(name, g.numel(), seed)] - ref["sample"]).abs().max())
    if err > tol:
        return "%s: max abs error %.3e > %.3e" % (name, err, tol)
    norm = float(g.double().norm())
    if abs(norm - ref["norm"]) > rel * ref["norm"] + 1e-9:
        return "%s: norm %.6e vs %.6e" % (name, norm, ref["norm"])
    s = float(g.double().sum())
    if abs(s - ref["sum"]) > rel * ref["absmax"] * ref["numel"] ** 0.5 + 1e-7:
        return "%s: sum %.6e vs %.6e" % (name, s, ref["sum"])
    return None
