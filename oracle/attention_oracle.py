"""CPU oracle for kernel (2): the per-decoder-step attention of the reference.

TEST INFRASTRUCTURE ONLY -- never imported by the product package.

Plain-PyTorch (CPU, fp32 or fp64) functional restatement of
`Attention.forward` in the reference (src/asr.py:410-462) and of the state it
caches between steps (src/asr.py:404-408, 412-419, 444-449).  Written as pure
functions over explicit weights so the same tensors can be fed to the CUDA
path.  Pinned against the real `src.asr.Attention` by tests/test_oracle_pins.py
(in the build container) and by the committed goldens in tests/golden/.

Weight naming follows the reference's parameter names (asr.py:383-400):
  psi.weight [A,D] psi.bias [A]     key projection        (asr.py:384,419)
  phi.weight [A,dec]                query projection      (asr.py:383,422)
  loc_conv.weight [C,1,2K+1]        location conv, pad K  (asr.py:398)
  loc_proj.weight [A,C]             location projection   (asr.py:399)
  gen_energy.weight [1,A] .bias [1] energy                (asr.py:400)
"""
import torch
import torch.nn.functional as F


def pad_mask(lens, t_max, device="cpu"):
    """True where frame t is padding: t >= len_b  (asr.py:415-418)."""
    ar = torch.arange(t_max, device=device)[None, :]
    return ar >= torch.as_tensor(list(lens), device=device)[:, None]


def project_keys(h, psi_w, psi_b):
    """K = tanh(h psi^T + b), computed once per batch (asr.py:419)."""
    return torch.tanh(F.linear(h, psi_w, psi_b))


def uniform_alignment(lens, t_max, dtype=torch.float32):
    """First-step 'previous alignment': 1/len_b on valid frames (asr.py:444-449)."""
    a = torch.zeros(len(lens), t_max, dtype=dtype)
    for b, n in enumerate(lens):
        a[b, :n] = 1.0 / n
    return a


def loc_step(s, h, keys, lens, prev_att, w, scale=2.0):
    """One location-aware step (asr.py:422, 451-457).

    s [B,dec], h [B,T,D], keys [B,T,A], prev_att [B,T] -> (alpha [B,T], ctx [B,D])
    w: dict with phi, loc_conv, loc_proj, gen_w, gen_b.
    """
    q = torch.tanh(F.linear(s, w["phi"]))                                   # asr.py:422
    half = (w["loc_conv"].shape[-1] - 1) // 2
    f = F.conv1d(prev_att[:, None, :], w["loc_conv"], padding=half)          # [B,C,T] asr.py:452
    loc = torch.tanh(F.linear(f.transpose(1, 2), w["loc_proj"]))             # [B,T,A]
    e = F.linear(torch.tanh(keys + q[:, None, :] + loc), w["gen_w"], w["gen_b"]).squeeze(-1)  # asr.py:453
    e = e.masked_fill(pad_mask(lens, h.shape[1], h.device), float("-inf"))   # asr.py:454
    alpha = torch.softmax(e * scale, dim=-1)                                  # asr.py:455
    ctx = torch.bmm(alpha[:, None, :], h).squeeze(1)                         # asr.py:457
    return alpha, ctx


def dot_step(s, h, keys, lens, w, scale=2.0, proj=True):
    """One dot-product step (asr.py:422, 426-432)."""
    q = torch.tanh(F.linear(s, w["phi"])) if proj else s
    e = torch.bmm(keys, q[:, :, None]).squeeze(-1)
    e = e.masked_fill(pad_mask(lens, h.shape[1], h.device), float("-inf"))
    alpha = torch.softmax(e * scale, dim=-1)
    ctx = torch.bmm(alpha[:, None, :], h).squeeze(1)
    return alpha, ctx


def weights_from_module(att):
    """Pull the weight dict out of a reference-shaped Attention module (ours or upstream's)."""
    w = {"phi": att.phi.weight.detach()}
    if getattr(att, "mode", "dot") == "loc":
        w.update(loc_conv=att.loc_conv.weight.detach(), loc_proj=att.loc_proj.weight.detach(),
                 gen_w=att.gen_energy.weight.detach(), gen_b=att.gen_energy.bias.detach())
    w.update(psi_w=att.psi.weight.detach(), psi_b=att.psi.bias.detach())
    return w
