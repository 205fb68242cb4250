"""TEST INFRASTRUCTURE ONLY (never imported by the product package).

CPU oracle for one Listener layer's recurrence, reference src/asr.py:466-501 (RNNLayer.forward):

    input_x = pack_padded_sequence(input_x, state_len, batch_first=True)
    output, hidden = self.layer(input_x, state)            # nn.LSTM(in, H, bidirectional, batch_first)
    output, state_len = pad_packed_sequence(output, batch_first=True)

The arithmetic lives in torch (nn.LSTM, unpinned upstream; 2.11.0 here), so `reference_call` executes that exact call
sequence on CPU in the requested precision, and `restated` is an independent frame-by-frame restatement of the
published LSTM equations (i,f,g,o gate order, c = f*c + i*g, h = o*tanh(c)) that the CPU tests check it against.
"""
import torch
from torch.nn.utils.rnn import pack_padded_sequence, pad_packed_sequence


def make_layer(n_in, H, bidir, seed, dtype=torch.float64):
    g = torch.Generator().manual_seed(seed)
    layer = torch.nn.LSTM(n_in, H, bidirectional=bidir, num_layers=1, batch_first=True).to(dtype)
    with torch.no_grad():
        for p in layer.parameters():
            p.copy_((torch.rand(p.shape, generator=g, dtype=torch.float64) * 2 - 1) * H ** -0.5)
    return layer


def reference_call(layer, x, lens):
    """asr.py:478-485 verbatim on a CPU layer; lens sorted descending (enforce_sorted default)."""
    packed = pack_padded_sequence(x, lens, batch_first=True)
    out, _ = layer(packed, None)
    out, _ = pad_packed_sequence(out, batch_first=True)
    return out


def restated(layer, x, lens):
    """Independent restatement: explicit loops over rows, directions and frames."""
    B, T, _ = x.shape
    H = layer.hidden_size
    nd = 2 if layer.bidirectional else 1
    tmax = max(lens)
    out = x.new_zeros(B, tmax, nd * H)
    for d in range(nd):
        sfx = "_l0" + ("_reverse" if d else "")
        w_ih, w_hh = getattr(layer, "weight_ih" + sfx), getattr(layer, "weight_hh" + sfx)
        b = getattr(layer, "bias_ih" + sfx) + getattr(layer, "bias_hh" + sfx)
        rows = []
        for r in range(B):
            h = x.new_zeros(H)
            c = x.new_zeros(H)
            frames = range(lens[r] - 1, -1, -1) if d else range(lens[r])
            hs = {}
            for t in frames:
                z = w_ih @ x[r, t] + w_hh @ h + b
                i, f, g, o = torch.sigmoid(z[:H]), torch.sigmoid(z[H:2 * H]), torch.tanh(z[2 * H:3 * H]), torch.sigmoid(z[3 * H:])
                c = f * c + i * g
                h = o * torch.tanh(c)
                hs[t] = h
            row = [hs[t] if t in hs else x.new_zeros(H) for t in range(tmax)]
            rows.append(torch.stack(row))
        out[:, :, d * H:(d + 1) * H] = torch.stack(rows)
    return out
