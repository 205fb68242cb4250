"""Time one Listener layer (forward, backward) on the persistent kernels against the library's packed path.
Usage: python tools/enc_bench.py [B] ; shapes are workload M2's five layers."""
import importlib
import os
import sys

import torch
from torch.nn.utils.rnn import pack_padded_sequence, pad_packed_sequence

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
PKG = "end-to-end-asr-pytorch_b200"
enc = importlib.import_module(PKG + "._enc")
lib = importlib.import_module(PKG + "._lib")


def timed(fn, n=5):
    fn(); fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    os.environ.setdefault("B2ASR_ENC_DEBUG", "1")
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    if len(sys.argv) > 2:
        lib.lib().b2asr_lstm_seq_path(1 if sys.argv[2] == "l2" else 0)
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    tot = {"native": 0.0, "cudnn": 0.0}
    for name, T, n_in, H in (("layer0", 1600, 120, 320), ("layer1", 800, 640, 320), ("layer2-4", 400, 640, 320)):
        lens = sorted(torch.randint(int(T * 0.625), T + 1, (B,)).tolist(), reverse=True)
        lens[0] = T
        layer = torch.nn.LSTM(n_in, H, bidirectional=True, batch_first=True).to(dev)
        x = torch.randn(B, T, n_in, device=dev)
        for b, n in enumerate(lens):
            x[b, n:] = 0
        x.requires_grad_(True)
        w = torch.randn(B, T, 2 * H, device=dev)

        def native():
            out = enc.lstm_layer(layer, x, lens)
            (out * w).sum().backward()

        def cudnn():
            out, _ = layer(pack_padded_sequence(x, lens, batch_first=True))
            out, _ = pad_packed_sequence(out, batch_first=True)
            (out * w).sum().backward()

        def native_fwd():
            with torch.no_grad():
                enc.lstm_layer(layer, x, lens)

        def cudnn_fwd():
            with torch.no_grad():
                pad_packed_sequence(layer(pack_padded_sequence(x, lens, batch_first=True))[0], batch_first=True)

        enc.status_word(dev).zero_()
        lib.profile_start()
        native()
        prof = lib.profile_stop()
        polls = enc.status_word(dev).tolist()
        a, b = timed(native), timed(cudnn)
        af, bf = timed(native_fwd), timed(cudnn_fwd)
        enc.check_status(dev)
        mult = 3 if name == "layer2-4" else 1
        tot["native"] += a * mult
        tot["cudnn"] += b * mult
        print("%-9s T=%4d in=%3d  fwd+bwd native %7.2f ms  cudnn %7.2f ms | fwd only native %6.2f  cudnn %6.2f | recurrence kernels: %s failed polls per CTA-frame fwd %.2f bwd %.2f"
              % (name, T, n_in, a, b, af, bf, {k: "%.2f ms" % v[1] for k, v in prof.items()},
                 polls[1] / (256.0 * T), polls[2] / (256.0 * 4 * T)), flush=True)
        print("          CTA 0 us per frame {exchange, MACs, rest}: fwd %s  bwd %s" % (
            ["%.2f" % (c / 1965.0 / T) for c in polls[4:7]], ["%.2f" % (c / 1965.0 / T) for c in polls[8:11]]), flush=True)
        print("          rest of the last kernel that wrote it (bwd: recv sum+cell, dG store+row max, global stores, barrier): %s" % ["%.2f" % (c / 1965.0 / T) for c in polls[11:15]], flush=True)
    print("five layers: native %.1f ms, cudnn %.1f ms" % (tot["native"], tot["cudnn"]))


if __name__ == "__main__":
    main()
