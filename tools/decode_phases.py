"""Wall-clock split of one M4 beam decode: encoder + CTC posteriors + keys | search set-up | the decode steps | results."""
import importlib, math, os, sys, time
import torch
import torch.nn.functional as F
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
PKG = "end-to-end-asr-pytorch_b200"
asr, rnnlm, syn, beam = (importlib.import_module(PKG + "." + m) for m in ("asr", "rnnlm", "synthetic", "beam"))
DEC = bench.DEC
w = syn.WORKLOADS[DEC["workload"]]
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = asr.Seq2Seq(torch.zeros(1, 8, w["d_in"]), w["vocab"], w["cfg"]).to(dev).eval()
model.joint_ctc, model.ctc_weight, model.decode_lm_weight = True, DEC["ctc_weight"], DEC["lm_weight"]
model.rnn_lm = rnnlm.RNN_LM(DEC["lm"][0], DEC["lm"][1], w["vocab"], layers=DEC["lm"][2]).to(dev).eval()
x = torch.randn(1, DEC["frames"], w["d_in"], generator=torch.Generator().manual_seed(1)).to(dev)
steps = int(math.ceil(DEC["frames"] * DEC["ratio"]))
n = DEC["beam"]


def sync():
    torch.cuda.synchronize()
    return time.perf_counter()


with torch.no_grad():
    for _ in range(2):
        model.beam_decode(x, steps, [DEC["frames"]], n)
    acc = [0.0] * 4
    R = 5
    for _ in range(R):
        t0 = sync()
        h, enc_len = model.encoder(x, [DEC["frames"]])
        x_ctc = F.log_softmax(model.ctc_layer(h), dim=-1)[0]
        h = h.contiguous()
        keys = model.attention.project_keys(h).contiguous()
        t1 = sync()
        bs = beam.BeamSearch(model.char_dim, n, steps, dev, x_ctc=x_ctc, ctc_weight=DEC["ctc_weight"], lm_weight=DEC["lm_weight"])
        t2 = sync()
        beam._native_decode(model, model.rnn_lm, bs, h, keys, int(enc_len[0]), steps)
        t3 = sync()
        out = bs.results(emb=model.embed)
        t4 = sync()
        for i, d in enumerate((t1 - t0, t2 - t1, t3 - t2, t4 - t3)):
            acc[i] += d
    names = ["encoder + ctc posteriors + keys", "BeamSearch set-up", "decode steps (b2asr_decode_run)", "results (D2H + Hypothesis list)"]
    for nm, v in zip(names, acc):
        print("  %-34s %7.2f ms" % (nm, v / R * 1e3))
    print("  per decode step %.1f us" % (acc[2] / R / steps * 1e6))
