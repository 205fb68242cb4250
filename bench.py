#!/usr/bin/env python
"""Benchmark of the hybrid CTC/attention train step (BASELINE.json metric: utterances/s).

    python bench.py --gpus N --steps K --warmup W            # this repo's sm_100a path
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's own CPU path (oracle/_ref)

A step = one pass of Trainer.exec's body (reference src/solver.py:132-182: forward, CE + CTC loss,
backward, [gradient all-reduce], clip, NaN guard, optimizer step) over one synthetic
libri_example.yaml-shaped batch (workload M2: B=32/GPU, T<=1600 frames of 120-dim
fbank+delta, V=31 chars, pBLSTM 5x320 -> T'=400, location attention 300, decoder 320, joint_ctc 0.5,
tf_rate 1.0, Adadelta).  Prints ONE JSON line (rank 0).
"""
import argparse
import gc
import importlib
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PKG = "end-to-end-asr-pytorch_b200"
METRIC = "utterances/sec hybrid CTC-attn train step"
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the kernel that dominates each call (profiles/, M2 shape)
NCU_DRAM_BYTES = {
    "b2asr_decoder_loop_bwd": (69.701e6 + 0.915e6, "profiles/r02_ncu_kernels.txt attn_step_bwd_kernel"),
    "b2asr_decoder_loop_fwd": (48.698e6 + 0.013e6, "profiles/r02_ncu_kernels.txt attn_step_fwd_kernel"),
}
WORKLOAD = "libri_m2"
N_BATCHES = 4                      # distinct synthetic batches rotated through the timed steps


def pkg(name):
    return importlib.import_module(PKG + "." + name)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


# ------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [v.strip() for v in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# algorithmic bytes per launch (DESIGN.md "Roofline")
# ------------------------------------------------------------------------------------------------
def algorithmic_bytes(kernel, enc_lens, B, Tp, A, D, dec, V, labels, C=10, KW1=201):
    s = sum(enc_lens)
    wts = 4 * (A * dec + C * KW1 + C * A + A + 1)
    if kernel == "b2asr_attn_step_fwd":      # keys + h read once, prev/att rows, state in, ctx out, weights
        return 4 * s * (A + D + 2) + 4 * B * (dec + D) + wts
    if kernel == "b2asr_attn_step_bwd":      # keys + h read, dK read-modify-write, 4 alignment-sized rows
        return 4 * s * (3 * A + D + 4) + 4 * B * (2 * A + D + dec) + wts
    if kernel == "b2asr_ctc_loss_fwd_bwd":   # every valid logit row read once, every grad row written once
        return 4 * V * (s + B * Tp) + 8 * sum(labels) + 16 * B
    if kernel == "b2asr_attn_keys":          # features read, keys written, psi weight + bias (padded batch rows)
        return 4 * B * Tp * (D + A) + 4 * A * (D + 1)
    if kernel == "b2asr_ce_loss_fwd_bwd":    # decoder logits read once, finished gradient written once
        return 8 * V * B * max(labels) + 16 * B * max(labels)
    if kernel == "b2asr_decoder_loop_fwd":   # L steps of (attention step + cell); GEMM weight reads not counted
        L = max(labels)
        return L * (algorithmic_bytes("b2asr_attn_step_fwd", enc_lens, B, Tp, A, D, dec, V, labels, C, KW1) +
                    algorithmic_bytes("b2asr_lstm_cell_fwd", enc_lens, B, Tp, A, D, dec, V, labels, C, KW1))
    if kernel == "b2asr_decoder_loop_bwd":
        L = max(labels)
        return L * (algorithmic_bytes("b2asr_attn_step_bwd", enc_lens, B, Tp, A, D, dec, V, labels, C, KW1) +
                    algorithmic_bytes("b2asr_lstm_cell_bwd", enc_lens, B, Tp, A, D, dec, V, labels, C, KW1))
    if kernel == "b2asr_lstm_cell_fwd":      # gates read + activations written, c_prev in, h and c out
        return 4 * B * dec * (4 + 4 + 3)
    if kernel == "b2asr_lstm_cell_bwd":      # activations, c_prev, c, d_h, d_c in; d_gates, d_c_prev out
        return 4 * B * dec * (4 + 4 + 4 + 1)
    return 0


def encoder_layer_bytes(frame_lens, dims, rates, nd, backward):
    """Algorithmic bytes of the Listener recurrence (csrc/enc*.cu), averaged over the layers (one launch each):
    forward  reads the pre-activations [sum T_b, ND, 4H], writes activations (same), cells [sum T_b, ND, H] and the
             padded output [B, T, ND*H]; backward reads activations, two cell rows and d_out, writes the padded
             d_gates [B, T, ND, 4H]; both read W_hh once.  Also returns the layers' recurrent flops."""
    lens = list(frame_lens)
    tot = flops = 0.0
    for H, r in zip(dims, rates):
        s, B, T = sum(lens), len(lens), max(lens)
        w = 4 * nd * 4 * H * H
        if backward:
            tot += 4 * nd * s * (4 * H + 2 * H + H) + 4 * B * T * nd * 4 * H + w
        else:
            tot += 4 * nd * s * (4 * H + 4 * H + H) + 4 * B * T * nd * H + w
        flops += 2.0 * s * nd * 4 * H * H
        lens = [int(n / r) for n in lens]
    return tot / len(dims), flops / len(dims)


# ------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the oracle port of the reference's train step on the host cores
# ------------------------------------------------------------------------------------------------
def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU arm runs on rank 0 alone and may use the box."""
    import torch
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    if torch.get_num_threads() < n:
        torch.set_num_threads(n)
    return torch.get_num_threads()


CPU_SAMPLE_STRIDE = 8      # the CPU arms time utterances 0, 8, 16, 24 of the 32-utterance batch (every 8th: the batch is
                           # sorted by length, so a strided pick keeps its length distribution); always the same ones


def cpu_sample(workload, seed=1234):
    syn = pkg("synthetic")
    w = syn.WORKLOADS[workload]
    x, y, lens = syn.make_batch(workload, seed)
    stride = max(1, w["batch"] // 4)
    xs, ys = x[::stride].contiguous(), y[::stride].contiguous()
    ans = int((ys != 0).sum(-1).max())
    ys = ys[:, :ans + 1].contiguous()
    xs = xs[:, :lens[0]].contiguous()
    what = "utterances %s of the %d-utterance %s batch (seed %d; T=%d frames, longest label %d)" % (
        list(range(0, w["batch"], stride)), w["batch"], workload, seed, xs.shape[1], ans)
    return xs, ys, what


def cpu_train_step_rate(steps, warmup, workload=None):
    """Times the reference's train step (oracle/reference_step.py: the reference's own Seq2Seq.forward, losses,
    backward, clip, Adadelta step -- reference src/solver.py:132-182) on the host cores, on a FIXED sample of the
    workload: every 8th utterance of the first synthetic batch at full length, the same 4 in every run."""
    from oracle import reference_step
    syn = pkg("synthetic")
    workload = workload or WORKLOAD
    w = syn.WORKLOADS[workload]
    cores = use_all_host_threads()
    ref = reference_step.ReferenceStep(w["d_in"], w["vocab"], w["cfg"], "cpu")
    xs, ys, what = cpu_sample(workload)
    for _ in range(warmup):
        ref.step(xs, ys, tf_rate=1.0)
    t0 = time.perf_counter()
    for _ in range(steps):
        ref.step(xs, ys, tf_rate=1.0)
    dt = time.perf_counter() - t0
    sample = "%s; %d warm-up + %d timed steps" % (what, warmup, steps)
    cpu_train_step_rate.n_params = sum(p.numel() for p in ref.model.parameters())
    return xs.shape[0] * steps / dt, dt / steps * 1e3, cores, sample, ref.kind


def bench_config(workload, world, n_params=None):
    syn = pkg("synthetic")
    w = syn.WORKLOADS[workload]
    cfg = {"workload": workload, "batch_per_gpu": w["batch"], "global_batch": w["batch"] * world, "frames": w["t_hi"],
           "feature_dim": w["d_in"], "vocab": w["vocab"], "parallelism": "dp%d" % world,
           "l2": "inputs larger than L2: %d distinct batches rotate and one step touches >1 GB of "
                 "activations (126 MB L2)" % N_BATCHES}
    if world > 1:
        cfg["rank_batches"] = "same utterance/label lengths on every rank (fixed work per GPU), different content"
    if n_params is not None:
        cfg["params"] = n_params
    return cfg


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 10)), max(1, min(args.warmup, 3))
    value, ms, cores, sample, kind = cpu_train_step_rate(steps, warmup)
    cfg = bench_config(WORKLOAD, max(1, args.gpus), cpu_train_step_rate.n_params)
    cfg["note"] = ("the reference's own CPU path (its Seq2Seq.forward + torch.nn.CTCLoss + CrossEntropyLoss + backward + "
                   "clip + Adadelta, imported unmodified from oracle/_ref) on the host cores of rank 0; a step = the fixed "
                   "4-utterance sample of the workload's batch, value = utterances of the sample / time")
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "utterances/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": value, "unit": "utterances/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "utterances/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(json.dumps(out))


# ------------------------------------------------------------------------------------------------
# joint beam decode leg (BASELINE.json metric part 2: beam-decode RTF; SURVEY.md section 8d workload M4)
# ------------------------------------------------------------------------------------------------
DEC = dict(workload="libri_m2s", beam=20, ctc_weight=0.5, lm_weight=0.5, ratio=0.1, lm=(650, 650, 2), frames=1600)


def decode_leg(dev, rank, world, n_utt, dist):
    """Each rank decodes its own `n_utt` synthetic 16-s utterances (utterance-sharded, no collective):
    shipped libri_example.yaml-shaped model (VGGBiRNN, V=5000) + RNN_LM(650,650,2), beam 20,
    decode_ctc_weight 0.5, decode_lm_weight 0.5, max_decode_step_ratio 0.1 -> 160 steps."""
    import torch
    asr, rnnlm, syn = pkg("asr"), pkg("rnnlm"), pkg("synthetic")
    lib = pkg("_lib")
    w = syn.WORKLOADS[DEC["workload"]]
    torch.manual_seed(0)
    model = asr.Seq2Seq(torch.zeros(1, 8, w["d_in"]), w["vocab"], w["cfg"]).to(dev).eval()
    model.joint_ctc, model.ctc_weight, model.decode_lm_weight = True, DEC["ctc_weight"], DEC["lm_weight"]
    model.rnn_lm = rnnlm.RNN_LM(DEC["lm"][0], DEC["lm"][1], w["vocab"], layers=DEC["lm"][2]).to(dev).eval()
    g = torch.Generator().manual_seed(4321 + rank)
    utts = [torch.randn(1, DEC["frames"], w["d_in"], generator=g).pin_memory() for _ in range(n_utt + 1)]
    steps = int(math.ceil(DEC["frames"] * DEC["ratio"]))

    def one(xh):
        x = xh.to(dev, non_blocking=True)                          # host features in, token ids out
        hyps = model.beam_decode(x, steps, [DEC["frames"]], DEC["beam"])
        return hyps[0].outIndex

    with torch.no_grad():
        one(utts[-1])                                              # warm-up utterance
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        l0 = lib.LAUNCHES
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for u in range(n_utt):
            one(utts[u])
        e1.record()
        torch.cuda.synchronize()
        launches = lib.LAUNCHES - l0
        # the same utterances, up to four in flight on streams of their own (beam.beam_decode_many): one decode keeps
        # only a fraction of the SMs busy, so throughput (not latency) multiplies
        beam_mod = pkg("beam")
        many = lambda: beam_mod.beam_decode_many(
            model, [(utts[u].to(dev, non_blocking=True), steps, [DEC["frames"]]) for u in range(n_utt)], DEC["beam"], streams=4)
        many()
        torch.cuda.synchronize()
        # this leg is host-sensitive (one Python thread enqueues every utterance's launches): best of three passes
        ms_c = float("inf")
        for _ in range(3):
            c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            c0.record()
            outs = many()
            c1.record()
            torch.cuda.synchronize()
            assert len(outs) == n_utt and all(len(o) > 0 for o in outs)
            ms_c = min(ms_c, c0.elapsed_time(c1))
    ms = torch.tensor([e0.elapsed_time(e1), ms_c], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    sec, sec_c = float(ms[0]) * 1e-3, float(ms[1]) * 1e-3
    audio = n_utt * world * DEC["frames"] * 0.01
    return {"rtf": sec / audio * world, "rtf_aggregate": sec / audio, "utterances_per_s": n_utt * world / sec,
            "concurrent": {"streams": 4, "rtf": sec_c / audio * world, "utterances_per_s": n_utt * world / sec_c,
                           "ms_per_utterance": sec_c / n_utt * 1e3,
                           "note": "the same utterances with up to 4 in flight per GPU (throughput, best of 3 passes; `rtf` above is one at a time, single pass)"},
            "ms_per_utterance": sec / n_utt * 1e3, "decode_steps": steps, "utterances_per_gpu": n_utt,
            "beam": DEC["beam"], "ctc_weight": DEC["ctc_weight"], "lm_weight": DEC["lm_weight"], "vocab": w["vocab"],
            "audio_s_per_utterance": DEC["frames"] * 0.01, "gpu_launches": launches,
            "note": "rtf = per-GPU wall clock / audio seconds (end to end: host features in, token ids out); "
                    "rtf_aggregate divides by the audio all ranks decoded"}


def cpu_decode_rate(n_steps=2):
    """Oracle port of Seq2Seq.beam_decode (oracle/decode_oracle.py, restating asr.py:155-258 + ctc.py) on the host
    cores for `n_steps` decode steps of one M4 utterance; RTF extrapolated to the full 160 steps."""
    import torch
    from oracle import decode_oracle as do
    from oracle import model_oracle as mo
    syn = pkg("synthetic")
    w = syn.WORKLOADS[DEC["workload"]]
    use_all_host_threads()
    torch.manual_seed(0)
    model = mo.OracleModel(w["d_in"], w["vocab"], w["cfg"]).eval()
    lm = do.OracleLM(DEC["lm"][0], DEC["lm"][1], w["vocab"], layers=DEC["lm"][2]).eval()
    x = torch.randn(1, DEC["frames"], w["d_in"], generator=torch.Generator().manual_seed(4321))
    t0 = time.perf_counter()
    do.beam_decode(model, lm, x, 1, [DEC["frames"]], DEC["beam"], DEC["ctc_weight"], DEC["lm_weight"])
    t1 = time.perf_counter()
    do.beam_decode(model, lm, x, 1 + n_steps, [DEC["frames"]], DEC["beam"], DEC["ctc_weight"], DEC["lm_weight"])
    t2 = time.perf_counter()
    per_step = ((t2 - t1) - (t1 - t0)) / n_steps            # steps 2.. carry the full beam; step 1 only the root
    enc = t1 - t0
    steps = int(math.ceil(DEC["frames"] * DEC["ratio"]))
    rtf = (enc + per_step * steps) / (DEC["frames"] * 0.01)
    return {"rtf": rtf, "ms_per_decode_step": per_step * 1e3, "cores": torch.get_num_threads(), "kind": "port",
            "sample": "1 utterance: encoder + root step timed once, then %d full-beam decode steps; RTF extrapolated to %d steps"
                      % (n_steps, steps)}


# ------------------------------------------------------------------------------------------------
# the other BASELINE.json configurations, timed in the same run (reported under `workloads` / `ctc_m5`)
# ------------------------------------------------------------------------------------------------
def train_workload_leg(name, dev, steps, warmup, tf_rate):
    """BASELINE.json config 1 (timit_m1) / the shipped libri yaml (libri_m2s) through the same TrainStep, one GPU:
    -> resident and end-to-end utterances/s (the definitions of the headline's `value` and `e2e`)."""
    import torch
    lib, asr, solver, syn = pkg("_lib"), pkg("asr"), pkg("solver"), pkg("synthetic")
    w = syn.WORKLOADS[name]
    torch.manual_seed(0)
    model = asr.Seq2Seq(torch.zeros(1, 8, w["d_in"]), w["vocab"], w["cfg"]).to(dev).train()
    opt = solver.make_optimizer(model, w["cfg"]["optimizer"])
    ts = solver.TrainStep(model, opt, w["cfg"]["optimizer"]["joint_ctc"])
    host = []
    for i in range(N_BATCHES):
        x, y, _ = syn.make_batch(name, 1234 + 1000 * i)
        host.append((x.pin_memory(), y.pin_memory()))
    resident = [(x.to(dev), y.to(dev)) for x, y in host]

    def timed(fn, n):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    def step_resident(i):
        x, y = resident[i % N_BATCHES]
        ts.step(x, y, tf_rate=tf_rate)

    def step_e2e(i):
        xh, yh = host[i % N_BATCHES]
        out = ts.step(xh.to(dev, non_blocking=True), yh.to(dev, non_blocking=True), tf_rate=tf_rate)
        float(out["loss"])

    for i in range(warmup):
        step_resident(i)
    l0 = lib.LAUNCHES
    ms = timed(step_resident, steps)
    launches = lib.LAUNCHES - l0
    step_e2e(0)

    def run_e2e(n):
        torch.cuda.synchronize()
        gc.collect()
        gc.disable()                            # a generation-2 collection inside the loop is tens of ms of host time,
        try:                                    # which the per-step loss read-back turns into GPU idle time
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for x, y in solver.DevicePrefetcher((host[i % N_BATCHES] for i in range(n)), dev):
                out = ts.step(x, y, tf_rate=tf_rate)
                float(out["loss"])
            e1.record()
            torch.cuda.synchronize()
        finally:
            gc.enable()
        return e0.elapsed_time(e1)

    run_e2e(2)                                  # untimed: the prefetcher's stream and the allocator's blocks for the per-step copies
    ms_e = min(run_e2e(steps), run_e2e(steps))  # these short legs (5 steps) are at the mercy of one host hiccup: best of two passes
    B = w["batch"]
    out = {"value": B * steps / (ms * 1e-3), "unit": "utterances/s", "ms_per_step": ms / steps,
           "e2e": {"value": B * steps / (ms_e * 1e-3), "unit": "utterances/s", "ms_per_step": ms_e / steps,
                   "h2d_bytes_per_step": host[0][0].numel() * 4 + host[0][1].numel() * 8, "passes": "best of 2"},
           "steps": steps, "warmup": warmup, "tf_rate": tf_rate, "batch": B, "frames": w["t_hi"], "vocab": w["vocab"],
           "feature_dim": w["d_in"], "encoder": w["cfg"]["encoder"]["enc_type"], "gpu_launches": launches,
           "params": sum(p.numel() for p in model.parameters())}
    del ts, opt, model, resident
    torch.cuda.empty_cache()
    return out


M5_POINTS = [(500, 50), (1000, 150), (2000, 300), (3000, 600)]


def ctc_m5_leg(dev, iters=5, warm=3):
    """BASELINE.json config 5 (pure CTC stress sweep; main.py cannot train pure CTC, reference solver.py:188-190, so
    kernels only): B=64, V=5000, logits ~N(0,1) [B,T',V] read through (stride_t, stride_b), input lengths
    ~U[0.6T',T'] (max T'), label lengths ~U[0.5L,L] (max L), 10 % forced adjacent repeats.  One b2asr_ctc_loss_fwd_bwd
    call (fused log-softmax, 'mean', loss + gradient) per iteration, CUDA events on the launching stream; the 2x
    (0.6..3.8) GB of logits + gradient exceed the 126 MB L2, so nothing stays cached between iterations.
    achieved = algorithmic bytes 4*V*(sum_b T'_b + B*T') + 8*sum_b L_b + 16*B  /  time."""
    import torch
    lib = pkg("_lib")
    L = lib.lib()
    peak, how = measured_peaks()
    B, V = 64, 5000
    points = []
    for T, Lmax in M5_POINTS:
        g = torch.Generator().manual_seed(T)
        x = torch.randn(B, T, V, device=dev)
        il = torch.randint(int(0.6 * T), T + 1, (B,), generator=g); il[0] = T
        tl = torch.randint(Lmax // 2, Lmax + 1, (B,), generator=g); tl[0] = Lmax
        y = torch.randint(1, V, (B, Lmax), generator=g)
        y[:, 1::10] = y[:, 0::10][:, : y[:, 1::10].shape[1]]
        for b in range(B):
            y[b, tl[b]:] = 0
        yd, ild, tld = y.to(dev), il.to(dev), tl.to(dev)
        nbytes = L.b2asr_ctc_workspace_bytes(T, B, V, Lmax, 1)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        nll = torch.empty(B, device=dev)
        loss = torch.empty(1, device=dev)
        grad = torch.empty_like(x)
        st = lib.stream_ptr()

        def call():
            lib.call("b2asr_ctc_loss_fwd_bwd", 3, L.b2asr_ctc_loss_fwd_bwd, x.data_ptr(), x.stride(1), x.stride(0),
                     yd.data_ptr(), yd.stride(0), ild.data_ptr(), tld.data_ptr(), T, B, V, Lmax, 0, 1, 1, 1.0,
                     nll.data_ptr(), loss.data_ptr(), grad.data_ptr(), grad.stride(1), grad.stride(0),
                     ws.data_ptr(), nbytes, st)

        for _ in range(warm):
            call()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            call()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / iters * 1e3
        alg = 4 * V * (int(il.sum()) + B * T) + 8 * int(tl.sum()) + 16 * B
        points.append({"frames": T, "label_len": Lmax, "us": us, "alg_gb": alg / 1e9, "achieved_gbs": alg / us / 1e3,
                       "frac": alg / us / 1e3 / peak, "utterances_per_s": B / (us * 1e-6),
                       "workspace_mb": nbytes / 1e6, "finite": bool(torch.isfinite(loss).all())})
        del x, grad, ws
        torch.cuda.empty_cache()
    return {"batch": B, "vocab": V, "peak_gbs": peak, "peak_source": how, "iters": iters, "warmup": warm,
            "kernel": "b2asr_ctc_loss_fwd_bwd (from logits, loss + gradient)", "points": points}


def stock_torch_gpu_leg(dev, steps=3, warmup=2):
    """The reference's OWN GPU path (reference src/solver.py:29 runs on CUDA by default): its unmodified
    Seq2Seq.forward + torch.nn.CTCLoss + backward + Adadelta on cuda:0 through stock PyTorch ops (cuDNN LSTM, cuBLAS,
    ATen CTC), fp32 with TF32 off, on the SAME full-size batches as the headline.  A reported baseline: the
    like-for-like speed-up of the hand-written path over stock PyTorch on the same B200."""
    import torch
    from oracle import reference_step
    syn = pkg("synthetic")
    w = syn.WORKLOADS[WORKLOAD]
    ref = reference_step.ReferenceStep(w["d_in"], w["vocab"], w["cfg"], dev)
    batches = [syn.make_batch(WORKLOAD, 1234 + 1000 * i)[:2] for i in range(N_BATCHES)]
    batches = [(x.to(dev), y.to(dev)) for x, y in batches]
    for i in range(warmup):
        ref.step(*batches[i % N_BATCHES], tf_rate=1.0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        ref.step(*batches[i % N_BATCHES], tf_rate=1.0)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    out = {"value": w["batch"] / (ms * 1e-3), "unit": "utterances/s", "ms_per_step": ms, "steps": steps,
           "warmup": warmup, "kind": ref.kind, "workload": WORKLOAD,
           "note": "unmodified reference modules on cuda:0 (stock PyTorch/cuDNN/cuBLAS/ATen ops), inputs resident, "
                   "fp32 (cudnn.allow_tf32 = False), full 32-utterance batches"}
    del ref, batches
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------------
# this repo's arm
# ------------------------------------------------------------------------------------------------
def run_native(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = pkg("_lib")
    lib.lib()                                          # fail loudly if libb2asr.so is missing
    asr, solver, syn = pkg("asr"), pkg("solver"), pkg("synthetic")
    w = syn.WORKLOADS[WORKLOAD]
    cfg = w["cfg"]

    torch.manual_seed(0)                               # identical weights on every rank
    ex = torch.zeros(1, 8, w["d_in"])
    model = asr.Seq2Seq(ex, w["vocab"], cfg).to(dev)
    model.train()
    opt = solver.make_optimizer(model, cfg["optimizer"])
    ts = solver.TrainStep(model, opt, cfg["optimizer"]["joint_ctc"])
    n_params = sum(p.numel() for p in model.parameters())

    host = []
    for i in range(N_BATCHES):
        # every rank: the same utterance and label LENGTHS (fixed work per GPU = weak scaling; otherwise the step waits
        # for whichever rank drew the longest batch), its own content
        x, y, lens = syn.make_batch(WORKLOAD, 1234 + rank + 1000 * i, shape_seed=1234 + 1000 * i)
        host.append((x.pin_memory(), y.pin_memory(), lens))
    resident = [(x.to(dev), y.to(dev)) for x, y, _ in host]
    B = w["batch"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    def step_resident(i):
        x, y = resident[i % N_BATCHES]
        ts.step(x, y, tf_rate=1.0)

    d2h = [0]

    def step_e2e(i):
        xh, yh, _ = host[i % N_BATCHES]
        x = xh.to(dev, non_blocking=True)
        y = yh.to(dev, non_blocking=True)
        out = ts.step(x, y, tf_rate=1.0)
        float(out["loss"])                             # the step's result crosses back to the host
        d2h[0] = 4 + 4 + 8 * B + 8                      # loss, grad-norm, B lengths, max label length

    for i in range(args.warmup):
        step_resident(i)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = lib.LAUNCHES
    ms_total = timed(step_resident, args.steps)
    launches = lib.LAUNCHES - l0
    clocks = sampler.stop() if rank == 0 else None
    ms_step = ms_total / args.steps
    value = B * world * args.steps / (ms_total * 1e-3)

    for i in range(min(2, args.warmup)):
        step_e2e(i)

    def run_e2e(steps):
        """K steps through the public API from pinned host batches: solver.DevicePrefetcher issues batch i+1's
        host->device copies on a side stream under step i; every copy and every loss read-back is inside the timed
        region (the first copy is issued after the start event)."""
        barrier()
        gc.collect()
        gc.disable()                            # a generation-2 collection inside the loop is tens of ms of host time,
        try:                                    # which the per-step loss read-back turns into GPU idle time
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for x, y in solver.DevicePrefetcher(((host[i % N_BATCHES][0], host[i % N_BATCHES][1]) for i in range(steps)), dev):
                out = ts.step(x, y, tf_rate=1.0)
                float(out["loss"])
            e1.record()
        finally:
            gc.enable()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    run_e2e(2)                                  # untimed: the prefetcher's stream and the allocator's blocks for the per-step copies
    ms_e2e = run_e2e(args.steps)
    e2e_value = B * world * args.steps / (ms_e2e * 1e-3)
    h2d = host[0][0].numel() * 4 + host[0][1].numel() * 8

    # ---- per-kernel timing pass (same steps, CUDA events around every launch of libb2asr.so) -------
    gemm = pkg("_gemm")
    flops0 = dict(gemm.FLOPS)
    lib.profile_start()
    prof_steps = min(args.steps, 3)
    for i in range(prof_steps):
        step_resident(i)
    prof = lib.profile_stop()
    gemm_flops = {k: (gemm.FLOPS[k] - flops0[k]) / prof_steps for k in gemm.FLOPS}
    roofline = None
    kernels = {}
    if prof:
        peak, how = measured_peaks()
        att = cfg["attention"]
        D = int(cfg["encoder"]["dim"].split("_")[-1]) * 2
        enc_flops = {}
        for name, (calls, tot_ms) in prof.items():
            per = tot_ms / max(calls, 1)
            # algorithmic bytes use the first batches' actual lengths (averaged over the profiled steps)
            ab = 0.0
            for i in range(prof_steps):
                _, yb, lens = host[i % N_BATCHES]
                enc = syn.encoder_lengths(WORKLOAD, lens)
                labels = [int(v) for v in (yb != 0).sum(-1)]
                if name.startswith("b2asr_lstm_seq"):
                    e = cfg["encoder"]
                    eb, ef = encoder_layer_bytes([int(n) for n in lens], [int(v) for v in e["dim"].split("_")],
                                                 [int(v) for v in e["sample_rate"].split("_")],
                                                 2 if "BiRNN" in e["enc_type"] else 1, name.endswith("bwd"))
                    ab += eb
                    enc_flops[name] = enc_flops.get(name, 0.0) + ef / prof_steps
                else:
                    ab += algorithmic_bytes(name, enc, B, max(enc), att["dim"], D, cfg["decoder"]["dim"], w["vocab"], labels)
            ab /= prof_steps
            kernels[name] = {"calls_per_step": calls / prof_steps, "avg_us": per * 1e3,
                             "ms_per_step": tot_ms / prof_steps, "alg_bytes_per_launch": ab,
                             "achieved_gbs": (ab / (per * 1e-3) / 1e9) if per > 0 else None}
            if name in gemm_flops and tot_ms > 0:
                # tensor-core products: algorithmic 2MNK per step over the summed launch time (three times that is
                # issued: error-compensated TF32); the weight-gradient launches share the device with the recurrence
                kernels[name]["alg_gflop_per_step"] = gemm_flops[name] / 1e9
                kernels[name]["achieved_tflops"] = gemm_flops[name] / (tot_ms / prof_steps * 1e-3) / 1e12
        top = max(kernels, key=lambda k: kernels[k]["ms_per_step"])
        k = kernels[top]
        note = "M2 keys+features (48 MB) are L2-resident across decoder steps; see DESIGN.md"
        if top.startswith("b2asr_lstm_seq"):
            fp32_peak = 148 * 128 * 2 * 1.965e9 / 1e12          # B200 FFMA peak, TFLOP/s
            note = ("serial recurrence: T dependent frames per launch, bound by the fp32 FMA pipe of the 96 SMs the "
                    "7 resident 16-CTA clusters occupy and by the per-frame exchange; recurrent MACs run at %.1f of "
                    "%.1f fp32 TFLOP/s; see DESIGN.md section 4" % (enc_flops[top] / (k["avg_us"] * 1e-6) / 1e12, fp32_peak))
        # DRAM bytes of the call's dominant kernel from the committed `ncu --set full` captures (cold-cache replay
        # of one launch at the M2 shape), scaled to the launches one call makes
        traffic = None
        if WORKLOAD == "libri_m2" and top in NCU_DRAM_BYTES:
            per_launch, src = NCU_DRAM_BYTES[top]
            steps_per_call = 1.0
            if top.startswith("b2asr_decoder_loop"):
                steps_per_call = sum(max(int(v) for v in (host[i % N_BATCHES][1] != 0).sum(-1))
                                     for i in range(prof_steps)) / prof_steps
            traffic = per_launch * steps_per_call
            note += "; traffic = %s x %.0f launches per call (cold-cache replay: in the step the operands stay in L2)" % (src, steps_per_call)
        roofline = {"kernel": top, "bound": "hbm", "achieved": k["achieved_gbs"], "peak": peak, "unit": "GB/s",
                    "frac": k["achieved_gbs"] / peak, "traffic": traffic, "peak_source": how,
                    "share_of_step": k["ms_per_step"] / ms_step, "note": note}

    # ---- data-parallel breakdown (N > 1): where the step time goes on every rank -----------------------------
    # CUDA events on each rank: step start -> all-reduce start = that rank's own compute; the all-reduce interval =
    # transfer + waiting for the slowest rank (the NCCL kernel spins until every peer has arrived).
    dp = None
    if world > 1:
        ts.grads.timeline = []
        marks = []
        barrier()
        for i in range(min(args.steps, 5)):
            e0, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            step_resident(i)
            e2.record()
            marks.append((e0, e2))
        torch.cuda.synchronize()
        rows = []
        for (e0, e2), (a0, a1) in zip(marks, ts.grads.timeline):
            rows.append([e0.elapsed_time(a0), a0.elapsed_time(a1), a1.elapsed_time(e2)])
        ts.grads.timeline = None
        mine = torch.tensor(rows, device=dev).mean(0)                    # [compute, allreduce, tail] ms on this rank
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        if rank == 0:
            comp = [float(v[0]) for v in allr]
            ar = [float(v[1]) for v in allr]
            dp = {"compute_ms_per_rank": [round(v, 2) for v in comp], "allreduce_ms_per_rank": [round(v, 2) for v in ar],
                  "tail_ms_per_rank": [round(float(v[2]), 2) for v in allr],
                  "compute_ms_spread": round(max(comp) - min(comp), 2), "allreduce_ms_min": round(min(ar), 2),
                  "gradient_mb": ts.grads.flat.numel() * 4 / 1e6,
                  "note": "per rank: step start -> all-reduce start (own forward+backward on its own batch), the NCCL "
                          "all-reduce interval (transfer + wait for the slowest rank; its minimum over ranks is the "
                          "transfer itself), all-reduce end -> step end (norm + clipped optimizer launch)"}
    del ts, opt, model, resident
    torch.cuda.empty_cache()
    workloads, ctc_m5, stock = None, None, None
    if world == 1 and WORKLOAD == "libri_m2" and not args.no_extra:
        k = max(3, min(args.steps, 5))
        workloads = {"timit_m1": train_workload_leg("timit_m1", dev, k, 3, tf_rate=0.8),
                     "timit_m1_tf1": train_workload_leg("timit_m1", dev, k, 3, tf_rate=1.0),
                     "libri_m2s": train_workload_leg("libri_m2s", dev, k, 3, tf_rate=1.0)}
        ctc_m5 = ctc_m5_leg(dev)
        stock = stock_torch_gpu_leg(dev)

    decode = None
    if not args.no_decode:
        decode = decode_leg(dev, rank, world, args.decode_utts, dist)

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu:
        v, ms, cores, sample, kind = cpu_train_step_rate(steps=3, warmup=1)
        cpu_baseline = {"value": v, "unit": "utterances/s", "cores": cores, "kind": kind, "sample": sample,
                        "ms_per_step": ms}
        if decode is not None:
            decode["cpu_baseline"] = cpu_decode_rate()

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": "utterances/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": bench_config(WORKLOAD, world, n_params),
            "clocks": clocks, "gpu_launches": launches,
            "e2e": {"value": e2e_value, "unit": "utterances/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h[0], "ms_per_step": ms_e2e / args.steps},
            "roofline": roofline, "kernels": kernels, "cpu_baseline": cpu_baseline, "stock_torch_gpu": stock,
            "workloads": workloads, "ctc_m5": ctc_m5, "decode": decode, "dp_breakdown": dp,
        }
        emit(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def claim_stdout():
    """ONE JSON line on stdout: libraries print there too (NCCL's version banner comes from C code at communicator
    set-up), so file descriptor 1 is pointed at stderr for the whole run and the line goes to a saved copy."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT if _REAL_STDOUT is not None else sys.stdout
    out.write(line + "\n")
    out.flush()


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-decode", action="store_true", help="skip the beam-decode RTF leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the other BASELINE.json configurations "
                    "(timit_m1, libri_m2s, ctc_m5) and the stock-PyTorch GPU baseline")
    ap.add_argument("--decode-utts", type=int, default=4, help="utterances per GPU in the decode leg")
    ap.add_argument("--workload", default=WORKLOAD, help="train-step workload (synthetic.WORKLOADS); the default is the "
                    "configuration BASELINE.json's metric is quoted on, the others are reported in BASELINE.md only")
    args = ap.parse_args()
    globals()["WORKLOAD"] = args.workload
    if args.impl == "reference":
        run_reference(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
