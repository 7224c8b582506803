#!/usr/bin/env python
"""bench.py -- mel frames/sec of the B200-native TransformerTTS hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload fwd|ar]

A "step" is one teacher-forced forward (TransformerTTS.forward, model/model.py:260-320) over one batch of
synthetic LJSpeech-shaped input; at N=1 the workload is BASELINE.json configs[1] (B=64, Tp=150, Tm=800,
alignments returned).  With N>1 (torchrun, one rank per GPU) every rank runs the same per-GPU batch on its
own utterances: weak scaling, no collective on the data path; the timing is the max over ranks.
The line also carries AR-decode numbers (configs[2]) with their own end-to-end and CPU figures, the roofline of the
dominant kernel, the CPU baseline and clocks sampled during the timed region.

`--impl reference` times the reference's own CPU implementation of the same path with all host threads, rank 0 only:
the UNMODIFIED `model.TransformerTTS` staged under oracle/_ref by oracle/make_ref.py (kind "reference"; if that
directory did not travel, the oracle port, kind "port").  One step = one forward over BASELINE.json configs[0]
(B=16, Tp=150, Tm=800: a 16-utterance sample of the headline batch's workload, the reference's own CPU-runnable case).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

D, H, F, M, KS, L = 256, 4, 1024, 80, 5, 3
CFG2 = dict(B=64, Tp=150, Tm=800)


# ---------------------------------------------------------------------------------------------- helpers
def algorithmic_flops(B, Tp, Tm, causal_skip=True):
    """analytic FLOPs of one teacher-forced forward over the padded rectangle."""
    Np, Nm = B * Tp, B * Tm
    c = 0.5 if causal_skip else 1.0
    enc_pre = Np * (3 * 2 * D * D * KS + 2 * D * D)
    enc = Np * L * (8 * D * D + 4 * D * F + 4 * Tp * D)
    dec_pre = Nm * (2 * M * D + 2 * D * D)
    dec = Nm * L * (8 * D * D + 4 * D * D + 4 * D * F + c * 4 * Tm * D + 4 * Tp * D) + Np * L * 4 * D * D
    heads = Nm * (2 * D * M + 2 * D)
    post = Nm * (2 * 2 * M * D * KS + 3 * 2 * D * D * KS)
    return float(enc_pre + enc + dec_pre + dec + heads + post)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm_gbs=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sustained=d["bf16_tflops_sustained"],
                    source="measured")
    return dict(hbm_gbs=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms while the timed regions (device-resident steps and the
    end-to-end runs) execute."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = sorted(int(float(r[0])) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit())
        mx = [int(float(r[1])) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i].lower() == "active" for r in self.rows)]
        # median over the upper half = samples taken while the GPU was busy
        busy = sm[len(sm) // 2:] if sm else []
        return {"sm_mhz": busy[len(busy) // 2] if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---------------------------------------------------------------------------------------------- CPU arm
def _oracle():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import tts_oracle as O          # checker / CPU baseline only; never on the product path
    return O


CFG1 = dict(B=16, Tp=150, Tm=800)   # BASELINE.json configs[0]: the reference's own CPU-runnable case


class CpuReference:
    """The reference's CPU implementation of the path: the unmodified `model.TransformerTTS` from oracle/_ref
    (oracle/make_ref.py) when it travelled with the repo, else the oracle port.  fp32, eval, no_grad, all host threads,
    torch's default MHA fast path left on (the stock code path)."""

    def __init__(self, seed=1234):
        import torch
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import make_ref
        self.O = _oracle()
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        self.sd = self.O.make_state_dict(seed, stop_bias=-20.0)      # never-stop bias for the AR runs
        loaded = make_ref.load()
        if loaded is not None:
            self.kind = "reference"
            self.module = loaded[0](**self.O.CONFIG, device="cpu").eval()
            self.module.load_state_dict(self.sd, strict=True)
        else:
            self.kind, self.module = "port", None
        self.threads = torch.get_num_threads()

    def forward(self, ph, mel, pl, ml):
        import torch
        with torch.no_grad():
            if self.module is not None:
                return self.module(ph, mel, pl, ml)
            return self.O.forward(self.sd, ph, mel, pl, ml)

    def inference(self, ph, pl, max_len):
        """model/model.py:323-394 as-is: no KV cache, the whole prefix recomputed every step."""
        import torch
        with torch.no_grad():
            if self.module is not None:
                return self.module.inference(ph, pl, max_len=max_len)
            return self.O.inference(self.sd, ph, pl, max_len=max_len)

    def describe(self):
        return ("unmodified reference model.TransformerTTS (oracle/_ref), fp32, eval, no_grad" if self.kind == "reference"
                else "oracle port of the reference (oracle/_ref absent), fp32 torch")


def cpu_baselines(reps=3, ar_frames=200):
    """Bounded CPU sample for the `cpu_baseline` objects (rank 0, N=1): configs[0] forward + the AR loop at B=1 / B=32
    for `ar_frames` of the 1000 frames (the reference recomputes the prefix: O(T^2), so the full horizon takes minutes)."""
    ref = CpuReference()
    ph, mel, pl, ml = make_batch(CFG1["B"], CFG1["Tp"], CFG1["Tm"], seed=1)
    ref.forward(ph, mel, pl, ml)                              # warm-up (thread pool, allocator)
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        ref.forward(ph, mel, pl, ml)
        ts.append(time.perf_counter() - t0)
    sec = sorted(ts)[len(ts) // 2]
    fwd = {"value": int(ml.sum()) / sec, "unit": "valid mel frames/s", "cores": ref.threads, "kind": ref.kind,
           "sample": f"configs[0] exactly: B=16, Tp=150, Tm=800 (16 utterances of the headline batch's lens law), median of "
                     f"{reps} after 1 warm-up, {sec:.3f} s per forward; {ref.describe()}"}
    ar = {}
    for Bar in (1, 32):
        aph, _, apl, _ = make_batch(Bar, CFG1["Tp"], 8, seed=100)
        t0 = time.perf_counter()
        r = ref.inference(aph, apl, max_len=ar_frames + 1)
        sec = time.perf_counter() - t0
        T = r["pred_melspec"].shape[1]
        ar[f"B{Bar}"] = {"value": Bar * T / sec, "unit": "mel frames/s", "cores": ref.threads, "kind": ref.kind,
                         "sample": f"reference inference() as-is (no KV cache), B={Bar}, Tp=150, first {T} of the 1000 "
                                   f"frames, one run, {sec:.2f} s; throughput falls further with length (O(T^2))"}
    return fwd, ar


def our_config(world):
    """`config` of the JSON line: the workload only, so both arms print the same object truthfully; how each arm runs it
    (graph replay, L2 flush, weight precision) is in that arm's `run` object."""
    return {"workload": "configs[1]: teacher-forced forward, B=64 per GPU, Tp=150, Tm=800, alignments "
                        "returned (3x[B,4,Tm,Tp] fp32)",
            "batch_per_gpu": CFG2["B"], "parallelism": f"batch-sharded x{world}, no collective",
            "l2": "GPU arm: 256 MiB buffer written between timed iterations, and the working set (475 MB) exceeds L2; "
                  "CPU arm: n/a"}


def run_notes(arm, graph=True):
    if arm == "reference":
        return {"launch": "reference model.TransformerTTS.forward under torch.no_grad(), eval mode, CPU threads",
                "weights": "random init (reference init laws), fp32"}
    return {"launch": ("forward replayed as a CUDA graph (module option use_cuda_graph)" if graph else
                       "stream launches (--no-graph)"),
            "weights": "random init (reference init laws), bf16 on the GPU arm"}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation, same metric / unit / config; one step = one forward over
    configs[1] itself (B=64, Tp=150, Tm=800: the batch rank 0 of the GPU arm runs), about 1-2 s per step on 16 cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    ref = CpuReference()
    ph, mel, pl, ml = make_batch(CFG2["B"], CFG2["Tp"], CFG2["Tm"], seed=1)
    frames = int(ml.sum())
    times = []
    for _ in range(max(args.warmup, 1)):
        ref.forward(ph, mel, pl, ml)
    for _ in range(args.steps):
        t0 = time.perf_counter()
        ref.forward(ph, mel, pl, ml)
        times.append(time.perf_counter() - t0)
    ms = 1e3 * sum(times) / len(times)
    value = frames / (ms / 1e3)
    sample = (f"each step: configs[1] exactly (B=64, Tp=150, Tm=800; {frames} valid frames, the batch of the GPU arm's "
              f"rank 0), alignments returned; {ref.describe()}, torch.set_num_threads({ref.threads})")
    line = {
        "impl": "reference", "metric": "mel frames/sec (teacher-forced forward)", "value": value,
        "unit": "valid mel frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": our_config(1), "run": run_notes("reference"),
        "frames": {"batch": CFG2["B"], "valid_frames_per_step": frames, "padded_frames_per_step": CFG2["B"] * CFG2["Tm"]},
        "cpu_baseline": {"value": value, "unit": "valid mel frames/s", "cores": ref.threads, "kind": ref.kind,
                         "sample": sample},
        "e2e": {"value": value, "unit": "valid mel frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ---------------------------------------------------------------------------------------------- GPU arm
def make_batch(B, Tp, Tm, seed):
    """Synthetic LJSpeech-shaped batch: lens ~ U[2/3 Tp, Tp] / U[5/8 Tm, Tm], rows sorted by
    phoneme length descending (dataset.py:74-76), zero padding, mel ~ N(0,1)."""
    import torch
    g = torch.Generator().manual_seed(seed)
    pl = torch.randint(int(Tp * 2 / 3), Tp + 1, (B,), generator=g)
    ml = torch.randint(int(Tm * 5 / 8), Tm + 1, (B,), generator=g)
    pl[0] = Tp
    ml[int(torch.randint(0, B, (1,), generator=g))] = Tm
    pl, _ = torch.sort(pl, descending=True)
    ph = torch.randint(1, 100, (B, Tp), generator=g) * (torch.arange(Tp)[None, :] < pl[:, None])
    mel = torch.randn(B, Tm, M, generator=g) * (torch.arange(Tm)[None, :] < ml[:, None])[:, :, None]
    return ph.long(), mel.float(), pl.long(), ml.long()


class L2Flusher:
    def __init__(self, dev):
        import torch
        self.buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # 256 MiB > 126 MB L2

    def __call__(self):
        self.buf.zero_()


def time_kernel(fn, reps, flush):
    """Average device time (ms) of one launch of fn(), events on the launching (current) stream."""
    import torch
    for _ in range(3):
        fn()
    tot = 0.0
    for _ in range(reps):
        flush()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        b.synchronize()
        tot += a.elapsed_time(b)
    return tot / reps


def dominant_kernel_candidates(model, B, Tp, Tm, dev):
    """The kernel classes that carry the step: the fused FFN block (both GEMMs of a decoder layer's FFN, the largest
    GEMM work of the step), the causal self-attention and the cross-attention with its alignment store.  Each entry:
    name, FLOPs (or bytes) per launch, launches of that class per step, callable."""
    import ctypes as C
    import torch
    from transformertts_b200 import _ext
    lib = _ext.load()
    Nm = B * Tm
    st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
    A = torch.randn(Nm, D, device=dev).to(torch.bfloat16)
    W = (torch.randn(F, D, device=dev) / 16).to(torch.bfloat16)
    W2 = (torch.randn(D, F, device=dev) / 32).to(torch.bfloat16)
    bias = torch.zeros(F, device=dev)
    b2, g, be = torch.zeros(D, device=dev), torch.ones(D, device=dev), torch.zeros(D, device=dev)
    out = torch.empty(Nm, D, dtype=torch.bfloat16, device=dev)

    def gemm():     # the fused FFN block of one decoder layer: LayerNorm(x + W2 relu(W1 x + b1) + b2)
        _ext.check(lib.tts_op_ffn(_ext.ptr(A), _ext.ptr(W), _ext.ptr(bias), _ext.ptr(W2), _ext.ptr(b2), _ext.ptr(g),
                                  _ext.ptr(be), Nm, _ext.ptr(out), st()), "tts_op_ffn")

    qkv = (torch.randn(B, Tm, 3 * D, device=dev) * 0.5).to(torch.bfloat16)
    o = torch.empty(B, Tm, D, dtype=torch.bfloat16, device=dev)
    klen = torch.full((B,), Tm, dtype=torch.int32, device=dev)
    q, k, v = qkv[:, :, :D], qkv[:, :, D:2 * D], qkv[:, :, 2 * D:]

    def attn():
        _ext.check(lib.tts_op_attention(C.c_void_p(q.data_ptr()), 3 * D, C.c_void_p(k.data_ptr()), 3 * D,
                                        C.c_void_p(v.data_ptr()), 3 * D, B, Tm, Tm, _ext.ptr(klen), 1, _ext.ptr(o),
                                        None, st()), "tts_op_attention")

    # cross-attention with the alignment store: the kernel that writes the 3 x [B,4,Tm,Tp] fp32 probabilities
    cq = (torch.randn(B, Tm, D, device=dev) * 0.5).to(torch.bfloat16)
    ckv = (torch.randn(B, Tp, 2 * D, device=dev) * 0.5).to(torch.bfloat16)
    co = torch.empty(B, Tm, D, dtype=torch.bfloat16, device=dev)
    probs = torch.empty(B, H, Tm, Tp, device=dev)
    plen = torch.full((B,), Tp, dtype=torch.int32, device=dev)

    def cross():
        _ext.check(lib.tts_op_attention(_ext.ptr(cq), D, C.c_void_p(ckv.data_ptr()), 2 * D,
                                        C.c_void_p(ckv.data_ptr() + 2 * D), 2 * D, B, Tm, Tp, _ext.ptr(plen), 0,
                                        _ext.ptr(co), _ext.ptr(probs), st()), "tts_op_attention")

    # the persistent GEMM family, one shape per plan (profiles/README.md: 48 % of the step in round 1)
    Wq = (torch.randn(3 * D, D, device=dev) / 16).to(torch.bfloat16)
    bq = torch.zeros(3 * D, device=dev)
    oq = torch.empty(Nm, 3 * D, dtype=torch.bfloat16, device=dev)

    def qkv_proj():    # W-stationary plan: decoder self-attention in-projection
        _ext.check(lib.tts_op_linear(_ext.ptr(A), _ext.ptr(Wq), _ext.ptr(bq), Nm, 3 * D, D, 0, None, 0.0, 0, None, None,
                                     None, _ext.ptr(oq), None, st()), "tts_op_linear")

    Wo = (torch.randn(D, D, device=dev) / 16).to(torch.bfloat16)
    res = torch.randn(Nm, D, device=dev).to(torch.bfloat16)

    def out_ln():      # W-stationary + residual + LayerNorm on CTA pairs: attention out-projection
        _ext.check(lib.tts_op_linear(_ext.ptr(A), _ext.ptr(Wo), _ext.ptr(b2), Nm, D, D, 0, None, 0.0, 0, _ext.ptr(res),
                                     _ext.ptr(g), _ext.ptr(be), _ext.ptr(out), None, st()), "tts_op_linear")

    Wc = (torch.randn(D, KS * D, device=dev) / 36).to(torch.bfloat16)
    xc = torch.randn(B, Tm, D, device=dev).to(torch.bfloat16)
    oc = torch.empty(B, Tm, D, dtype=torch.bfloat16, device=dev)

    def conv():        # streamed-W plan on CTA pairs: PostNet Conv1d 256 -> 256, k = 5, folded BN, tanh
        _ext.check(lib.tts_op_conv1d(_ext.ptr(xc), _ext.ptr(Wc), _ext.ptr(b2), B, Tm, D, D, KS, 2, None, _ext.ptr(oc), None,
                                     st()), "tts_op_conv1d")

    # the remaining decoder-side shapes of the same kernel, so that the gemm2 class of the step is complete
    ocq = torch.empty(Nm, D, dtype=torch.bfloat16, device=dev)

    def cross_q():     # W-stationary: cross-attention query projection
        _ext.check(lib.tts_op_linear(_ext.ptr(A), _ext.ptr(Wo), _ext.ptr(b2), Nm, D, D, 0, None, 0.0, 0, None, None, None,
                                     _ext.ptr(ocq), None, st()), "tts_op_linear")

    pe_tab = model.pe.pe.detach().to(dev).float().contiguous()

    def prenet2():     # W-stationary + ReLU + alpha * PE epilogue: decoder prenet layer 2
        _ext.check(lib.tts_op_linear(_ext.ptr(A), _ext.ptr(Wo), _ext.ptr(b2), Nm, D, D, 1, _ext.ptr(pe_tab), 0.7, Tm, None, None,
                                     None, _ext.ptr(ocq), None, st()), "tts_op_linear")

    A128 = torch.randn(Nm, 128, device=dev).to(torch.bfloat16)
    W128 = (torch.randn(D, 128, device=dev) / 9).to(torch.bfloat16)

    def prenet1():     # W-stationary, K = 128 (80 mel bins zero padded) + ReLU
        _ext.check(lib.tts_op_linear(_ext.ptr(A128), _ext.ptr(W128), _ext.ptr(b2), Nm, D, 128, 1, None, 0.0, 0, None, None, None,
                                     _ext.ptr(ocq), None, st()), "tts_op_linear")

    Wh = (torch.randn(M, D, device=dev) / 16).to(torch.bfloat16)
    bh = torch.zeros(M, device=dev)
    oh32 = torch.empty(Nm, M, device=dev)
    oh16 = torch.empty(Nm, M, dtype=torch.bfloat16, device=dev)

    def heads():       # fp32-output plan: mel head (+ bf16 copy for the PostNet)
        _ext.check(lib.tts_op_linear(_ext.ptr(A), _ext.ptr(Wh), _ext.ptr(bh), Nm, M, D, 0, None, 0.0, 0, None, None, None,
                                     _ext.ptr(oh16), _ext.ptr(oh32), st()), "tts_op_linear")

    x80 = torch.randn(B, Tm, M, device=dev).to(torch.bfloat16)
    Wc0 = (torch.randn(D, KS * 128, device=dev) / 20).to(torch.bfloat16)

    def conv_first():  # streamed W: PostNet Conv1d 80 -> 256 (channels padded to 128 per tap), tanh
        _ext.check(lib.tts_op_conv1d(_ext.ptr(x80), _ext.ptr(Wc0), _ext.ptr(b2), B, Tm, M, D, KS, 2, None, _ext.ptr(oc), None,
                                     st()), "tts_op_conv1d")

    Wc4 = (torch.randn(M, KS * D, device=dev) / 36).to(torch.bfloat16)
    o80 = torch.empty(B, Tm, M, device=dev)
    r80 = torch.randn(B, Tm, M, device=dev)

    def conv_last():   # fp32-output plan: PostNet Conv1d 256 -> 80 + fp32 residual
        _ext.check(lib.tts_op_conv1d(_ext.ptr(xc), _ext.ptr(Wc4), _ext.ptr(bh), B, Tm, D, M, KS, 0, _ext.ptr(r80), None,
                                     _ext.ptr(o80), st()), "tts_op_conv1d")

    keep = (A, W, W2, bias, b2, g, be, out, qkv, o, klen, cq, ckv, co, probs, plen, Wq, bq, oq, Wo, res, Wc, xc, oc, ocq, pe_tab,
            A128, W128, Wh, bh, oh32, oh16, x80, Wc0, Wc4, o80, r80)
    extra = [
        dict(key="gemm2_cross_q", cls="gemm2_kernel", name="gemm2_kernel W-stationary: cross-attention Q projection (51200 x 256 x 256)",
             bound="tensor", work=2.0 * Nm * D * D, per_step=3, fn=cross_q, keep=keep, hbm_bytes=float(2 * Nm * D * 2)),
        dict(key="gemm2_prenet2", cls="gemm2_kernel", name="gemm2_kernel W-stationary + ReLU + PE: decoder prenet 2 (51200 x 256 x 256)",
             bound="tensor", work=2.0 * Nm * D * D, per_step=1, fn=prenet2, keep=keep, hbm_bytes=float(2 * Nm * D * 2)),
        dict(key="gemm2_prenet1", cls="gemm2_kernel", name="gemm2_kernel W-stationary + ReLU: decoder prenet 1 (51200 x 256 x 128)",
             bound="tensor", work=2.0 * Nm * D * M, per_step=1, fn=prenet1, keep=keep, hbm_bytes=float(Nm * M * 2 + Nm * D * 2)),
        dict(key="gemm2_heads", cls="gemm2_kernel", name="gemm2_kernel fp32-output plan: mel head (51200 x 80 x 256; the stop "
             "column rides along in the forward)", bound="tensor", work=2.0 * Nm * M * D, per_step=1, fn=heads, keep=keep,
             hbm_bytes=float(Nm * D * 2 + Nm * M * 6)),
        dict(key="gemm2_conv_first", cls="gemm2_kernel", name="gemm2_kernel streamed W: PostNet Conv1d 80 -> 256 k5 + BN + tanh",
             bound="tensor", work=2.0 * Nm * D * M * KS, per_step=1, fn=conv_first, keep=keep,
             hbm_bytes=float(Nm * M * 2 + Nm * D * 2)),
        dict(key="gemm2_conv_last", cls="gemm2_kernel", name="gemm2_kernel fp32-output plan: PostNet Conv1d 256 -> 80 k5 + BN + residual",
             bound="tensor", work=2.0 * Nm * D * M * KS, per_step=1, fn=conv_last, keep=keep,
             hbm_bytes=float(Nm * D * 2 + Nm * M * 8)),
    ]
    return extra + [
        dict(key="gemm2_qkv", cls="gemm2_kernel", name="gemm2_kernel W-stationary: decoder QKV projection (51200 x 768 x 256, bias)",
             bound="tensor", work=2.0 * Nm * 3 * D * D, per_step=3, fn=qkv_proj, keep=keep,
             hbm_bytes=float(Nm * D * 2 + Nm * 3 * D * 2)),
        dict(key="gemm2_out_ln", cls="gemm2_kernel", name="gemm2_kernel W-stationary + LN on CTA pairs: attention out-projection + residual + "
             "LayerNorm (51200 x 256 x 256)", bound="tensor", work=2.0 * Nm * D * D, per_step=6, fn=out_ln, keep=keep,
             hbm_bytes=float(3 * Nm * D * 2)),
        dict(key="gemm2_conv", cls="gemm2_kernel", name="gemm2_kernel streamed W on CTA pairs: PostNet Conv1d 256 -> 256 k5 + BN + tanh "
             "(51200 x 256 x 1280)", bound="tensor", work=2.0 * Nm * D * D * KS, per_step=3, fn=conv, keep=keep,
             hbm_bytes=float(2 * Nm * D * 2)),
        dict(key="ffn_fused", cls="ffn_kernel", name="ffn_kernel decoder FFN block (51200 rows, 256 -> 1024 -> 256, bias+ReLU, residual, "
             "LayerNorm; CTA pairs, hidden activation in tensor memory)", bound="tensor",
             work=4.0 * Nm * F * D, per_step=3, fn=gemm, keep=keep),
        dict(key="attn2_self", cls="attn2_kernel", name="attn2_kernel decoder causal self-attention (B x 4 heads, 800x800x64, "
             "causal-skipped flops)", bound="tensor", work=0.5 * 4.0 * Nm * Tm * D, per_step=3, fn=attn, keep=keep),
        dict(key="attn3_cross", cls="attn3_kernel", name="attn3_kernel cross-attention + alignment store (B x 4 heads, 800x150, "
             "algorithmic bytes = fp32 probabilities + bf16 q/k/v/o)", bound="hbm",
             work=float(B * H * Tm * Tp * 4 + 2 * Nm * D * 2 + B * Tp * 2 * D * 2), per_step=3, fn=cross, keep=keep),
    ]


def ncu_traffic(key):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` captures."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(path):
        return None
    with open(path) as f:
        return json.load(f).get(key)


def run_ours(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # one process per GPU on a shared host: give every rank its own slice of the cores, so the feeder / fetcher
        # threads of the ranks do not migrate over each other (the e2e path is host-side bound at N = 8)
        try:
            cores = sorted(os.sched_getaffinity(0))
            per = len(cores) // world
            if per >= 1:
                os.sched_setaffinity(0, cores[local * per:(local + 1) * per])
        except (AttributeError, OSError):
            pass
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # rank 0 prints ONE JSON line on stdout: NCCL's banner / debug lines go to stderr whatever NCCL_DEBUG says
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)

    from transformertts_b200 import TransformerTTS, _ext
    lib = _ext.load()
    torch.manual_seed(1234)
    model = TransformerTTS(device=dev).eval()          # random init, reference init laws (model/module.py:21-25)
    with torch.no_grad():
        model.linear2.linear.bias.fill_(-1.0e4)        # never-stop bias for the AR throughput runs
    model.refresh_weights()

    B, Tp, Tm = CFG2["B"], CFG2["Tp"], CFG2["Tm"]
    ph, mel, pl, ml = make_batch(B, Tp, Tm, seed=1 + rank)            # each rank: its own 64 utterances
    valid, padded = int(ml.sum()), B * Tm
    hph, hmel, hpl, hml = (t.pin_memory() for t in (ph, mel, pl, ml))
    dph, dmel, dpl, dml = (t.to(dev) for t in (ph, mel, pl, ml))
    flush = L2Flusher(dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- device-resident forward: the headline `value`.  The module replays the forward as a CUDA graph
    # (TransformerTTS.use_cuda_graph: ~50 launches on two streams captured once per shape / input buffers)
    model.use_cuda_graph = not args.no_graph
    if model.use_cuda_graph:                           # inputs resident in the graph's own static buffers: no per-step copy
        gin = model.graph_input_buffers(B, Tp, Tm)
        for dst, src in zip(gin, (dph, dmel, dpl, dml)):
            dst.copy_(src)
        dph, dmel, dpl, dml = gin
    for _ in range(max(args.warmup, 3)):
        out = model(dph, dmel, dpl, dml)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    n0 = lib.tts_launch_count() + model.graph_launches
    step_ms = []
    for _ in range(args.steps):
        flush()                                        # L2 flush between timed iterations (outside the events)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = model(dph, dmel, dpl, dml)
        b.record()
        b.synchronize()
        step_ms.append(a.elapsed_time(b))
    launches = int(lib.tts_launch_count() + model.graph_launches - n0)   # stream launches + kernels run by graph replays
    barrier()
    ms_step = max_over_ranks(sum(step_ms) / len(step_ms))
    total_valid = sum_over_ranks(valid)
    total_padded = padded * world
    value = total_valid / (ms_step / 1e3)

    # ---- end to end through the public API with host buffers: the package's own feeder / fetcher (transformertts_b200.data:
    # SURVEY 8(f) rank 3).  Every step: collated pinned host batch -> device on a copy stream, forward through the module,
    # pred | post | stop -> pinned host with ONE copy on a third stream; double-buffered, so the copies of neighbouring
    # steps overlap the forward.  Timed by wall clock around K steps incl. the final drain; median of 3 runs.
    from transformertts_b200 import TransformerTTSLoss
    from transformertts_b200.data import OutputFetcher, PinnedFeeder
    host_f32 = {"phoneme": hph, "melspec": hmel, "phoneme_lens": hpl, "melspec_lens": hml}
    host_b16 = dict(host_f32, melspec=hmel.to(torch.bfloat16).pin_memory())   # what data.collate_fn(mel_dtype=bf16) yields
    crit = TransformerTTSLoss(stop_weight=8.0).to(dev)
    loss_host = torch.empty(4).pin_memory()

    def e2e_variant(host_batch, out_dtype=torch.float32, with_align=False, loss_only=False, compute=True, check=False):
        fetch = None if loss_only else OutputFetcher(dev, dtype=out_dtype, with_alignments=with_align)
        static = {"o": None}
        bytes_io = [0, 0]
        feeder = PinnedFeeder((), dev)                   # one feeder per variant: its two device slots keep their addresses
        # graph replay writes into static output buffers: safe where the consumer (the loss) runs on the same stream; the
        # variants that copy outputs to the host on another stream keep stream launches (fresh outputs every step)
        graph = loss_only and not args.no_graph

        def run(n):
            model.use_cuda_graph = graph
            feeder.batches = (host_batch for _ in range(n))
            last = None
            for b in feeder:
                if compute or static["o"] is None:
                    o = model(b["phoneme"], b["melspec"], b["phoneme_lens"], b["melspec_lens"])
                    static["o"] = o
                else:
                    o = static["o"]                       # copy-only probe: the same device buffers travel every step
                if loss_only:
                    crit(o, b["melspec"], b["melspec_lens"])
                    loss_host.copy_(crit.values, non_blocking=True)       # total, pred_mel, post_mel, stop
                    bytes_io[1] = 16
                else:
                    last = fetch(o)
                    bytes_io[1] = fetch.d2h_bytes
                bytes_io[0] = feeder.h2d_bytes
            if fetch is not None:
                fetch.wait()
            torch.cuda.synchronize()
            return last

        run(3)
        runs = []
        for _ in range(3):
            barrier()
            t0 = time.perf_counter()
            last = run(args.steps)
            runs.append((time.perf_counter() - t0) / args.steps)
        barrier()
        if check:                                         # the fetched host tensors equal the device-resident run
            for k in ("pred_melspec", "post_melspec", "pred_stop"):
                assert torch.equal(last[k], out[k].cpu()), f"e2e result {k} differs from the device-resident run"
        sec = max_over_ranks(sorted(runs)[1])
        return {"value": total_valid / sec, "ms_per_step": sec * 1e3, "h2d_bytes_per_step": bytes_io[0],
                "d2h_bytes_per_step": bytes_io[1], "runs_ms_per_step": [x * 1e3 for x in runs]}

    def e2e_validation_step(host_batch):
        """harness.LightningModule.validation_step (the statement-by-statement stand-in of lightning_module.py:102-130)
        driven as Trainer.validate drives it: PinnedFeeder -> prepare_batch -> forward (graph replay) -> criterion ->
        loss['total'].item() EVERY step (a host sync per step, as in the reference)."""
        from transformertts_b200.harness import LightningModule
        cfg = {"model": {}, "loss": {"stop_weight": 8.0}, "training": {"num_epochs": 300, "teacher_forcing_mode": "linear",
                                                                       "log_interval": 10 ** 9}}
        lm = LightningModule.__new__(LightningModule)    # the module under test is the bench's model, not a fresh one
        lm.device, lm.model, lm.criterion, lm.config, lm.exp_dir = dev, model, crit, cfg, None
        lm.train_losses, lm.valid_losses, lm.example_batch, lm.log_interval = [], [], None, 10 ** 9
        lm.current_epoch, lm.total_train_batches, lm.total_valid_batches = 0, 0, 0
        lm.plots, lm.log_fn, lm.logged = {}, None, {}
        feeder = PinnedFeeder((), dev)

        def run(n):
            model.use_cuda_graph = not args.no_graph
            feeder.batches = (host_batch for _ in range(n))
            lm.valid_losses.clear()
            for i, b in enumerate(feeder):
                lm.validation_step(b, i + 1)              # batch_idx >= 1: the plots of batch 0 are UI, out of scope
            torch.cuda.synchronize()
            return sum(lm.valid_losses) / len(lm.valid_losses)

        run(3)
        runs = []
        for _ in range(3):
            barrier()
            t0 = time.perf_counter()
            mean_loss = run(args.steps)
            runs.append((time.perf_counter() - t0) / args.steps)
        barrier()
        sec = max_over_ranks(sorted(runs)[1])
        return {"value": total_valid / sec, "ms_per_step": sec * 1e3, "h2d_bytes_per_step": feeder.h2d_bytes,
                "d2h_bytes_per_step": 4, "runs_ms_per_step": [x * 1e3 for x in runs], "mean_total_loss": mean_loss}

    # headline: the reference's own step around forward -- LightningModule.validation_step (lightning_module.py:102-130):
    # batch to the device, forward, criterion on the device, the loss scalars back to the host (`.item()`); the outputs
    # and alignments stay on the device (utils/plot.py pulls single maps once per epoch)
    e2e_main = e2e_validation_step(host_f32)
    e2e_outputs = e2e_variant(host_b16, check=True)      # round-1 definition: pred | post | stop to the host every step
    e2e_variants = {"outputs_f32_d2h": e2e_outputs,
                    "loss_scalars_async": e2e_variant(host_f32, loss_only=True)}   # no per-step host sync (r02's headline)
    if not args.no_extra:
        e2e_variants.update({
            "outputs_f32_d2h_f32_mel_in": e2e_variant(host_f32, check=True),
            "outputs_bf16_packed_d2h": e2e_variant(host_b16, out_dtype=torch.bfloat16),
            "outputs_and_alignments_d2h": e2e_variant(host_b16, with_align=True),
            "copy_only_probe": e2e_variant(host_b16, compute=False),
        })
    clocks = sampler.stop() if rank == 0 else None
    model.use_cuda_graph = False                         # the remaining sections time single calls of changing shapes

    # ---- configs[0] (B=16, Tp=150, Tm=800) on the GPU: the same-config counterpart of the CPU arm's step
    c1 = [t.to(dev) for t in make_batch(CFG1["B"], CFG1["Tp"], CFG1["Tm"], seed=1)]
    ms1 = max_over_ranks(time_kernel(lambda: model(*c1), 10, flush))
    cfg1 = {"workload": "configs[0]: B=16, Tp=150, Tm=800, alignments returned (the CPU arm's step, same batch)",
            "ms_per_forward": ms1, "valid_frames_per_s": world * float(c1[3].sum()) / (ms1 / 1e3)}

    # ---- AR decode (configs[2]): B=1 and B=32, Tp=150, 1000 frames, never-stop bias; whole inference() call
    ar = {}
    if not args.no_ar:
        for Bar in (1, 32):
            aph, _, apl, _ = make_batch(Bar, Tp, 8, seed=100 + rank)
            haph, hapl = aph.pin_memory(), apl.pin_memory()
            aph, apl = aph.to(dev), apl.to(dev)
            model.inference(aph, apl, max_len=1001)
            torch.cuda.synchronize()
            ts = []
            for _ in range(3):
                flush()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                r = model.inference(aph, apl, max_len=1001)
                b.record()
                b.synchronize()
                ts.append(a.elapsed_time(b))
            ts.sort()
            T = r["pred_melspec"].shape[1]
            ms = max_over_ranks(ts[1])
            # end to end: host phonemes in (pinned), mel_pre / mel_post / stop back in pinned host memory, wall clock
            hres = {k: torch.empty_like(v, device="cpu").pin_memory() for k, v in r.items()}
            es = []
            for _ in range(3):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                r2 = model.inference(haph, hapl, max_len=1001)
                for k, v in r2.items():
                    hres[k].copy_(v, non_blocking=True)
                torch.cuda.synchronize()
                es.append(time.perf_counter() - t0)
            e_s = max_over_ranks(sorted(es)[1])
            # algorithmic bytes: self-KV reads sum_t B*L*2*t*d*2 + cross-KV reads + weights once per step
            kv = Bar * L * 2 * D * 2 * (T * (T - 1) / 2) + T * Bar * L * 2 * Tp * D * 2
            wbytes = T * 5.75e6
            ar[f"B{Bar}"] = {"frames_per_s": world * Bar * T / (ms / 1e3), "ms_total": ms, "frames": T,
                             "us_per_step": 1e3 * ms / T,
                             "achieved_gbs": (kv + wbytes) / (ms / 1e3) / 1e9,
                             "frac_of_hbm_peak": (kv + wbytes) / (ms / 1e3) / 1e9 / measured_peaks()["hbm_gbs"],
                             "e2e": {"value": world * Bar * T / e_s, "unit": "mel frames/s", "ms_total": e_s * 1e3,
                                     "h2d_bytes_per_step": int(haph.numel() * 8 + hapl.numel() * 8),
                                     "d2h_bytes_per_step": int(sum(v.numel() * 4 for v in hres.values())),
                                     "note": "one step = one inference() call: pinned host phonemes in, encoder + 1000 "
                                             "decode steps + PostNet, pred / post / stop copied to pinned host memory"}}

    # ---- batch-sharded AR decode (configs[3]): 512 utterances dealt in contiguous blocks to the ranks
    # (sharding.shard_bounds), micro-batches of 32 / 64 / 128 (each one inference() call), 1000 frames each, no collective
    ar_sharded = None
    if not args.no_ar and not args.no_extra:
        from transformertts_b200.sharding import shard_bounds, micro_batches
        NU = 512
        uph, _, upl, _ = make_batch(NU, Tp, 8, seed=300)               # the same 512 utterances on every rank
        lo, hi = shard_bounds(NU, world, rank)
        uph, upl = uph.to(dev), upl.to(dev)
        ar_sharded = {"utterances": NU, "frames_each": 1000, "per_rank": hi - lo, "scaling": "strong"}
        for mb in (32, 64, 128):
            parts = micro_batches(lo, hi, mb)

            def run_block():
                n = 0
                for s_, e_ in parts:
                    tp_ = int(upl[s_:e_].max())
                    r_ = model.inference(uph[s_:e_, :tp_].contiguous(), upl[s_:e_], max_len=1001)
                    n += (e_ - s_) * r_["pred_melspec"].shape[1]
                return n
            s_, e_ = parts[0]
            model.inference(uph[s_:e_, :int(upl[s_:e_].max())].contiguous(), upl[s_:e_], max_len=1001)   # warm-up
            barrier()
            flush()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            nframes = run_block()
            b.record()
            b.synchronize()
            ms = max_over_ranks(a.elapsed_time(b))
            tot = sum_over_ranks(float(nframes))
            ar_sharded[f"micro_batch_{mb}"] = {"frames_per_s": tot / (ms / 1e3), "ms_total": ms, "frames": int(tot),
                                               "calls_per_rank": len(parts)}

    # ---- long-utterance stress (configs[4]): Tp=400, Tm=2000, B=1 and B=16, alignments materialised
    long_utt = None
    if not args.no_extra:
        long_utt = {}
        for Bl in (1, 16):
            lph, lmel, lpl, lml = (t.to(dev) for t in make_batch(Bl, 400, 2000, seed=500 + rank))
            fn = lambda: model(lph, lmel, lpl, lml)
            ms = max_over_ranks(time_kernel(fn, 5, flush))
            lfl = algorithmic_flops(Bl, 400, 2000, True)
            long_utt[f"B{Bl}"] = {"ms_per_forward": ms, "valid_frames_per_s": sum_over_ranks(float(lml.sum())) / (ms / 1e3),
                                  "padded_frames_per_s": world * Bl * 2000 / (ms / 1e3),
                                  "achieved_tflops": lfl / (ms / 1e3) / 1e12,
                                  "align_bytes": 3 * Bl * H * 2000 * 400 * 4}
            del lph, lmel, lpl, lml
        torch.cuda.empty_cache()

    # ---- roofline of the dominant kernel, timed alone on this stream (burst peak)
    peaks = measured_peaks()
    roofline, kernels = None, []
    if rank == 0:
        for c in dominant_kernel_candidates(model, B, Tp, Tm, dev):
            ms = time_kernel(c["fn"], 10, flush)
            k = {"name": c["name"], "kernel": c["cls"], "bound": c["bound"], "ms_per_launch": ms,
                 "launches_per_step": c["per_step"], "share_of_step": c["per_step"] * ms / ms_step,
                 "work_per_launch": c["work"], "traffic": ncu_traffic(c["key"])}
            if "hbm_bytes" in c:      # the K = 256 GEMMs are memory bound by construction: show that roofline too
                k["hbm_frac"] = c["hbm_bytes"] / (ms / 1e3) / 1e9 / peaks["hbm_gbs"]
            if c["bound"] == "tensor":
                k.update(achieved=c["work"] / (ms / 1e3) / 1e12, peak=peaks["tf_burst"], unit="TFLOP/s")
            else:
                k.update(achieved=c["work"] / (ms / 1e3) / 1e9, peak=peaks["hbm_gbs"], unit="GB/s")
            k["frac"] = k["achieved"] / k["peak"]
            kernels.append(k)
        # The dominant kernel is the __global__ function with the largest share of the step: its shapes are aggregated
        # (gemm2_kernel runs every projection / conv of the decoder side: nine shapes, four plans), achieved = the
        # algorithmic work of one step's launches / their summed duration.
        classes = {}
        for k in kernels:
            c = classes.setdefault(k["kernel"], {"share": 0.0, "work": 0.0, "ms": 0.0, "bound": k["bound"], "peak": k["peak"],
                                                 "unit": k["unit"], "traffic": 0, "traffic_ok": True, "shapes": 0})
            c["share"] += k["share_of_step"]
            c["work"] += k["work_per_launch"] * k["launches_per_step"]
            c["ms"] += k["ms_per_launch"] * k["launches_per_step"]
            c["shapes"] += 1
            if k["traffic"] is None:
                c["traffic_ok"] = False
            else:
                c["traffic"] += k["traffic"] * k["launches_per_step"]
        name, top = max(classes.items(), key=lambda kv: kv[1]["share"])
        scale = 1e12 if top["bound"] == "tensor" else 1e9
        achieved = top["work"] / (top["ms"] / 1e3) / scale
        roofline = {"bound": top["bound"], "kernel": f"{name} ({top['shapes']} shapes of the step aggregated, "
                                                     f"{100 * top['share']:.0f} % of the step)",
                    "achieved": achieved, "peak": top["peak"], "unit": top["unit"], "frac": achieved / top["peak"],
                    "traffic": top["traffic"] if top["traffic_ok"] else None,
                    "traffic_note": "dram bytes of one step's launches of this kernel (sum over the captured shapes)"
                                    if top["traffic_ok"] else "not every shape of the class has an ncu capture",
                    "classes": {n: {"share_of_step": c["share"], "frac": c["work"] / (c["ms"] / 1e3) / (1e12 if c["bound"] == "tensor" else 1e9) / c["peak"],
                                    "bound": c["bound"]} for n, c in classes.items()},
                    "peak_source": peaks["source"] + " (burst figure: kernels timed alone between L2 flushes)"}
    fl = algorithmic_flops(B, Tp, Tm, True) * world
    align_bytes = 3 * B * H * Tm * Tp * 4
    # write-only HBM bandwidth, probed live (torch fill of 1 GiB, best of 5): the alignment store is pure write
    # traffic and pure writes run well below the copy figure in MEASURED_PEAKS.json
    wbuf = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
    wbest = 1e9
    for _ in range(6):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); wbuf.zero_(); b.record(); b.synchronize()
        wbest = min(wbest, a.elapsed_time(b))
    del wbuf
    write_gbs = (1 << 30) / (wbest / 1e3) / 1e9
    t_tensor = fl / world / (peaks["tf_sustained"] * 1e12) * 1e3               # ms per GPU
    t_store = align_bytes / (peaks["hbm_gbs"] * 1e9) * 1e3
    step_roofline = {"flops_causal_skipped": fl, "achieved_tflops": fl / (ms_step / 1e3) / 1e12,
                     "peak_tflops_sustained": peaks["tf_sustained"] * world,
                     "frac": fl / (ms_step / 1e3) / 1e12 / (peaks["tf_sustained"] * world),
                     "align_store_bytes": align_bytes * world,
                     "roofline_ms": t_tensor + t_store, "frac_with_align_store": (t_tensor + t_store) / ms_step,
                     "hbm_write_only_gbs_probe": write_gbs,
                     # whole-forward DRAM bytes per GPU from one ncu pass over a graph-free forward (profiles/r03_step_dram.txt)
                     # against the bytes the step must move: inputs (phonemes, mel, lens) + outputs + alignments
                     "step_dram_bytes": ncu_traffic("step_forward_cfg2"),
                     "algorithmic_bytes": int(B * Tp * 8 + 2 * B * 8 + B * Tm * 80 * 4 + 2 * B * Tm * 80 * 4 + B * Tm * 4
                                              + align_bytes),
                     "note": "roofline_ms = tensor term at the sustained bf16 peak + alignment store at the HBM "
                             "copy peak (the kernels are serial); the write-only probe shows what pure stores reach"}

    # ---- CPU baseline on this box's host cores (rank 0, N=1 only): bounded sample of the same workload
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu, cpu_ar = cpu_baselines(reps=3, ar_frames=200)
        for k, v in cpu_ar.items():
            if k in ar:
                ar[k]["cpu_baseline"] = v

    if rank == 0:
        line = {
            "metric": "mel frames/sec (teacher-forced forward)", "value": value, "unit": "valid mel frames/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": our_config(world), "run": run_notes("ours", graph=not args.no_graph),
            "frames": {"valid_frames_per_step": total_valid, "padded_frames_per_step": total_padded,
                       "padded_frames_per_s": total_padded / (ms_step / 1e3)},
            "e2e": dict(e2e_main, unit="valid mel frames/s",
                        note="median of 3 runs of K steps of transformertts_b200.harness.LightningModule.validation_step "
                             "(lightning_module.py:102-130) as Trainer.validate drives it: every step a collated pinned host "
                             "batch (fp32 mel) -> device on a copy stream (data.PinnedFeeder, double-buffered), "
                             "TransformerTTS.forward with alignments (graph replay), TransformerTTSLoss on the device, "
                             "loss['total'].item() -- one host sync per step, as in the reference; outputs stay on the device. "
                             "e2e_outputs_to_host is the other first-class definition (pred | post | stop fp32 copied to the host "
                             "every step, checked bit-equal against the device-resident run): it is host-copy bound at N = 8. "
                             "e2e_variants.loss_scalars_async is round 2's headline (no per-step sync); copy_only_probe runs "
                             "the copies without the forward"),
            "e2e_outputs_to_host": dict(e2e_outputs, unit="valid mel frames/s"),
            "e2e_variants": e2e_variants, "cfg1_b16": cfg1,
            "gpu_launches": launches,
            "roofline": roofline, "kernels": kernels, "step_roofline": step_roofline,
            "ar_decode": ar, "ar_sharded_512": ar_sharded, "long_utterance": long_utt,
            "cpu_baseline": cpu, "clocks": clocks,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-ar", action="store_true", help="skip the AR-decode section")
    ap.add_argument("--no-graph", action="store_true", help="stream launches instead of CUDA-graph replay of the forward")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline")
    ap.add_argument("--no-extra", action="store_true",
                    help="skip the configs[3] (512 sharded AR utterances) and configs[4] (long utterances) sections")
    args = ap.parse_args()
    # stdout carries ONE JSON line: while the run is in progress file descriptor 1 points at stderr, so banners that
    # libraries write straight to fd 1 (NCCL's version line) cannot get in front of it
    global _STDOUT_FD
    sys.stdout.flush()
    _STDOUT_FD = os.dup(1)
    os.dup2(2, 1)
    try:
        if args.impl == "reference":
            return run_reference(args)
        return run_ours(args)
    finally:
        sys.stdout.flush()
        os.dup2(_STDOUT_FD, 1)


_STDOUT_FD = None


def emit(line):
    """Print the result line on the real stdout."""
    sys.stdout.flush()
    if _STDOUT_FD is not None:
        os.write(_STDOUT_FD, (json.dumps(line) + "\n").encode())
    else:
        print(json.dumps(line))


if __name__ == "__main__":
    sys.exit(main())
