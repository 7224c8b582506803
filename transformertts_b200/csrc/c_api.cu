// extern "C" boundary (include/tts_b200.h) and the orchestration of the hot path:
// TransformerTTS.forward (model/model.py:260-320) and .inference (model/model.py:323-394) as sequences
// of the kernels in gemm2.cu / ffn.cu / attention2.cu / attention3.cu / decode.cu over a caller-provided workspace.
#include "../../include/tts_b200.h"

#include <math.h>
#include <stdlib.h>
#include <stdio.h>
#include <string.h>

#include "decode.cuh"
#include "kernels.h"

using namespace tts;

static thread_local char g_err[512] = "";

static int fail(const char* msg) {
  snprintf(g_err, sizeof(g_err), "%s", msg);
  return 1;
}
#define TRY(expr)                                  \
  do {                                             \
    const char* _e = (expr);                       \
    if (_e != nullptr) return fail(_e);            \
  } while (0)

// ------------------------------------------------------------------ small kernels
#include <atomic>
static std::atomic<unsigned long long> g_launches{0};
namespace tts {

bool pdl_enabled() {
  static const bool on = getenv("TTS_NO_PDL") == nullptr;
  return on;
}

void note_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

__global__ void embed_kernel(const int64_t* __restrict__ ph, const uint4* __restrict__ table, uint4* __restrict__ out,
                             int n_tok, int n_phon) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;      // one 16-byte chunk (8 channels) per thread
  if (i >= n_tok * 32) return;
  const int tok = i >> 5, ch = i & 31;
  long long id = ph[tok];
  id = id < 0 ? 0 : (id >= n_phon ? n_phon - 1 : id);
  out[i] = __ldg(table + id * 32 + ch);
}

__global__ void cast_mel_kernel(const float* __restrict__ in, bf16* __restrict__ out, size_t n) {
  const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i + 4 <= n) {
    const float4 v = *reinterpret_cast<const float4*>(in + i);
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    uint2 u;
    u.x = *reinterpret_cast<uint32_t*>(&a);
    u.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(out + i) = u;
  } else {
    for (size_t j = i; j < n; ++j) out[j] = __float2bfloat16(in[j]);
  }
}

__global__ void lens_kernel(const int64_t* __restrict__ in, int* __restrict__ out, int n, int lo, int hi) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    long long v = in[i];
    out[i] = (int)(v < lo ? lo : (v > hi ? hi : v));
  }
}

const char* embed_launch(const int64_t* phoneme, const bf16* table, bf16* out, int n_tok, int n_phon, cudaStream_t s) {
  const int n = n_tok * 32;
  embed_kernel<<<(n + 255) / 256, 256, 0, s>>>(phoneme, reinterpret_cast<const uint4*>(table),
                                                reinterpret_cast<uint4*>(out), n_tok, n_phon);
  note_launch();
  return cudaGetLastError() == cudaSuccess ? nullptr : "embedding kernel launch failed";
}
const char* cast_mel_launch(const float* mel, bf16* out, size_t n, cudaStream_t s) {
  const size_t thr = (n + 3) / 4;
  cast_mel_kernel<<<(unsigned)((thr + 255) / 256), 256, 0, s>>>(mel, out, n);
  note_launch();
  return cudaGetLastError() == cudaSuccess ? nullptr : "cast kernel launch failed";
}
const char* lens_launch(const int64_t* in, int* out, int n, int lo, int hi, cudaStream_t s) {
  lens_kernel<<<(n + 127) / 128, 128, 0, s>>>(in, out, n, lo, hi);
  note_launch();
  return cudaGetLastError() == cudaSuccess ? nullptr : "lens kernel launch failed";
}

}  // namespace tts

// ------------------------------------------------------------------ argument builders
static GemmArgs linear_args(const bf16* A, int M, int K, const void* W, const float* bias, int N) {
  GemmArgs g;
  memset(&g, 0, sizeof(g));
  g.A = A; g.B = 1; g.T = M; g.C = K;
  g.W = reinterpret_cast<const bf16*>(W); g.N = N;
  g.taps = 1; g.tap_off = 0; g.kc_per_tap = (K + 63) / 64;
  g.bias = bias;
  return g;
}
static GemmArgs conv_args(const bf16* x, int B, int T, int C, const void* W, const float* bias, int N, int taps,
                          int tap_off) {
  GemmArgs g;
  memset(&g, 0, sizeof(g));
  g.A = x; g.B = B; g.T = T; g.C = C;
  g.W = reinterpret_cast<const bf16*>(W); g.N = N;
  g.taps = taps; g.tap_off = tap_off; g.kc_per_tap = (C + 63) / 64;
  g.bias = bias;
  return g;
}

// carve 1 KiB aligned regions out of the caller's workspace
struct Carver {
  uint8_t* base; size_t off; size_t cap;
  void* take(size_t bytes) {
    void* p = base + off;
    off += (bytes + 1023) & ~size_t(1023);
    return p;
  }
};

struct Workspace {
  int* ph_len; int* mel_len;
  bf16 *e0, *e1, *qkv_e, *att_e, *ffn_e;         // encoder side  [Np][...]
  bf16 *mel_in, *d0, *d1, *qkv_d, *att_d, *ffn_d, *cq, *ckv, *mel_pre16;
  size_t total;
};

static Workspace carve(void* base, int B, int Tp, int Tm) {
  const size_t Np = (size_t)B * Tp, Nm = (size_t)B * (Tm > 0 ? Tm : 1);
  Carver c{reinterpret_cast<uint8_t*>(base), 0, 0};
  Workspace w;
  w.ph_len = (int*)c.take(sizeof(int) * B);
  w.mel_len = (int*)c.take(sizeof(int) * B);
  w.e0 = (bf16*)c.take(Np * kD * 2);
  w.e1 = (bf16*)c.take(Np * kD * 2);
  w.qkv_e = (bf16*)c.take(Np * 3 * kD * 2);
  w.att_e = (bf16*)c.take(Np * kD * 2);
  w.ffn_e = (bf16*)c.take(Np * kF * 2);
  w.ckv = (bf16*)c.take(3 * Np * 2 * kD * 2);           // cross K/V of the memory, one [Np][512] block per decoder layer
  w.mel_in = (bf16*)c.take(Nm * kMel * 2);
  w.mel_pre16 = (bf16*)c.take(Nm * kMel * 2);
  w.d0 = (bf16*)c.take(Nm * kD * 2);
  w.d1 = (bf16*)c.take(Nm * kD * 2);
  w.qkv_d = (bf16*)c.take(Nm * 3 * kD * 2);
  w.att_d = (bf16*)c.take(Nm * kD * 2);
  w.cq = (bf16*)c.take(Nm * kD * 2);
  w.ffn_d = (bf16*)c.take(Nm * kF * 2);
  w.total = c.off;
  return w;
}

extern "C" int tts_abi_version(void) { return TTS_ABI_VERSION; }
extern "C" const char* tts_last_error(void) { return g_err; }
extern "C" int tts_decode_max_clusters(int rows_per_cluster, int cluster_size) {
  return decode_max_clusters(rows_per_cluster, cluster_size);
}
extern "C" int tts_decode_capacity(void) { return decode_capacity(); }
extern "C" uint64_t tts_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
extern "C" size_t tts_workspace_bytes(int B, int Tp, int Tm) {
  if (B <= 0 || Tp <= 0 || Tm < 0) return 0;
  return carve(nullptr, B, Tp, Tm).total;
}

// ------------------------------------------------------------------ operator-level entry points
extern "C" int tts_op_linear(const void* A, const void* W, const float* bias, int M, int N, int K, int act,
                             const float* pe, float alpha, int pe_period, const void* residual, const float* ln_g,
                             const float* ln_b, void* out_bf16, float* out_f32, void* stream) {
  if (K % 64 != 0) return fail("tts_op_linear: K must be a multiple of 64");
  GemmArgs g = linear_args((const bf16*)A, M, K, W, bias, N);
  g.act = act;
  g.pe = pe; g.alpha = alpha; g.pe_period = pe_period;
  g.residual = (const bf16*)residual; g.ld_res = N;
  g.ln_g = ln_g; g.ln_b = ln_b;
  g.out_bf16 = (bf16*)out_bf16; g.ld_bf16 = N;
  g.out_f32 = out_f32; g.ld_f32 = N;
  TRY(gemm_launch(g, (cudaStream_t)stream));
  return 0;
}

extern "C" int tts_op_linear_chain(const void* A, const void* W1, const float* b1, int M, int K1, int act1, const void* residual,
                                   const float* ln_g, const float* ln_b, const void* W2, const float* b2, int act2,
                                   const float* pe, float alpha, int pe_period, void* out1_bf16, void* out2_bf16,
                                   void* stream) {
  if (K1 % 64 != 0) return fail("tts_op_linear_chain: K1 must be a multiple of 64");
  GemmArgs g = linear_args((const bf16*)A, M, K1, W1, b1, kD);
  g.act = act1;
  g.residual = (const bf16*)residual; g.ld_res = kD;
  g.ln_g = ln_g; g.ln_b = ln_b;
  g.out_bf16 = (bf16*)out1_bf16; g.ld_bf16 = kD;
  g.W2 = (const bf16*)W2; g.bias2 = b2; g.act2 = act2;
  g.pe = pe; g.alpha = alpha; g.pe_period = pe_period;
  g.out2 = (bf16*)out2_bf16; g.ld_out2 = kD;
  if (!gemm2_chain_supports(g)) return fail("tts_op_linear_chain: unsupported shape (K1 <= 256, N = 256, >= 129 rows, aligned)");
  TRY(gemm2_chain_launch(g, (cudaStream_t)stream));
  return 0;
}

extern "C" int tts_op_conv1d(const void* x, const void* W, const float* bias, int B, int T, int C, int N, int taps,
                             int act, const float* residual_f32, void* out_bf16, float* out_f32, void* stream) {
  GemmArgs g = conv_args((const bf16*)x, B, T, C, W, bias, N, taps, -((taps - 1) / 2));
  g.act = act;
  g.residual_f32 = residual_f32; g.ld_res32 = N;
  g.out_bf16 = (bf16*)out_bf16; g.ld_bf16 = N;
  g.out_f32 = out_f32; g.ld_f32 = N;
  TRY(gemm_launch(g, (cudaStream_t)stream));
  return 0;
}

extern "C" int tts_op_ffn(const void* x, const void* w1, const float* b1, const void* w2, const float* b2,
                          const float* ln_g, const float* ln_b, int M, void* out, void* stream) {
  FfnArgs f;
  f.x = (const bf16*)x; f.M = M;
  f.w1 = (const bf16*)w1; f.b1 = b1; f.w2 = (const bf16*)w2; f.b2 = b2;
  f.ln_g = ln_g; f.ln_b = ln_b; f.out = (bf16*)out;
  if (!ffn_supports(f)) return fail("tts_op_ffn: pointers must be 16-byte aligned and x != out");
  TRY(ffn_launch(f, (cudaStream_t)stream));
  return 0;
}

extern "C" int tts_op_attention(const void* q, int ldq, const void* k, int ldk, const void* v, int ldv, int B, int Tq,
                                int Tk, const int32_t* klen, int causal, void* out, float* probs, void* stream) {
  AttnArgs a;
  a.Q = (const bf16*)q; a.ldq = ldq;
  a.K = (const bf16*)k; a.ldk = ldk;
  a.V = (const bf16*)v; a.ldv = ldv;
  a.O = (bf16*)out; a.ldo = kD;
  a.probs = probs; a.klen = klen;
  a.B = B; a.Tq = Tq; a.Tk = Tk; a.causal = causal;
  TRY(attn_launch(a, (cudaStream_t)stream));
  return 0;
}

// ------------------------------------------------------------------ shared pieces of forward / inference
static const char* run_attention(const bf16* q, int ldq, const bf16* k, const bf16* v, int ldkv, bf16* out, float* probs,
                                 const int* klen, int B, int Tq, int Tk, int causal, cudaStream_t s) {
  AttnArgs a;
  a.Q = q; a.ldq = ldq; a.K = k; a.ldk = ldkv; a.V = v; a.ldv = ldkv;
  a.O = out; a.ldo = kD; a.probs = probs; a.klen = klen;
  a.B = B; a.Tq = Tq; a.Tk = Tk; a.causal = causal;
  return attn_launch(a, s);
}

// TTS_GEMM_CHAIN=1: run  out-projection + LayerNorm -> cross-attention Q projection  and  prenet 1 -> prenet 2 + PE  as
// ONE chained launch each (gemm2.cu CHAIN plan: the intermediate activation stays in tensor memory).  Measured and NOT the
// default (profiles/r02_gemm_chain_ab.txt): alone, between L2 flushes, a chained launch beats its two kernels (41.0 vs
// 26.2 + 19.2 us; 42.2 vs 15.9 + 32.6 us), but inside the forward the two-kernel sequence wins (1.128 vs 1.171 ms per
// cfg-2 forward): the 26 MB intermediate is L2-resident for the second kernel anyway, and the chained tile is one serial
// chain (GEMM 1 -> two-pass LayerNorm epilogue -> GEMM 2 -> epilogue 2) on a 3-stage ring where the separate kernels
// overlap tile i's epilogue with tile i+1's MMAs on a 7-stage ring.  Kept under test (tests/test_gpu_ops.py,
// tests/test_gpu_pair_plans.py) as the starting point for a chained plan with two tiles in flight.
static bool chain_enabled() {
  static const char* env = getenv("TTS_GEMM_CHAIN");
  static const bool on = env != nullptr && env[0] == '1';
  return on;
}

// out = LayerNorm(res + A . W^T + b)
static const char* proj_res_ln(const bf16* A, int M, int K, const void* W, const float* b, const bf16* res,
                               const TtsLnW& ln, bf16* out, cudaStream_t s) {
  GemmArgs g = linear_args(A, M, K, W, b, kD);
  g.residual = res; g.ld_res = kD;
  g.ln_g = ln.g; g.ln_b = ln.b;
  g.out_bf16 = out; g.ld_bf16 = kD;
  return gemm_launch(g, s);
}

static const char* proj(const bf16* A, int M, int K, const void* W, const float* b, int N, int act, bf16* out,
                        cudaStream_t s) {
  GemmArgs g = linear_args(A, M, K, W, b, N);
  g.act = act;
  g.out_bf16 = out; g.ld_bf16 = N;
  return gemm_launch(g, s);
}

// x_out = LayerNorm(x + FFN(x)): one fused launch on CTA pairs (ffn.cu; the hidden activation stays in tensor
// memory), or with TTS_FFN_FUSE=0 the two-GEMM path through `hidden`.  Measured at cfg 2: 64.5 us against
// 37.6 + 46.1 us per decoder layer in isolation, 1.193 against 1.235 ms for the whole forward (profiles/README.md).
static const char* ffn_block(const bf16* x, int M, const TtsFfnW& w, const TtsLnW& ln, bf16* hidden, bf16* out,
                             cudaStream_t s) {
  static const char* fuse_env = getenv("TTS_FFN_FUSE");
  static const bool no_fuse = fuse_env != nullptr && fuse_env[0] == '0';
  FfnArgs f;
  f.x = x; f.M = M;
  f.w1 = (const bf16*)w.w1; f.b1 = w.b1; f.w2 = (const bf16*)w.w2; f.b2 = w.b2;
  f.ln_g = ln.g; f.ln_b = ln.b; f.out = out;
  if (!no_fuse && ffn_supports(f)) return ffn_launch(f, s);
  const char* e = proj(x, M, kD, w.w1, w.b1, kF, ACT_RELU, hidden, s);
  if (e) return e;
  return proj_res_ln(hidden, M, kF, w.w2, w.b2, x, ln, out, s);
}

// emb -> EncoderPreNet -> PE -> 3 post-norm encoder layers (model/model.py:288-292 / 345-348).
// Result (memory) is left in ws.e0.
static const char* run_encoder(const TtsWeights* w, const int64_t* phoneme, int B, int Tp, bool masked, Workspace& ws,
                               cudaStream_t s) {
  const int Np = B * Tp;
  const char* e;
  if ((e = embed_launch(phoneme, (const bf16*)w->emb, ws.e0, Np, w->n_phon, s))) return e;
  bf16* cur = ws.e0; bf16* nxt = ws.e1;
  for (int i = 0; i < 3; ++i) {                       // Conv1d k5 + folded BN, no activation
    GemmArgs g = conv_args(cur, B, Tp, kD, w->enc_conv_w[i], w->enc_conv_b[i], kD, kTaps, -(kTaps - 1) / 2);
    g.out_bf16 = nxt; g.ld_bf16 = kD;
    if ((e = gemm_launch(g, s))) return e;
    bf16* t = cur; cur = nxt; nxt = t;
  }
  {                                                   // Linear + alpha * PE  (cur == e1 -> e0)
    GemmArgs g = linear_args(cur, Np, kD, w->enc_lin_w, w->enc_lin_b, kD);
    g.pe = w->pe; g.alpha = w->alpha; g.pe_period = Tp;
    g.out_bf16 = nxt; g.ld_bf16 = kD;
    if ((e = gemm_launch(g, s))) return e;
    bf16* t = cur; cur = nxt; nxt = t;
  }
  for (int l = 0; l < 3; ++l) {
    const TtsEncLayerW& L = w->enc[l];
    if ((e = proj(cur, Np, kD, L.sa.w_in, L.sa.b_in, 3 * kD, ACT_NONE, ws.qkv_e, s))) return e;
    if ((e = run_attention(ws.qkv_e, 3 * kD, ws.qkv_e + kD, ws.qkv_e + 2 * kD, 3 * kD, ws.att_e, nullptr,
                           masked ? ws.ph_len : nullptr, B, Tp, Tp, 0, s))) return e;
    if ((e = proj_res_ln(ws.att_e, Np, kD, L.sa.w_out, L.sa.b_out, cur, L.ln1, nxt, s))) return e;
    if ((e = ffn_block(nxt, Np, L.ffn, L.ln2, ws.ffn_e, cur, s))) return e;
  }
  return cur == ws.e0 ? nullptr : "internal: encoder output buffer parity";
}

// cross-attention K/V of the memory for decoder layer l: [Np][512]
static const char* run_cross_kv(const TtsWeights* w, int l, const bf16* memory, int Np, bf16* out, cudaStream_t s) {
  const TtsAttnW& ca = w->dec[l].ca;
  return proj(memory, Np, kD, (const bf16*)ca.w_in + (size_t)kD * kD, ca.b_in + kD, 2 * kD, ACT_NONE, out, s);
}

// PostNet (model/model.py:129-135) + residual (:310): 4 x tanh(BN(conv)), BN(conv), + mel_pre
static const char* run_postnet(const TtsWeights* w, const bf16* mel16, const float* mel_pre, int B, int T, bf16* p0,
                               bf16* p1, float* mel_post, cudaStream_t s) {
  const char* e;
  GemmArgs g = conv_args(mel16, B, T, kMel, w->post_w[0], w->post_b[0], kD, kTaps, -(kTaps - 1) / 2);
  g.act = ACT_TANH; g.out_bf16 = p0; g.ld_bf16 = kD;
  if ((e = gemm_launch(g, s))) return e;
  bf16* cur = p0; bf16* nxt = p1;
  for (int i = 1; i < 4; ++i) {
    GemmArgs h = conv_args(cur, B, T, kD, w->post_w[i], w->post_b[i], kD, kTaps, -(kTaps - 1) / 2);
    h.act = ACT_TANH; h.out_bf16 = nxt; h.ld_bf16 = kD;
    if ((e = gemm_launch(h, s))) return e;
    bf16* t = cur; cur = nxt; nxt = t;
  }
  GemmArgs f = conv_args(cur, B, T, kD, w->post_w[4], w->post_b[4], kMel, kTaps, -(kTaps - 1) / 2);
  f.residual_f32 = mel_pre; f.ld_res32 = kMel;
  f.out_f32 = mel_post; f.ld_f32 = kMel;
  return gemm_launch(f, s);
}

static int check_common(const TtsWeights* w, int B, int Tp, int Tm, const void* ws, size_t ws_bytes) {
  if (w == nullptr) return fail("weights pointer is NULL");
  if (B <= 0 || Tp <= 0 || Tm <= 0) return fail("B, Tp and Tm must be positive");
  if (Tp > kPeMax || Tm > kPeMax) return fail("sequence longer than the positional-encoding table (5000)");
  if (ws == nullptr || ws_bytes < tts_workspace_bytes(B, Tp, Tm)) return fail("workspace too small");
  if ((reinterpret_cast<uintptr_t>(ws) & 1023) != 0) return fail("workspace must be 1024-byte aligned");
  return 0;
}

// ------------------------------------------------------------------ side stream for the encoder branch
// The encoder (+ the cross K/V projections of its output) and the decoder front (prenet, layer-0 self-attention
// block) are independent until the first cross-attention; tts_forward runs the encoder branch on a per-device side
// stream, forked from and joined back into the caller's stream with events, so its many small launches fill the tails
// and launch gaps of the decoder kernels.
// Measured and NOT kept (profiles/r02_fwd_split_ab.txt): the decoder stack as 2 / 3 / 4 independent sub-batches on their
// own streams (utterances are independent), hoping that the SMs one chain's kernel frees while it drains run the other
// chain's kernel: 1.13 -> 1.21 / 1.37 / 1.40 ms per cfg-2 forward -- every kernel is persistent with one CTA per SM, so
// two chains only take SMs away from each other.
#include <mutex>
// `use` is held by a tts_forward call from its fork to its join (host-side enqueue only, ~0.1 ms): the event pair and the
// side stream are shared by every caller on the device, and cudaStreamWaitEvent binds to the event's LATEST record, so
// two host threads (two models on one GPU) must not interleave their record / wait pairs.  The GPU work of the two
// calls still overlaps: only the enqueueing is serialised.
struct SideStream { cudaStream_t enc; cudaEvent_t fork, enc_join; bool ok; std::mutex use; };
static SideStream* side_stream() {
  static SideStream table[kMaxDevices] = {};
  static std::mutex mu;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return nullptr;
  std::lock_guard<std::mutex> lock(mu);
  SideStream& s = table[dev];
  if (!s.ok) {
    if (cudaStreamCreateWithFlags(&s.enc, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&s.enc_join, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    s.ok = true;
  }
  return &s;
}

// Decoder stack + heads + PostNet for utterances [b0, b0 + nb) of the batch on stream s (model/model.py:297-313).
// `enc_join` (nullable): event after which the encoder branch's memory / cross K/V are complete.
static const char* run_decoder(const TtsWeights* w, const void* mel, int flags, int B, int Tp, int Tm, int b0, int nb,
                               float* mel_pre, float* mel_post, float* stop, float* align, Workspace& ws,
                               cudaEvent_t enc_join, cudaStream_t s) {
  const size_t r0 = (size_t)b0 * Tm;                 // first decoder row of this part
  const int Nm = nb * Tm;
  const char* e;
  bf16* mel_in = ws.mel_in + r0 * kMel;
  const bf16* mel16 = mel_in;
  if (flags & TTS_FLAG_MEL_BF16) {
    mel16 = reinterpret_cast<const bf16*>(mel) + r0 * kMel;       // read in place by the prenet's TMA loads
  } else {
    if ((e = cast_mel_launch(reinterpret_cast<const float*>(mel) + r0 * kMel, mel_in, (size_t)Nm * kMel, s))) return e;
  }
  bf16 *d0 = ws.d0 + r0 * kD, *d1 = ws.d1 + r0 * kD, *qkv = ws.qkv_d + r0 * 3 * kD, *att = ws.att_d + r0 * kD;
  bf16 *cq = ws.cq + r0 * kD, *ffn = ws.ffn_d + r0 * kF, *pre16 = ws.mel_pre16 + r0 * kMel;
  const int* mel_len = ws.mel_len + b0;
  const int* ph_len = ws.ph_len + b0;
  {
    // decoder prenet on the go-frame-shifted target (model/model.py:278-279, 297): frame t reads mel[t-1]
    GemmArgs g = conv_args(mel16, nb, Tm, kMel, w->pre1_w, w->pre1_b, kD, 1, -1);
    g.act = ACT_RELU;
    // prenet 1 -> prenet 2 + PE: two launches by default; one chained launch (hidden activation stays in tensor memory)
    // only with TTS_GEMM_CHAIN=1
    g.W2 = (const bf16*)w->pre2_w; g.bias2 = w->pre2_b; g.act2 = ACT_RELU;
    g.pe = w->pe; g.alpha = w->alpha; g.pe_period = Tm;
    g.out2 = d1; g.ld_out2 = kD;
    if (chain_enabled() && gemm2_chain_supports(g)) {
      if ((e = gemm2_chain_launch(g, s))) return e;
    } else {
      g.W2 = nullptr; g.pe = nullptr; g.out2 = nullptr;
      g.out_bf16 = d0; g.ld_bf16 = kD;
      if ((e = gemm_launch(g, s))) return e;
      GemmArgs h = linear_args(d0, Nm, kD, w->pre2_w, w->pre2_b, kD);
      h.act = ACT_RELU; h.pe = w->pe; h.alpha = w->alpha; h.pe_period = Tm;
      h.out_bf16 = d1; h.ld_bf16 = kD;
      if ((e = gemm_launch(h, s))) return e;
    }
  }
  bf16* x = d1; bf16* y = d0;
  bool joined = enc_join == nullptr;
  const bool ckv_stacked = w->ckv_w != nullptr && w->ckv_b != nullptr;
  const size_t Np_all = (size_t)B * Tp;
  for (int l = 0; l < 3; ++l) {                      // model/layers.py:46-50 (post-norm)
    const TtsDecLayerW& L = w->dec[l];
    if ((e = proj(x, Nm, kD, L.sa.w_in, L.sa.b_in, 3 * kD, ACT_NONE, qkv, s))) return e;
    if ((e = run_attention(qkv, 3 * kD, qkv + kD, qkv + 2 * kD, 3 * kD, att, nullptr, mel_len, nb, Tm, Tm, 1, s))) return e;
    {                                                // y = LN1(x + att Wo^T + b), cq = y Wq^T + bq (chained launch: opt-in)
      GemmArgs g = linear_args(att, Nm, kD, L.sa.w_out, L.sa.b_out, kD);
      g.residual = x; g.ld_res = kD;
      g.ln_g = L.ln1.g; g.ln_b = L.ln1.b;
      g.out_bf16 = y; g.ld_bf16 = kD;
      g.W2 = (const bf16*)L.ca.w_in; g.bias2 = L.ca.b_in; g.act2 = ACT_NONE;
      g.out2 = cq; g.ld_out2 = kD;
      if (chain_enabled() && gemm2_chain_supports(g)) {
        if ((e = gemm2_chain_launch(g, s))) return e;
      } else {
        if ((e = proj_res_ln(att, Nm, kD, L.sa.w_out, L.sa.b_out, x, L.ln1, y, s))) return e;
        if ((e = proj(y, Nm, kD, L.ca.w_in, L.ca.b_in, kD, ACT_NONE, cq, s))) return e;
      }
    }
    if (!joined) {                                   // first use of the encoder branch's results
      if (cudaStreamWaitEvent(s, enc_join, 0) != cudaSuccess) return "side stream join failed";
      joined = true;
    }
    const bf16* ckv = ckv_stacked ? ws.ckv + (size_t)b0 * Tp * 6 * kD + (size_t)l * 2 * kD
                                  : ws.ckv + (size_t)l * Np_all * 2 * kD + (size_t)b0 * Tp * 2 * kD;
    const int ld_ckv = ckv_stacked ? 6 * kD : 2 * kD;
    float* probs = (flags & TTS_FLAG_WANT_ALIGN) ? align + ((size_t)l * B + b0) * kH * Tm * Tp : nullptr;
    if ((e = run_attention(cq, kD, ckv, ckv + kD, ld_ckv, att, probs, ph_len, nb, Tm, Tp, 0, s))) return e;
    if ((e = proj_res_ln(att, Nm, kD, L.ca.w_out, L.ca.b_out, y, L.ln2, x, s))) return e;            // x = LN2
    if ((e = ffn_block(x, Nm, L.ffn, L.ln3, ffn, y, s))) return e;                                   // y = LN3
    bf16* t = x; x = y; y = t;
  }
  {                                                  // mel + stop heads (model/model.py:309,313)
    GemmArgs g = linear_args(x, Nm, kD, w->head_w, w->head_b, kMel + 1);
    g.out_f32 = mel_pre + r0 * kMel; g.ld_f32 = kMel;
    g.out_bf16 = pre16; g.ld_bf16 = kMel;
    g.out_last = stop + r0;
    if ((e = gemm_launch(g, s))) return e;
  }
  return run_postnet(w, pre16, mel_pre + r0 * kMel, nb, Tm, qkv, ffn, mel_post + r0 * kMel, s);
}

// ------------------------------------------------------------------ TransformerTTS.forward
extern "C" int tts_forward(const TtsWeights* w, const int64_t* phoneme, const void* mel, const int64_t* ph_lens,
                           const int64_t* mel_lens, int B, int Tp, int Tm, int flags, float* mel_pre, float* mel_post,
                           float* stop, float* align, void* workspace, size_t ws_bytes, void* stream) {
  if (check_common(w, B, Tp, Tm, workspace, ws_bytes)) return 1;
  if ((flags & TTS_FLAG_WANT_ALIGN) && align == nullptr) return fail("WANT_ALIGN set but align is NULL");
  if ((flags & TTS_FLAG_MEL_BF16) && (reinterpret_cast<uintptr_t>(mel) & 15) != 0) return fail("bf16 mel must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  Workspace ws = carve(workspace, B, Tp, Tm);
  const int Np = B * Tp;
  TRY(lens_launch(ph_lens, ws.ph_len, B, 1, Tp, s));
  TRY(lens_launch(mel_lens, ws.mel_len, B, 1, Tm, s));
  static const bool env_serial = getenv("TTS_SERIAL") != nullptr;
  SideStream* side = (env_serial || (flags & TTS_FLAG_SERIAL)) ? nullptr : side_stream();
  cudaStream_t se = s;                               // stream of the encoder branch
  std::unique_lock<std::mutex> side_hold;
  if (side != nullptr) {
    side_hold = std::unique_lock<std::mutex>(side->use);
    if (cudaEventRecord(side->fork, s) != cudaSuccess || cudaStreamWaitEvent(side->enc, side->fork, 0) != cudaSuccess)
      return fail("side stream fork failed");
    se = side->enc;
  }
  const char* err = run_encoder(w, phoneme, B, Tp, (flags & TTS_FLAG_ENC_MASK) != 0, ws, se);
  const bf16* memory = ws.e0;
  if (w->ckv_w != nullptr && w->ckv_b != nullptr) {  // one [Np][1536] projection for all layers
    if (err == nullptr) err = proj(memory, Np, kD, w->ckv_w, w->ckv_b, 6 * kD, ACT_NONE, ws.ckv, se);
  } else {
    for (int l = 0; l < 3 && err == nullptr; ++l)
      err = run_cross_kv(w, l, memory, Np, ws.ckv + (size_t)l * Np * 2 * kD, se);
  }
  if (side != nullptr && cudaEventRecord(side->enc_join, se) != cudaSuccess && err == nullptr) err = "side stream join failed";
  if (err == nullptr)
    err = run_decoder(w, mel, flags, B, Tp, Tm, 0, B, mel_pre, mel_post, stop, align, ws,
                      side != nullptr ? side->enc_join : nullptr, s);
  // whatever path was taken, the caller's stream ends up ordered after the encoder branch
  if (side != nullptr) cudaStreamWaitEvent(s, side->enc_join, 0);
  if (err != nullptr) return fail(err);
  return 0;
}

// ------------------------------------------------------------------ TransformerTTS.inference pieces
extern "C" int tts_encode(const TtsWeights* w, const int64_t* phoneme, const int64_t* ph_lens, int B, int Tp, int flags,
                          void* cross_kv, int32_t* ph_len_i32, void* workspace, size_t ws_bytes, void* stream) {
  if (check_common(w, B, Tp, 1, workspace, ws_bytes)) return 1;
  if (cross_kv == nullptr || ph_len_i32 == nullptr) return fail("tts_encode: output pointer is NULL");
  cudaStream_t s = (cudaStream_t)stream;
  Workspace ws = carve(workspace, B, Tp, 1);
  TRY(lens_launch(ph_lens, ws.ph_len, B, 1, Tp, s));
  TRY(lens_launch(ph_lens, ph_len_i32, B, 1, Tp, s));
  TRY(run_encoder(w, phoneme, B, Tp, (flags & TTS_FLAG_ENC_MASK) != 0, ws, s));
  const int Np = B * Tp;
  for (int l = 0; l < 3; ++l)
    TRY(run_cross_kv(w, l, ws.e0, Np, (bf16*)cross_kv + (size_t)l * Np * 2 * kD, s));
  return 0;
}

extern "C" int tts_decode(const TtsWeights* w, const void* cross_kv, const int32_t* ph_len_i32, int B, int Tp,
                          int max_frames, float stop_threshold, void* self_kv, float* mel_out, float* stop_out,
                          int32_t* stop_count, int32_t* n_frames, int rows_per_cluster, int cluster_size,
                          const float* forced_mel, void* stream) {
  if (w == nullptr) return fail("weights pointer is NULL");
  if (B <= 0 || Tp <= 0 || max_frames <= 0) return fail("B, Tp and max_frames must be positive");
  if (max_frames > kPeMax) return fail("max_frames exceeds the positional-encoding table (5000)");
  if (w->pre2_kt == nullptr) return fail("tts_decode needs the K-blocked transposes (TtsWeights.dec_kt / pre2_kt)");
  for (int l = 0; l < 3; ++l)
    for (int j = 0; j < 3; ++j)
      if (w->dec_kt[l][j] == nullptr) return fail("tts_decode needs the K-blocked transposes (TtsWeights.dec_kt / pre2_kt)");
  cudaStream_t s = (cudaStream_t)stream;
  DecodeArgs a;
  memset(&a, 0, sizeof(a));
  for (int l = 0; l < 3; ++l) {
    const TtsDecLayerW& L = w->dec[l];
    DecLayerW& d = a.layer[l];
    d.w_qkv = (const bf16*)L.sa.w_in; d.b_qkv = L.sa.b_in;
    d.w_so = (const bf16*)w->dec_kt[l][0]; d.b_so = L.sa.b_out;
    d.ln1_g = L.ln1.g; d.ln1_b = L.ln1.b;
    d.w_cq = (const bf16*)L.ca.w_in; d.b_cq = L.ca.b_in;
    d.w_co = (const bf16*)w->dec_kt[l][1]; d.b_co = L.ca.b_out;
    d.ln2_g = L.ln2.g; d.ln2_b = L.ln2.b;
    d.w_f1 = (const bf16*)L.ffn.w1; d.b_f1 = L.ffn.b1;
    d.w_f2 = (const bf16*)w->dec_kt[l][2]; d.b_f2 = L.ffn.b2;
    d.ln3_g = L.ln3.g; d.ln3_b = L.ln3.b;
  }
  a.w_pre1 = (const bf16*)w->pre1_w; a.b_pre1 = w->pre1_b;
  a.w_pre2 = (const bf16*)w->pre2_kt; a.b_pre2 = w->pre2_b;
  a.w_head = (const bf16*)w->head_w; a.b_head = w->head_b;
  a.pe = w->pe; a.alpha = w->alpha;
  a.cross_kv = (const bf16*)cross_kv;
  a.self_kv = (bf16*)self_kv;
  a.ph_len = ph_len_i32;
  a.B = B; a.Tp = Tp; a.Tmax = max_frames;
  // sigmoid(z) >= thr  <=>  z >= logit(thr)   (model/model.py:382-383)
  if (stop_threshold <= 0.f) a.stop_logit = -INFINITY;
  else if (stop_threshold >= 1.f) a.stop_logit = INFINITY;
  else a.stop_logit = (float)log((double)stop_threshold / (1.0 - (double)stop_threshold));
  a.mel_out = mel_out; a.stop_out = stop_out; a.stop_count = stop_count; a.n_frames = n_frames;
  a.forced_mel = forced_mel;
  if (cudaMemsetAsync(stop_count, 0, sizeof(int) * (size_t)max_frames, s) != cudaSuccess)
    return fail("cudaMemsetAsync(stop_count) failed");
  int C = cluster_size, R = rows_per_cluster;
  if (C <= 0 && R <= 0) {
    decode_plan(B, &R, &C);
  } else if (C <= 0) {                               // largest cluster whose landing buffers fit and that runs in one wave
    C = 16;
    while (C > 4 && (B + R - 1) / R > decode_max_clusters(R, C)) C >>= 1;
  } else if (R <= 0) {                               // fewest rows per cluster that still fits one wave
    R = 1;
    while (R < 8 && (B + R - 1) / R > decode_max_clusters(R, C)) R *= 2;
  }
  TRY(decode_launch(a, R, C, s));
  return 0;
}

extern "C" int tts_postnet(const TtsWeights* w, const float* mel_pre, int B, int T, float* mel_post, void* workspace,
                           size_t ws_bytes, void* stream) {
  if (check_common(w, B, 1, T, workspace, ws_bytes)) return 1;
  cudaStream_t s = (cudaStream_t)stream;
  Workspace ws = carve(workspace, B, 1, T);
  TRY(cast_mel_launch(mel_pre, ws.mel_pre16, (size_t)B * T * kMel, s));
  TRY(run_postnet(w, ws.mel_pre16, mel_pre, B, T, ws.qkv_d, ws.ffn_d, mel_post, s));
  return 0;
}

// ------------------------------------------------------------------ forward-only training / validation helpers
extern "C" int tts_teacher_forcing_mix(const float* pred, const float* mel, const uint8_t* seed, const int64_t* mel_lens,
                                       int B, int T, int L_bar, float* out, void* stream) {
  if (pred == nullptr || mel == nullptr || seed == nullptr || mel_lens == nullptr || out == nullptr)
    return fail("tts_teacher_forcing_mix: NULL pointer");
  if (B <= 0 || T <= 0 || L_bar <= 0) return fail("tts_teacher_forcing_mix: B, T and L_bar must be positive");
  if (((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(mel) | reinterpret_cast<uintptr_t>(out)) & 15) != 0)
    return fail("tts_teacher_forcing_mix: tensors must be 16-byte aligned");
  if (out == pred || out == mel) return fail("tts_teacher_forcing_mix: out must not alias an input");
  TRY(tf_mix_launch(pred, mel, seed, mel_lens, B, T, L_bar, out, (cudaStream_t)stream));
  return 0;
}

extern "C" size_t tts_loss_scratch_bytes(void) {
  DevInfo di;
  if (device_info(&di) != nullptr) return 0;
  return (size_t)loss_blocks(di.sm_count) * 24;
}

extern "C" int tts_loss(const float* pred, const float* post, const float* stop, const float* mel, const int64_t* mel_lens,
                        int B, int T, float pos_weight, float* out4, void* scratch, size_t scratch_bytes, void* stream) {
  if (pred == nullptr || post == nullptr || stop == nullptr || mel == nullptr || mel_lens == nullptr || out4 == nullptr)
    return fail("tts_loss: NULL pointer");
  if (B <= 0 || T <= 0) return fail("tts_loss: B and T must be positive");
  if (((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(post) | reinterpret_cast<uintptr_t>(mel)) & 15) != 0)
    return fail("tts_loss: tensors must be 16-byte aligned");
  if ((reinterpret_cast<uintptr_t>(scratch) & 7) != 0) return fail("tts_loss: scratch must be 8-byte aligned");
  TRY(loss_launch(pred, post, stop, mel, mel_lens, B, T, pos_weight, out4, scratch, scratch_bytes, (cudaStream_t)stream));
  return 0;
}

// Variant of tts_loss for callers that cannot allocate device memory: the partial sums go to a fixed array inside the
// library (train_aux.cu) instead of caller scratch.  No cudaMalloc / cudaFree here or below.
extern "C" int tts_loss_static(const float* pred, const float* post, const float* stop, const float* mel,
                               const int64_t* mel_lens, int B, int T, float pos_weight, float* out4, void* stream) {
  if (pred == nullptr || post == nullptr || stop == nullptr || mel == nullptr || mel_lens == nullptr || out4 == nullptr)
    return fail("tts_loss_static: NULL pointer");
  if (B <= 0 || T <= 0) return fail("tts_loss_static: B and T must be positive");
  if (((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(post) | reinterpret_cast<uintptr_t>(mel)) & 15) != 0)
    return fail("tts_loss_static: tensors must be 16-byte aligned");
  TRY(loss_static_launch(pred, post, stop, mel, mel_lens, B, T, pos_weight, out4, (cudaStream_t)stream));
  return 0;
}

extern "C" int tts_pack_outputs(const float* pred, const float* post, const float* stop, long long rows, void* packed,
                                void* stream) {
  if (pred == nullptr || post == nullptr || stop == nullptr || packed == nullptr) return fail("tts_pack_outputs: NULL pointer");
  if (rows <= 0) return fail("tts_pack_outputs: rows must be positive");
  if (((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(post) | reinterpret_cast<uintptr_t>(packed)) & 15) != 0)
    return fail("tts_pack_outputs: tensors must be 16-byte aligned");
  TRY(pack_out_launch(pred, post, stop, rows, packed, (cudaStream_t)stream));
  return 0;
}
