/* tts_b200.h -- C ABI of the B200-native TransformerTTS hot path.
 *
 * The reference (Orca0917/TransformerTTS) has no FFI layer: its boundary for this path is the Python
 * class model.TransformerTTS (model/model.py:138-394).  This header is what a binding of that class
 * calls: plain pointers and sizes, no C++ or torch types.  All data pointers are DEVICE pointers owned
 * by the caller (PyTorch in the shipped host code); the library never allocates or frees device memory,
 * launches asynchronously on the given stream (a cudaStream_t passed as void*), and is re-entrant across
 * distinct weight/workspace sets (one per GPU).  Every entry point returns 0 on success or a non-zero
 * status with a message available from tts_last_error() (thread-local).
 *
 * Reference interface replaced by each entry point:
 *   tts_forward          TransformerTTS.forward                model/model.py:260-320
 *   tts_encode           encoder half of TransformerTTS.inference   model/model.py:342-348
 *   tts_decode           AR loop of TransformerTTS.inference    model/model.py:351-384
 *   tts_postnet          PostNet + residual                     model/model.py:129-135, 310, 388
 *   tts_teacher_forcing_mix   block_mask + apply_teacher_forcing  utils/util.py:103-120
 *   tts_loss             TransformerTTSLoss.forward              loss.py:15-55
 *   tts_loss_static      (no reference counterpart) tts_loss over scratch built into the library: allocates nothing
 *   tts_pack_outputs     (no reference counterpart) pred/post/stop -> one bf16 buffer for a single D2H copy
 *   tts_op_linear / tts_op_linear_chain / tts_op_conv1d / tts_op_attention / tts_op_ffn
 *                        the torch.nn operators underneath (model/module.py:4-53,
 *                        torch/nn/functional.py:5798,6620-6665), exported for per-operator parity tests
 */
#ifndef TTS_B200_H
#define TTS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TTS_ABI_VERSION 3

/* flags */
#define TTS_FLAG_WANT_ALIGN 1 /* tts_forward: store cross-attention probabilities into `align` */
#define TTS_FLAG_SERIAL 4     /* tts_forward: run the encoder branch on the caller's stream instead of the side stream */
#define TTS_FLAG_ENC_MASK 2   /* apply the phoneme key-padding mask in the encoder (forward: yes, inference: no) */
#define TTS_FLAG_MEL_BF16 8   /* tts_forward: `mel` is bf16 [B][Tm][80] (16-byte aligned) instead of f32: halves the
                                 host -> device bytes of a step and skips the cast kernel (SURVEY 8(b) mel_bf16_or_f32) */

/* activation codes for the operator-level entry points */
#define TTS_ACT_NONE 0
#define TTS_ACT_RELU 1
#define TTS_ACT_TANH 2

typedef struct TtsAttnW {
  const void* w_in;   /* bf16 [768][256]  rows 0-255 Wq (pre-scaled by 1/8), 256-511 Wk, 512-767 Wv */
  const float* b_in;  /* f32  [768]       (q part pre-scaled) */
  const void* w_out;  /* bf16 [256][256] */
  const float* b_out; /* f32  [256] */
} TtsAttnW;

typedef struct TtsFfnW {
  const void* w1;  const float* b1; /* bf16 [1024][256], f32 [1024] */
  const void* w2;  const float* b2; /* bf16 [256][1024], f32 [256]  */
} TtsFfnW;

typedef struct TtsLnW { const float* g; const float* b; } TtsLnW; /* f32 [256] each */

typedef struct TtsEncLayerW { TtsAttnW sa; TtsFfnW ffn; TtsLnW ln1, ln2; } TtsEncLayerW;
typedef struct TtsDecLayerW { TtsAttnW sa, ca; TtsFfnW ffn; TtsLnW ln1, ln2, ln3; } TtsDecLayerW;

/* All matrices bf16, K-major ([out][in]); conv weights are [out][tap][in_padded] with eval-mode
 * BatchNorm folded in (model/module.py:19,30); biases / LayerNorm / PE in f32. */
typedef struct TtsWeights {
  const void* emb; int32_t n_phon; int32_t reserved0;   /* bf16 [n_phon][256] */
  const void* enc_conv_w[3]; const float* enc_conv_b[3];/* bf16 [256][5*256] */
  const void* enc_lin_w; const float* enc_lin_b;        /* bf16 [256][256] */
  TtsEncLayerW enc[3];
  const void* pre1_w; const float* pre1_b;              /* bf16 [256][128] (80 mel bins zero padded) */
  const void* pre2_w; const float* pre2_b;              /* bf16 [256][256] */
  TtsDecLayerW dec[3];
  const void* head_w; const float* head_b;              /* bf16 [81][256]: rows 0-79 mel, row 80 stop */
  const void* post_w[5]; const float* post_b[5];        /* [256][5*128], 3x[256][5*256], [80][5*256] */
  const float* pe; float alpha; int32_t reserved1;      /* f32 [5000][256]; shared scalar alpha */
  /* Cross-attention K/V projections of the three decoder layers stacked: bf16 [1536][256], rows 512 l + [0,256) = Wk of
   * layer l, 512 l + [256,512) = Wv of layer l (copies of dec[l].ca.w_in rows 256-767), f32 [1536] likewise.  Lets
   * tts_forward project the memory once for all layers.  May be NULL: the per-layer matrices are used then. */
  const void* ckv_w; const float* ckv_b;
  /* K-blocked transposes for the AR decode kernel's K-split projections: bf16 [K/8][256][8], element [kb][n][i] =
   * W[n][8 kb + i], so that 256 threads (one per output n) read consecutive 16-byte words of a K slice.
   * dec_kt[l][0] = dec[l].sa.w_out (K = 256), [l][1] = dec[l].ca.w_out (K = 256), [l][2] = dec[l].ffn.w2 (K = 1024);
   * pre2_kt = pre2_w (K = 256).  Required by tts_decode. */
  const void* dec_kt[3][3];
  const void* pre2_kt;
} TtsWeights;

int tts_abi_version(void);
const char* tts_last_error(void);
/* Number of CUDA kernels this library has launched in this process so far (monotonic). */
uint64_t tts_launch_count(void);

/* Bytes of scratch the calls below need for these sizes (Tm = frames; for tts_decode pass max frames). */
size_t tts_workspace_bytes(int B, int Tp, int Tm);

/* Teacher-forced forward.  phoneme i64 [B][Tp], mel f32 [B][Tm][80] (bf16 with TTS_FLAG_MEL_BF16), lens i64 [B]
 * (1 <= len <= T).  The encoder branch uses one side stream and one fork/join event pair per device; concurrent callers
 * on one device (two host threads, two weight / workspace sets) are serialised by the library while they ENQUEUE
 * (a per-device mutex from fork to join); their GPU work still overlaps.  TTS_FLAG_SERIAL avoids the side stream.
 * Outputs f32: mel_pre [B][Tm][80], mel_post [B][Tm][80], stop [B][Tm], align [3][B][4][Tm][Tp]
 * (align may be NULL unless TTS_FLAG_WANT_ALIGN). */
int tts_forward(const TtsWeights* w, const int64_t* phoneme, const void* mel, const int64_t* ph_lens,
                const int64_t* mel_lens, int B, int Tp, int Tm, int flags, float* mel_pre, float* mel_post,
                float* stop, float* align, void* workspace, size_t ws_bytes, void* stream);

/* Encoder + per-layer cross-attention K/V of the memory.
 * cross_kv bf16 [3][B][Tp][512] (K cols 0-255, V cols 256-511); ph_len_i32 int32 [B] (clamped copy). */
int tts_encode(const TtsWeights* w, const int64_t* phoneme, const int64_t* ph_lens, int B, int Tp, int flags,
               void* cross_kv, int32_t* ph_len_i32, void* workspace, size_t ws_bytes, void* stream);

/* Greedy AR decode, all steps in one launch.  self_kv bf16 [3][B][max_frames][512] scratch;
 * mel_out f32 [B][max_frames][80], stop_out f32 [B][max_frames]; stop_count int32 [max_frames] scratch;
 * n_frames int32 [1] (device) receives the number of frames produced under the reference's rule
 * (first step at which sigmoid(stop) >= threshold for ALL rows, else max_frames).
 * rows_per_cluster / cluster_size: 0 = choose automatically.
 * forced_mel: NULL for the reference's greedy loop; else f32 [B][max_frames][80], and step t consumes frame t-1 of it
 * instead of the kernel's own prediction (teacher-forced decode: every step is computed through the same kernel and
 * K/V cache but errors do not compound through the feedback; used for full-horizon parity). */
int tts_decode(const TtsWeights* w, const void* cross_kv, const int32_t* ph_len_i32, int B, int Tp,
               int max_frames, float stop_threshold, void* self_kv, float* mel_out, float* stop_out,
               int32_t* stop_count, int32_t* n_frames, int rows_per_cluster, int cluster_size,
               const float* forced_mel, void* stream);

/* Number of thread-block clusters of the decode kernel the current device holds at once for a given plan
 * (cudaOccupancyMaxActiveClusters); 0 if the plan cannot run. */
int tts_decode_max_clusters(int rows_per_cluster, int cluster_size);

/* Utterances one tts_decode launch holds co-resident on the current device (largest one-wave plan).  A larger batch is
 * still decoded correctly, in waves, but then the stop rule cannot end the launch early: callers should split larger
 * batches into micro-batches of at most this many rows (transformertts_b200.sharding does). */
int tts_decode_capacity(void);

/* PostNet with residual: mel_post = PostNet(mel_pre) + mel_pre, f32 [B][T][80] both. */
int tts_postnet(const TtsWeights* w, const float* mel_pre, int B, int T, float* mel_post, void* workspace,
                size_t ws_bytes, void* stream);

/* ---- forward-only pieces of the training / validation step ---- */

/* Block-wise scheduled sampling (utils/util.py:103-120).  seed u8 [B][T]: 1 where the reference's
 * `torch.rand(B,1,T) < 1 - p_tf` is true (the caller draws it, so parity tests are deterministic).  The seed is dilated
 * like max_pool1d(kernel L_bar, stride 1, padding L_bar/2)[:T]; out = where(mask, pred, mel) * (t < mel_lens[b]).
 * pred, mel, out f32 [B][T][80]; out may alias neither input. */
int tts_teacher_forcing_mix(const float* pred, const float* mel, const uint8_t* seed, const int64_t* mel_lens, int B,
                            int T, int L_bar, float* out, void* stream);

/* Bytes of scratch tts_loss needs on the current device (per-block partial sums). */
size_t tts_loss_scratch_bytes(void);

/* TransformerTTSLoss.forward (loss.py:15-55) over the valid frames t < mel_lens[b]:
 * out4 f32 [4] (device) = { total, pred_mel, post_mel, stop } with pred_mel = MSE(pred, mel), post_mel = MSE(post, mel),
 * stop = BCE-with-logits(stop, [t == len-1], pos_weight), total = pred_mel + 0.5 post_mel + stop.
 * pred/post/mel f32 [B][T][80], stop f32 [B][T].  Deterministic (fixed reduction order). */
int tts_loss(const float* pred, const float* post, const float* stop, const float* mel, const int64_t* mel_lens, int B,
             int T, float pos_weight, float* out4, void* scratch, size_t scratch_bytes, void* stream);

/* tts_loss without caller scratch and without any device allocation.  (No counterpart in the reference: it exists for
 * hosts that embed this library where device memory may not be allocated -- a plugin inside a framework that owns and
 * budgets the memory pool, a latency-critical serving thread that cannot take cudaMalloc's device-wide synchronisation,
 * a process whose HBM is already fully reserved by a caching allocator.)  Same arguments, same kernels and bit-identical
 * out4 as tts_loss; the per-block partial sums go to a fixed 24 KiB array inside the library (1024 partials of 24 bytes,
 * enough for 4 blocks per SM on devices of up to 256 SMs; each GPU that loads the library holds it until the process
 * exits).  A device that needs more partials is refused with an error -- use tts_loss there.  Calls cudaMalloc,
 * cudaFree or an allocating library routine never.
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress, with that device current)
 * and caches it; make that first call outside any stream capture.  All calls on a device share the array and it keeps
 * the previous call's partial sums (every launch overwrites the slots it then reads, so nothing needs clearing): calls
 * whose kernels may overlap on the GPU -- two streams, two host threads -- race on it and must be serialised by the
 * caller. */
int tts_loss_static(const float* pred, const float* post, const float* stop, const float* mel, const int64_t* mel_lens,
                    int B, int T, float pos_weight, float* out4, void* stream);

/* pred / post f32 [rows][80] + stop f32 [rows] -> packed bf16 [rows][168] (cols 0-79 pred, 80-159 post, 160 stop,
 * 161-167 zero): one buffer, one device -> host copy per step at half the bytes. */
int tts_pack_outputs(const float* pred, const float* post, const float* stop, long long rows, void* packed_bf16,
                     void* stream);

/* ---- operator-level entry points (used by the parity tests; same kernels as above) ---- */

/* out = epilogue(A[M][K] . W[N][K]^T): +bias, act, +alpha*pe[row % pe_period], +residual, LayerNorm.
 * A, W, residual, out_bf16 are bf16; bias/pe/ln f32; any of pe/residual/ln_g/out_bf16/out_f32 may be NULL.
 * K must be a multiple of 64 in memory (zero padded). */
int tts_op_linear(const void* A, const void* W, const float* bias, int M, int N, int K, int act,
                  const float* pe, float alpha, int pe_period, const void* residual, const float* ln_g,
                  const float* ln_b, void* out_bf16, float* out_f32, void* stream);

/* Two chained GEMMs in one launch (N = 256 both, K1 <= 256, M >= 129):
 *   y1 = LayerNorm(residual + A W1^T + b1) when ln_g != NULL (act1 must be 0), else act1(A W1^T + b1);
 *   out2 = act2(y1_bf16 W2^T + b2) + alpha * pe[row % pe_period]   (pe may be NULL).
 * y1 stays on the SM (bf16, tensor memory) as the A operand of the second GEMM; it is also stored to out1_bf16 unless
 * that is NULL.  tts_forward runs out-projection + LayerNorm -> cross-attention Q projection and prenet 1 -> prenet 2 + PE
 * this way only when the process sets TTS_GEMM_CHAIN=1; by default it launches the two GEMMs separately (measured
 * faster inside the whole forward, profiles/r02_gemm_chain_ab.txt). */
int tts_op_linear_chain(const void* A, const void* W1, const float* b1, int M, int K1, int act1, const void* residual,
                        const float* ln_g, const float* ln_b, const void* W2, const float* b2, int act2,
                        const float* pe, float alpha, int pe_period, void* out1_bf16, void* out2_bf16, void* stream);

/* Conv1d(kernel `taps`, zero padding (taps-1)/2) over x bf16 [B][T][C] with W bf16 [N][taps][Cpad]
 * (Cpad = C rounded up to 64), + bias, act; optional f32 residual [B][T][N]. */
int tts_op_conv1d(const void* x, const void* W, const float* bias, int B, int T, int C, int N, int taps,
                  int act, const float* residual_f32, void* out_bf16, float* out_f32, void* stream);

/* Fused FFN block of a post-norm layer (torch/nn/modules/transformer.py:951-957, 980-982):
 * out = LayerNorm(x + W2 relu(W1 x + b1) + b2); x, out bf16 [M][256] (distinct), w1 bf16 [1024][256],
 * w2 bf16 [256][1024], b1 f32 [1024], b2 / ln_g / ln_b f32 [256].  The hidden activation stays on the SM. */
int tts_op_ffn(const void* x, const void* w1, const float* b1, const void* w2, const float* b2, const float* ln_g,
               const float* ln_b, int M, void* out, void* stream);

/* Multi-head attention core, 4 heads x 64: q [B][Tq][ldq], k/v [B][Tk][ld] bf16 (q pre-scaled),
 * klen int32 [B] or NULL, causal 0/1, out bf16 [B][Tq][256], probs f32 [B][4][Tq][Tk] or NULL. */
int tts_op_attention(const void* q, int ldq, const void* k, int ldk, const void* v, int ldv, int B, int Tq,
                     int Tk, const int32_t* klen, int causal, void* out, float* probs, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TTS_B200_H */
