// Internal launch API between c_api.cu (orchestration) and the kernel translation units.
// Not part of the C ABI (include/tts_b200.h is).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace tts {

typedef __nv_bfloat16 bf16;

// model constants (config.yaml:25-42 of the reference; the kernels are specialised to them)
constexpr int kD = 256;      // d_model
constexpr int kH = 4;        // heads
constexpr int kDh = 64;      // head dim
constexpr int kF = 1024;     // FFN dim
constexpr int kMel = 80;     // mel bins
constexpr int kMelPad = 128; // mel bins padded to two 64-wide k-blocks
constexpr int kTaps = 5;     // conv kernel size
constexpr int kPeMax = 5000; // model/model.py:76

void note_launch(int n = 1);      // counts kernel launches (tts_launch_count in the C ABI)

// Per-device host state (host.cu).  Function attributes (the > 48 KB shared-memory opt-in) and the SM count belong to
// a DEVICE, and one process may drive several GPUs, so every launcher keys them by the current device ordinal.
constexpr int kMaxDevices = 64;
struct DevInfo { int dev; int sm_count; };
const char* device_info(DevInfo* out);            // current device + its SM count (cached); nullptr or an error string
class PerDeviceOnce {                              // "has this kernel's attribute been set on device d yet"
 public:
  bool need(int dev);
  void mark(int dev);
 private:
  bool done_[kMaxDevices] = {};
};
// Opt a kernel into `bytes` of dynamic shared memory once per device.
template <class K>
inline const char* ensure_smem(K kernel, int bytes, PerDeviceOnce& once, int dev, const char* what) {
  if (once.need(dev)) {
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess) return what;
    once.mark(dev);
  }
  return nullptr;
}

enum Act { ACT_NONE = 0, ACT_RELU = 1, ACT_TANH = 2 };

// One dense contraction  out[row, n] = epi( sum_k A[row(+tap), k] * W[n, k] ).
// A is a bf16 activation viewed as [B][T][C] (a plain matrix is B=1, T=M); the K loop runs over
// `taps` row-shifted copies of A (conv1d as implicit GEMM; rows outside [0,T) of the same utterance
// read as zero through TMA out-of-bounds fill) times `kc_per_tap` 64-wide channel blocks.
// W is bf16 [N][taps*kc_per_tap*64], K-major.
struct GemmArgs {
  const bf16* A;  int B, T, C;            // activation and its logical shape (C = true channels)
  const bf16* W;  int N;                  // weight rows actually present in memory
  int taps, tap_off, kc_per_tap;
  // epilogue, applied in this order: +bias, act, +alpha*pe[t], +residual, LayerNorm, store
  const float* bias;                      // [N]
  int act;
  const float* pe; float alpha; int pe_period;  // pe table [kPeMax][256] fp32 or null; position = row % pe_period
  const bf16* residual; int ld_res;       // bf16 [B*T][ld_res] or null
  const float* residual_f32; int ld_res32;// fp32 residual (PostNet tail) or null
  const float* ln_g; const float* ln_b;   // LayerNorm over N == 256 or null
  bf16* out_bf16; int ld_bf16;            // nullable
  float* out_f32; int ld_f32;             // nullable
  float* out_last;                        // if set, column N-1 goes to out_last[row] instead (stop head)
  // chained second GEMM (gemm2.cu CHAIN plan; N == 256): out2 = epi2( epi1(...) W2^T ) with epi2 = +bias2, act2,
  // +alpha*pe (pe / alpha / pe_period then belong to the SECOND epilogue); out_bf16 may be null (first result not stored)
  const bf16* W2; const float* bias2; int act2;
  bf16* out2; int ld_out2;
};

const char* gemm_launch(const GemmArgs& a, cudaStream_t s);   // returns nullptr or an error string
bool gemm2_supports(const GemmArgs& a);                        // persistent bf16-out kernel (gemm2.cu)
const char* gemm2_launch(const GemmArgs& a, cudaStream_t s);
bool gemm2_chain_supports(const GemmArgs& a);                  // two chained GEMMs in one launch (CTA pairs, K <= 256)
const char* gemm2_chain_launch(const GemmArgs& a, cudaStream_t s);

// Fused FFN block: out = LayerNorm(x + W2 relu(W1 x + b1) + b2) over M rows of 256 channels (ffn.cu).
// x, out bf16 [M][256] (distinct buffers), w1 bf16 [1024][256], w2 bf16 [256][1024], biases / LN fp32.
struct FfnArgs {
  const bf16* x; int M;
  const bf16* w1; const float* b1;
  const bf16* w2; const float* b2;
  const float* ln_g; const float* ln_b;
  bf16* out;
};
bool ffn_supports(const FfnArgs& a);
const char* ffn_launch(const FfnArgs& a, cudaStream_t s);

// Fused masked-softmax attention over packed projections.
// Q rows: q[(b*Tq + i)*ldq + h*64 + d], K/V likewise with Tk, ldk/ldv.  Keys j >= klen[b] are masked
// when klen != nullptr; causal masks j > i.  Q is expected pre-scaled by 1/sqrt(64).
// out[(b*Tq+i)*ldo + h*64 + d] bf16.  If probs != nullptr the normalised probabilities are stored as
// fp32 [B][H][Tq][Tk] (model/layers.py:68-74).
struct AttnArgs {
  const bf16* Q; int ldq;
  const bf16* K; int ldk;
  const bf16* V; int ldv;
  bf16* O; int ldo;
  float* probs;
  const int* klen;
  int B, Tq, Tk, causal;
};
const char* attn_launch(const AttnArgs& a, cudaStream_t s);
bool attn2_supports(const AttnArgs& a);                        // tcgen05 kernel (attention2.cu)
const char* attn2_launch(const AttnArgs& a, cudaStream_t s);
bool attn3_supports(const AttnArgs& a);                        // single-pass kernel, Tk <= 256 (attention3.cu)
const char* attn3_launch(const AttnArgs& a, cudaStream_t s);

// small elementwise kernels
const char* embed_launch(const int64_t* phoneme, const bf16* table, bf16* out, int n_tok, int n_phon,
                         cudaStream_t s);
const char* cast_mel_launch(const float* mel, bf16* out, size_t n, cudaStream_t s);
const char* lens_launch(const int64_t* in, int* out, int n, int lo, int hi, cudaStream_t s);

// forward-only training / validation helpers (train_aux.cu)
const char* tf_mix_launch(const float* pred, const float* mel, const uint8_t* seed, const int64_t* lens, int B, int T,
                          int L_bar, float* out, cudaStream_t s);
int loss_blocks(int sm_count);                                 // partial sums the loss keeps in its scratch (24 B each)
const char* loss_launch(const float* pred, const float* post, const float* stop, const float* mel, const int64_t* lens,
                        int B, int T, float pos_weight, float* out4, void* scratch, size_t scratch_bytes, cudaStream_t s);
// same two kernels over a scratch array that is part of the library image (no caller scratch, nothing allocated)
const char* loss_static_launch(const float* pred, const float* post, const float* stop, const float* mel,
                               const int64_t* lens, int B, int T, float pos_weight, float* out4, cudaStream_t s);
const char* pack_out_launch(const float* pred, const float* post, const float* stop, long long rows, void* out,
                            cudaStream_t s);

// Launch with programmatic stream serialization: the kernel's prologue (barrier init, TMEM allocation, tensor-map
// prefetch) may overlap the tail of the previous kernel in the stream; the kernel itself calls griddepcontrol.wait
// before it touches global memory.  TTS_NO_PDL=1 falls back to plain stream order.
bool pdl_enabled();
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
// same, as clusters of two CTAs (grid.x even): the CTA pairs of the cta_group::2 plans
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl_pairs(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// tensor-map helper (driver entry point fetched through the runtime; no libcuda link dependency)
const char* make_tmap_3d(CUtensorMap* m, const void* base, uint64_t d0, uint64_t d1, uint64_t d2,
                         uint64_t stride1_bytes, uint64_t stride2_bytes, uint32_t box0, uint32_t box1,
                         int swizzle_bytes = 128, int elem_bytes = 2);   // dims / boxes in elements of elem_bytes

}  // namespace tts
