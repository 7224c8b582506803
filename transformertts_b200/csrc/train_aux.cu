// Forward-only pieces of the reference's training / validation step that consume the hot path's outputs in place
// (SURVEY 8(f) rank 1):
//   * block-wise scheduled-sampling mix   utils/util.py:103-120  (block_mask + apply_teacher_forcing)
//   * TransformerTTSLoss.forward           loss.py:15-55          (masked MSE x2 + BCE-with-logits, pos_weight)
//   * packing of pred / post / stop into one bf16 buffer for a single device -> host copy
// All three are HBM-bound streaming kernels: 16-byte coalesced loads, padded rows skipped (their bytes are never
// read), grids sized in multiples of the SM count.  The loss is reduced in two deterministic stages (per-block fp64
// partials in a caller-owned scratch, then one block in fixed order), so a loss value is bit-reproducible.
#include <mutex>

#include "kernels.h"

namespace tts {

constexpr int AUX_THREADS = 256;

// ------------------------------------------------------------------ scheduled-sampling mix
// mask[b][t] = any(seed[b][t-L/2 .. t-L/2+L-1] inside [0,T))      (max_pool1d(kernel L, stride 1, padding L/2)[:T])
// out[b][t][:] = (mask ? pred : mel)[b][t][:] * (t < mel_lens[b])
__global__ void __launch_bounds__(AUX_THREADS)
tf_mix_kernel(const float4* __restrict__ pred, const float4* __restrict__ mel, const uint8_t* __restrict__ seed,
              const int64_t* __restrict__ lens, int B, int T, int L_bar, float4* __restrict__ out) {
  constexpr int V = kMel / 4;                                    // 20 float4 per frame
  const long long n = (long long)B * T * V;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / V;
    const int b = (int)(row / T), t = (int)(row - (long long)b * T);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t < lens[b]) {
      const int lo = max(t - L_bar / 2, 0), hi = min(t - L_bar / 2 + L_bar - 1, T - 1);
      bool m = false;
      for (int j = lo; j <= hi; ++j) m |= seed[(long long)b * T + j] != 0;
      v = __ldg((m ? pred : mel) + i);
    }
    out[i] = v;
  }
}

const char* tf_mix_launch(const float* pred, const float* mel, const uint8_t* seed, const int64_t* lens, int B, int T,
                          int L_bar, float* out, cudaStream_t s) {
  DevInfo di;
  if (const char* e = device_info(&di)) return e;
  const long long n = (long long)B * T * (kMel / 4);
  long long blocks = (n + AUX_THREADS - 1) / AUX_THREADS;
  const long long cap = (long long)di.sm_count * 8;
  if (blocks > cap) blocks = cap;
  tf_mix_kernel<<<(unsigned)blocks, AUX_THREADS, 0, s>>>(reinterpret_cast<const float4*>(pred),
                                                         reinterpret_cast<const float4*>(mel), seed, lens, B, T, L_bar,
                                                         reinterpret_cast<float4*>(out));
  note_launch();
  return cudaGetLastError() == cudaSuccess ? nullptr : "scheduled-sampling mix kernel launch failed";
}

// ------------------------------------------------------------------ loss
// stage 1: every block accumulates  sum (pred-mel)^2,  sum (post-mel)^2,  sum bce(stop)  over its share of the valid
// frames and writes three fp64 partials; stage 2: one block adds the partials in index order and forms the four scalars.
struct LossPartial { double pred, post, stop; };

__device__ __forceinline__ float sq4(const float4 a, const float4 b) {
  const float x = a.x - b.x, y = a.y - b.y, z = a.z - b.z, w = a.w - b.w;
  return fmaf(x, x, fmaf(y, y, fmaf(z, z, w * w)));
}

__global__ void __launch_bounds__(AUX_THREADS)
loss_partial_kernel(const float4* __restrict__ pred, const float4* __restrict__ post, const float* __restrict__ stop,
                    const float4* __restrict__ mel, const int64_t* __restrict__ lens, int B, int T, float pos_weight,
                    LossPartial* __restrict__ partial) {
  constexpr int V = kMel / 4;
  const long long n = (long long)B * T * V;
  float a_pred = 0.f, a_post = 0.f, a_stop = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / V;
    const int c = (int)(i - row * V);
    const int b = (int)(row / T), t = (int)(row - (long long)b * T);
    const long long len = lens[b];
    if (t >= len) continue;                                        // padded frame: nothing is read
    const float4 m = __ldg(mel + i);
    a_pred += sq4(__ldg(pred + i), m);
    a_post += sq4(__ldg(post + i), m);
    if (c == 0) {
      // F.binary_cross_entropy_with_logits(x, y, pos_weight=w):  (1 - y) x + (1 + (w - 1) y) softplus(-x)
      const float x = __ldg(stop + row);
      const float y = (t == len - 1) ? 1.f : 0.f;
      const float sp = log1pf(expf(-fabsf(x))) + fmaxf(-x, 0.f);
      a_stop += (1.f - y) * x + (1.f + (pos_weight - 1.f) * y) * sp;
    }
  }
  __shared__ double red[3][AUX_THREADS / 32];
  double d[3] = {(double)a_pred, (double)a_post, (double)a_stop};
#pragma unroll
  for (int k = 0; k < 3; ++k)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) d[k] += __shfl_xor_sync(0xffffffffu, d[k], o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { red[0][warp] = d[0]; red[1][warp] = d[1]; red[2][warp] = d[2]; }
  __syncthreads();
  if (threadIdx.x == 0) {
    LossPartial p = {0.0, 0.0, 0.0};
    for (int w = 0; w < AUX_THREADS / 32; ++w) { p.pred += red[0][w]; p.post += red[1][w]; p.stop += red[2][w]; }
    partial[blockIdx.x] = p;
  }
}

__global__ void loss_final_kernel(const LossPartial* __restrict__ partial, int n_partial, const int64_t* __restrict__ lens,
                                  int B, int T, float* __restrict__ out4) {
  if (threadIdx.x != 0) return;                                   // a few hundred additions: one thread, fixed order
  double p = 0.0, q = 0.0, s = 0.0;
  for (int i = 0; i < n_partial; ++i) { p += partial[i].pred; q += partial[i].post; s += partial[i].stop; }
  long long valid = 0;
  for (int b = 0; b < B; ++b) { const long long l = lens[b]; valid += l < 0 ? 0 : (l > T ? T : l); }
  const double nv = (double)valid;
  const double pred_mel = p / (nv * kMel), post_mel = q / (nv * kMel), stop = s / nv;   // nn.MSELoss / 'mean' reductions
  out4[0] = (float)(pred_mel + 0.5 * post_mel + stop);            // total      loss.py:39,48
  out4[1] = (float)pred_mel;
  out4[2] = (float)post_mel;
  out4[3] = (float)stop;
}

int loss_blocks(int sm_count) { return sm_count * 4; }

const char* loss_launch(const float* pred, const float* post, const float* stop, const float* mel, const int64_t* lens,
                        int B, int T, float pos_weight, float* out4, void* scratch, size_t scratch_bytes, cudaStream_t s) {
  DevInfo di;
  if (const char* e = device_info(&di)) return e;
  const int blocks = loss_blocks(di.sm_count);
  if (scratch == nullptr || scratch_bytes < sizeof(LossPartial) * (size_t)blocks) return "loss: scratch too small";
  LossPartial* part = reinterpret_cast<LossPartial*>(scratch);
  loss_partial_kernel<<<blocks, AUX_THREADS, 0, s>>>(reinterpret_cast<const float4*>(pred),
                                                     reinterpret_cast<const float4*>(post), stop,
                                                     reinterpret_cast<const float4*>(mel), lens, B, T, pos_weight, part);
  loss_final_kernel<<<1, 32, 0, s>>>(part, blocks, lens, B, T, out4);
  note_launch(2);
  return cudaGetLastError() == cudaSuccess ? nullptr : "loss kernel launch failed";
}

// ------------------------------------------------------------------ loss with library-owned scratch (tts_loss_static)
// Some callers of this library are not allowed to allocate device memory at the point where they want a loss value:
// a plugin running inside a host framework whose memory pool is owned and budgeted by someone else, a real-time or
// serving thread where cudaMalloc's implicit device synchronisation is unacceptable, a process that has handed every
// free byte of HBM to a caching allocator and would fail a late cudaMalloc.  For them the stage-1 partial sums live in
// the array below, which is part of the library image: every GPU that loads libtts_b200.so carries these 24 KiB from
// module load until the process exits, and nothing is allocated or freed per call.
//
// kLossStaticPartials bounds the stage-1 grid (4 blocks per SM, as in loss_launch, so both entry points add the same
// numbers in the same order): 1024 partials cover devices of up to 256 SMs.  A device that would need more is refused.
// Nothing in the array has to start at zero: each stage-1 block overwrites its own slot and stage 2 reads exactly the
// slots that were just written.  The array is shared by every call on a device and keeps the previous call's partials,
// so two calls whose kernels can overlap on the GPU (different streams, different host threads) race on it.
constexpr int kLossStaticPartials = 1024;
static __device__ __align__(256) LossPartial g_loss_static[kLossStaticPartials];

// In host code `g_loss_static` is only a handle for the runtime, not a device address; the address a kernel can use is
// looked up once per device (with that device current) and cached here.
static std::mutex g_loss_static_mu;
static LossPartial* g_loss_static_ptr[kMaxDevices] = {};

const char* loss_static_launch(const float* pred, const float* post, const float* stop, const float* mel,
                               const int64_t* lens, int B, int T, float pos_weight, float* out4, cudaStream_t s) {
  DevInfo di;
  if (const char* e = device_info(&di)) return e;
  const int blocks = loss_blocks(di.sm_count);
  if (blocks > kLossStaticPartials)
    return "tts_loss_static: this device needs more than the 1024 partial sums (256 SMs) the built-in scratch holds; "
           "use tts_loss with caller-provided scratch";
  LossPartial* part = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_loss_static_mu);
    part = g_loss_static_ptr[di.dev];
    if (part == nullptr) {                                          // first call on this device: resolve the symbol
      void* p = nullptr;
      if (cudaGetSymbolAddress(&p, g_loss_static) != cudaSuccess || p == nullptr) {
        (void)cudaGetLastError();
        return "tts_loss_static: cudaGetSymbolAddress failed for the built-in scratch";
      }
      part = g_loss_static_ptr[di.dev] = static_cast<LossPartial*>(p);
    }
  }
  loss_partial_kernel<<<blocks, AUX_THREADS, 0, s>>>(reinterpret_cast<const float4*>(pred),
                                                     reinterpret_cast<const float4*>(post), stop,
                                                     reinterpret_cast<const float4*>(mel), lens, B, T, pos_weight, part);
  loss_final_kernel<<<1, 32, 0, s>>>(part, blocks, lens, B, T, out4);
  note_launch(2);
  return cudaGetLastError() == cudaSuccess ? nullptr : "loss kernel launch failed";
}

// ------------------------------------------------------------------ output packing (one D2H copy per step)
// packed bf16 [B][T][168]: cols 0-79 pred, 80-159 post, 160 stop logit, 161-167 zero (row = 336 B, 16-byte multiple)
__global__ void __launch_bounds__(AUX_THREADS)
pack_out_kernel(const float4* __restrict__ pred, const float4* __restrict__ post, const float* __restrict__ stop,
                long long rows, uint2* __restrict__ out) {
  constexpr int V = kMel / 4, W = 2 * V + 2;                      // 42 uint2 (4 bf16 each) per row
  const long long n = rows * W;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / W;
    const int c = (int)(i - row * W);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < V) v = __ldg(pred + row * V + c);
    else if (c < 2 * V) v = __ldg(post + row * V + (c - V));
    else if (c == 2 * V) v.x = __ldg(stop + row);
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    uint2 u;
    u.x = *reinterpret_cast<uint32_t*>(&a);
    u.y = *reinterpret_cast<uint32_t*>(&b);
    out[i] = u;
  }
}

const char* pack_out_launch(const float* pred, const float* post, const float* stop, long long rows, void* out,
                            cudaStream_t s) {
  DevInfo di;
  if (const char* e = device_info(&di)) return e;
  const long long n = rows * 42;
  long long blocks = (n + AUX_THREADS - 1) / AUX_THREADS;
  const long long cap = (long long)di.sm_count * 8;
  if (blocks > cap) blocks = cap;
  pack_out_kernel<<<(unsigned)blocks, AUX_THREADS, 0, s>>>(reinterpret_cast<const float4*>(pred),
                                                           reinterpret_cast<const float4*>(post), stop, rows,
                                                           reinterpret_cast<uint2*>(out));
  note_launch();
  return cudaGetLastError() == cudaSuccess ? nullptr : "output pack kernel launch failed";
}

}  // namespace tts
