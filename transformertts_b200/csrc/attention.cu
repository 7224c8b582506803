// Fused masked-softmax attention (head dim 64), probabilities written only on request.
//
// Replaces nn.MultiheadAttention's core as the reference calls it: scores = (q/8) k^T + (-inf masks),
// softmax over keys, P v (torch/nn/functional.py:6620-6665), with the masks of
// TransformerTTS._get_mask (model/model.py:229-257) derived from per-utterance lengths instead of
// materialised tensors, and the per-head cross-attention probabilities of
// TransformerDecoderLayer._mha_block (model/layers.py:54-74) stored fp32 [B,H,Tq,Tk] when asked for.
//
// Round-1 implementation: flash-style online softmax, 128 query rows per CTA (8 warps x 16 rows),
// 64-key tiles double-buffered with cp.async, bf16 mma.sync m16n8k16 with fp32 accumulate.
// (tcgen05/TMEM version of this kernel is the next step; see DESIGN.md.)
#include <stdlib.h>

#include "kernels.h"
#include "ptx.cuh"

namespace tts {

constexpr int AQ = 128;        // query rows per CTA
constexpr int AK = 64;         // keys per tile
constexpr int APITCH = 72;     // bf16 elements per smem row (64 + 8 pad -> conflict-free ldmatrix)
constexpr int ATT_THREADS = 256;
constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(p)));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], const void* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(p)));
}
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

struct AttnSmem {
  bf16 q[AQ * APITCH];
  bf16 k[2][AK * APITCH];
  bf16 v[2][AK * APITCH];
};

// rows [r0, r0+nrows) x 64 dims of a packed projection -> padded smem tile; rows >= rmax are zero-filled
__device__ __forceinline__ void load_tile(bf16* dst, const bf16* src, int ld, int r0, int nrows, int rmax, int tid) {
  for (int c = tid; c < nrows * 8; c += ATT_THREADS) {
    const int r = c >> 3, ch = c & 7;
    const int gr = r0 + r;
    const bool ok = gr < rmax;
    cp_async16(dst + r * APITCH + ch * 8, src + (size_t)(ok ? gr : 0) * ld + ch * 8, ok ? 16 : 0);
  }
}

// S[16 x 64] = Q_warp[16 x 64] * K_tile[64 x 64]^T for this warp
__device__ __forceinline__ void qk_tile(float (&s)[8][4], const uint32_t (&qf)[4][4], const bf16* ks, int lane) {
#pragma unroll
  for (int nb = 0; nb < 8; ++nb) { s[nb][0] = s[nb][1] = s[nb][2] = s[nb][3] = 0.f; }
#pragma unroll
  for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
    for (int np = 0; np < 4; ++np) {                 // 16 keys per ldmatrix.x4
      uint32_t kf[4];
      const int key = np * 16 + (lane & 7) + ((lane >> 4) << 3);
      const int col = kk * 16 + (((lane >> 3) & 1) << 3);
      ldsm_x4(kf, ks + key * APITCH + col);
      mma16816(s[np * 2 + 0], qf[kk], kf[0], kf[1]);
      mma16816(s[np * 2 + 1], qf[kk], kf[2], kf[3]);
    }
  }
}

__global__ void __launch_bounds__(ATT_THREADS)
attn_kernel(const AttnArgs a) {
  extern __shared__ uint8_t smem_raw[];
  AttnSmem& sm = *reinterpret_cast<AttnSmem*>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int qt = gridDim.x - 1 - blockIdx.x;          // heavy (late, causal) tiles first
  const int h = blockIdx.y, b = blockIdx.z;
  const int q0 = qt * AQ;
  const int klen = a.klen != nullptr ? min(a.klen[b], a.Tk) : a.Tk;
  int kv_end = klen;
  if (a.causal) kv_end = min(kv_end, q0 + AQ);
  const int n_tiles = (kv_end + AK - 1) / AK;

  const bf16* Qg = a.Q + (size_t)b * a.Tq * a.ldq + h * kDh;
  const bf16* Kg = a.K + (size_t)b * a.Tk * a.ldk + h * kDh;
  const bf16* Vg = a.V + (size_t)b * a.Tk * a.ldv + h * kDh;

  load_tile(sm.q, Qg, a.ldq, q0, AQ, a.Tq, tid);
  load_tile(sm.k[0], Kg, a.ldk, 0, AK, a.Tk, tid);
  load_tile(sm.v[0], Vg, a.ldv, 0, AK, a.Tk, tid);
  cp_async_commit();

  const int g = lane >> 2, tq = lane & 3;
  const int row_a = q0 + warp * 16 + g;               // query index of accumulator rows c0/c1
  const int row_b = row_a + 8;                        // ... and c2/c3
  uint32_t qf[4][4];
  float o[8][4];
#pragma unroll
  for (int nb = 0; nb < 8; ++nb) { o[nb][0] = o[nb][1] = o[nb][2] = o[nb][3] = 0.f; }
  float m_a = -INFINITY, m_b = -INFINITY, l_a = 0.f, l_b = 0.f;

  for (int it = 0; it < n_tiles; ++it) {
    const int buf = it & 1;
    if (it + 1 < n_tiles) {
      load_tile(sm.k[buf ^ 1], Kg, a.ldk, (it + 1) * AK, AK, a.Tk, tid);
      load_tile(sm.v[buf ^ 1], Vg, a.ldv, (it + 1) * AK, AK, a.Tk, tid);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    if (it == 0) {
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const int r = warp * 16 + (lane & 7) + (((lane >> 3) & 1) << 3);
        const int c = kk * 16 + ((lane >> 4) << 3);
        ldsm_x4(qf[kk], sm.q + r * APITCH + c);
      }
    }
    float s[8][4];
    qk_tile(s, qf, sm.k[buf], lane);
    const int kv0 = it * AK;
    float mx_a = -INFINITY, mx_b = -INFINITY;
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int j = kv0 + nb * 8 + tq * 2 + e;
        const bool dead = j >= klen;
        if (dead || (a.causal && j > row_a)) s[nb][e] = -INFINITY;
        if (dead || (a.causal && j > row_b)) s[nb][2 + e] = -INFINITY;
        mx_a = fmaxf(mx_a, s[nb][e]);
        mx_b = fmaxf(mx_b, s[nb][2 + e]);
      }
    }
    mx_a = fmaxf(mx_a, __shfl_xor_sync(0xffffffffu, mx_a, 1));
    mx_a = fmaxf(mx_a, __shfl_xor_sync(0xffffffffu, mx_a, 2));
    mx_b = fmaxf(mx_b, __shfl_xor_sync(0xffffffffu, mx_b, 1));
    mx_b = fmaxf(mx_b, __shfl_xor_sync(0xffffffffu, mx_b, 2));
    const float mn_a = fmaxf(m_a, mx_a), mn_b = fmaxf(m_b, mx_b);
    const float ref_a = mn_a == -INFINITY ? 0.f : mn_a;      // fully masked so far: avoid inf-inf
    const float ref_b = mn_b == -INFINITY ? 0.f : mn_b;
    const float corr_a = exp2f((m_a - ref_a) * kLog2e);
    const float corr_b = exp2f((m_b - ref_b) * kLog2e);
    m_a = mn_a; m_b = mn_b;
    float sum_a = 0.f, sum_b = 0.f;
    uint32_t pf[4][4];
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) {
      const float p0 = exp2f((s[nb][0] - ref_a) * kLog2e), p1 = exp2f((s[nb][1] - ref_a) * kLog2e);
      const float p2 = exp2f((s[nb][2] - ref_b) * kLog2e), p3 = exp2f((s[nb][3] - ref_b) * kLog2e);
      sum_a += p0 + p1; sum_b += p2 + p3;
      pf[nb >> 1][(nb & 1) * 2 + 0] = pack_bf16(p0, p1);
      pf[nb >> 1][(nb & 1) * 2 + 1] = pack_bf16(p2, p3);
    }
    l_a = l_a * corr_a + sum_a;
    l_b = l_b * corr_b + sum_b;
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) { o[nb][0] *= corr_a; o[nb][1] *= corr_a; o[nb][2] *= corr_b; o[nb][3] *= corr_b; }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {                  // 16 keys per step
#pragma unroll
      for (int np = 0; np < 4; ++np) {                // 16 head dims per ldmatrix.x4.trans
        uint32_t vf[4];
        const int key = kk * 16 + (lane & 7) + (((lane >> 3) & 1) << 3);
        const int col = np * 16 + ((lane >> 4) << 3);
        ldsm_x4_t(vf, sm.v[buf] + key * APITCH + col);
        mma16816(o[np * 2 + 0], pf[kk], vf[0], vf[1]);
        mma16816(o[np * 2 + 1], pf[kk], vf[2], vf[3]);
      }
    }
    __syncthreads();
  }
  l_a += __shfl_xor_sync(0xffffffffu, l_a, 1); l_a += __shfl_xor_sync(0xffffffffu, l_a, 2);
  l_b += __shfl_xor_sync(0xffffffffu, l_b, 1); l_b += __shfl_xor_sync(0xffffffffu, l_b, 2);
  const float inv_a = l_a > 0.f ? 1.f / l_a : 0.f;
  const float inv_b = l_b > 0.f ? 1.f / l_b : 0.f;

  // ---- O: stage this warp's 16 x 64 block in its own rows of the Q tile, then 16-byte row stores
  {
    bf16* st = sm.q + warp * 16 * APITCH;
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) {
      *reinterpret_cast<uint32_t*>(st + g * APITCH + nb * 8 + tq * 2) = pack_bf16(o[nb][0] * inv_a, o[nb][1] * inv_a);
      *reinterpret_cast<uint32_t*>(st + (g + 8) * APITCH + nb * 8 + tq * 2) = pack_bf16(o[nb][2] * inv_b, o[nb][3] * inv_b);
    }
    __syncwarp();
    bf16* Og = a.O + (size_t)b * a.Tq * a.ldo + h * kDh;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = i * 32 + lane;
      const int r = c >> 3, ch = c & 7;
      const int qi = q0 + warp * 16 + r;
      if (qi < a.Tq)
        *reinterpret_cast<uint4*>(Og + (size_t)qi * a.ldo + ch * 8) = *reinterpret_cast<const uint4*>(st + r * APITCH + ch * 8);
    }
  }
  if (a.probs == nullptr) return;

  // ---- second pass: recompute the scores and store the normalised probabilities (fp32, all Tk keys)
  __syncthreads();
  float* stage = reinterpret_cast<float*>(&sm.k[0][0]) + warp * 16 * 65;     // 8 warps x 16 x 65 fp32 = 33 KB over k[]+v[]
  bf16* kbuf = sm.q;                                                         // Q lives in registers by now
  static_assert(sizeof(sm.k) + sizeof(sm.v) >= 8 * 16 * 65 * sizeof(float), "probability staging does not fit");
  float* Pg = a.probs + ((size_t)(b * kH + h) * a.Tq) * a.Tk;
  const int all_tiles = (a.Tk + AK - 1) / AK;
  const float ref_a = m_a == -INFINITY ? 0.f : m_a, ref_b = m_b == -INFINITY ? 0.f : m_b;
  for (int it = 0; it < all_tiles; ++it) {
    const int kv0 = it * AK;
    load_tile(kbuf, Kg, a.ldk, kv0, AK, a.Tk, tid);
    cp_async_commit();
    cp_async_wait<0>();
    __syncthreads();
    float s[8][4];
    qk_tile(s, qf, kbuf, lane);
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int jl = nb * 8 + tq * 2 + e;
        const int j = kv0 + jl;
        const bool dead = j >= klen;
        const float pa = (dead || (a.causal && j > row_a)) ? 0.f : exp2f((s[nb][e] - ref_a) * kLog2e) * inv_a;
        const float pb = (dead || (a.causal && j > row_b)) ? 0.f : exp2f((s[nb][2 + e] - ref_b) * kLog2e) * inv_b;
        stage[g * 65 + jl] = pa;
        stage[(g + 8) * 65 + jl] = pb;
      }
    }
    __syncwarp();
#pragma unroll 4
    for (int r = 0; r < 16; ++r) {
      const int qi = q0 + warp * 16 + r;
      if (qi < a.Tq) {
        float* dst = Pg + (size_t)qi * a.Tk + kv0;
        if (kv0 + lane < a.Tk) dst[lane] = stage[r * 65 + lane];
        if (kv0 + 32 + lane < a.Tk) dst[32 + lane] = stage[r * 65 + 32 + lane];
      }
    }
    __syncthreads();
  }
}

const char* attn_launch(const AttnArgs& a, cudaStream_t s) {
  if (a.B <= 0 || a.Tq <= 0 || a.Tk <= 0) return "attention: empty problem";
  if ((a.ldq | a.ldk | a.ldv | a.ldo) % 8 != 0) return "attention: leading dimensions must be multiples of 8";
  {
    static const bool force_v1 = getenv("TTS_ATTN_V1") != nullptr;
    static const bool no_v3 = getenv("TTS_ATTN_NO_V3") != nullptr;
    // probabilities wanted and short keys: single-pass softmax with the block store (without the store the
    // streaming kernel is as fast: 18.8 vs 22.6 us on the encoder shape)
    if (!force_v1 && !no_v3 && a.probs != nullptr && attn3_supports(a)) return attn3_launch(a, s);
    if (!force_v1 && attn2_supports(a)) return attn2_launch(a, s);
  }
  static bool attr_done = false;
  if (!attr_done) {
    if (cudaFuncSetAttribute(attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(AttnSmem)) !=
        cudaSuccess)
      return "cudaFuncSetAttribute(attention) failed";
    attr_done = true;
  }
  dim3 grid((a.Tq + AQ - 1) / AQ, kH, a.B);
  attn_kernel<<<grid, ATT_THREADS, sizeof(AttnSmem), s>>>(a);
  note_launch();
  return cudaGetLastError() == cudaSuccess ? nullptr : "attention kernel launch failed";
}

}  // namespace tts
