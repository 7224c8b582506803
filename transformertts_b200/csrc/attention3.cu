// Single-pass masked-softmax attention for short key sequences (Tk <= 256, non-causal) on tcgen05 / TMEM / TMA.
//
// This is the cross-attention of TransformerDecoderLayer._mha_block (model/layers.py:54-74 ->
// torch/nn/functional.py:6620-6665) -- the one whose per-head probabilities the reference returns as
// `alignments` -- and the encoder self-attention (torch/nn/modules/transformer.py:961-975).  Phoneme sequences
// are short (Tp ~ 150), so a whole score row fits in TMEM: S = Q K^T is produced once (one 128 x 64*NT MMA
// group), the softmax is exact (no online rescale), P V follows, and the normalised probabilities leave the
// SM as ONE contiguous block: rows q0..q0+127 of probs[b][h] are adjacent in memory, so the [128][Tk] fp32
// tile is staged in shared memory in its final layout and written with cp.async.bulk (fully coalesced, async).
//
// One CTA = 128 queries of one (utterance, head), 320 threads:
//   warp 0      TMA producer : Q tile, K rows [0, 64*NT), V rows [0, 64*NT)  (SWIZZLE_128B, OOB rows zero)
//   warp 1      MMA issuer   : S (128 x 64NT x 64) -> TMEM, then O = P V (V consumed MN-major)
//   warps 2..9  softmax      : thread == (query row, parity of the 32-column chunk); row max and row sum are
//                              exchanged once per CTA through shared memory; unnormalised exp2 goes to smem as
//                              bf16 (A operand of P V) and, when probabilities are wanted, back into TMEM as fp32
// Shared memory (NT = 3: 96 KB, two CTAs per SM): [Q / P tile 0][P tiles 1..][K][V]; after P V everything is dead
// and the same bytes stage the probability block.
#include <stdlib.h>

#include "kernels.h"
#include "ptx.cuh"

namespace tts {

namespace a3 {
constexpr int BQ = 128;
constexpr int THREADS = 320;
constexpr int TILE_P = BQ * 128;       // 16 KB: 128 rows x 64 bf16 (Q, or one 64-key tile of P)
constexpr int TILE_KV = 64 * 128;      // 8 KB : 64 keys x 64 dims
constexpr float kLog2e = 1.4426950408889634f;

struct Dev {
  bf16* O; int ldo;
  float* probs;
  const int* klen;
  int Tq, Tk, NT;
};

__host__ __device__ inline int smem_main(int NT) { return NT * TILE_P + 2 * NT * TILE_KV; }
__host__ __device__ inline int smem_total(int NT) { return smem_main(NT) + 64 + 4 * BQ * 4 + 1024; }
}  // namespace a3

using namespace a3;

// The probabilities are written once and not read again by the forward: store them with an L2 evict-first policy so
// 123 MB of alignments per layer do not push the 26 MB activations the next kernels read out of L2.
__device__ __forceinline__ void bulk_store_1d(void* gdst, const void* smem_src, uint32_t bytes) {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;"
               :
               : "l"(reinterpret_cast<uint64_t>(gdst)), "r"(smem_u32(smem_src)), "r"(bytes), "l"(pol)
               : "memory");
}

__global__ void __launch_bounds__(THREADS, 2)
attn3_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
             const __grid_constant__ CUtensorMap tmV, const Dev p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  const int NT = p.NT;
  uint8_t* s_p = smem;                                   // P tiles (tile 0 aliases Q)
  uint8_t* s_k = smem + NT * TILE_P;
  uint8_t* s_v = s_k + NT * TILE_KV;
  uint64_t* qk_full = reinterpret_cast<uint64_t*>(smem + smem_main(NT));
  uint64_t* v_full = qk_full + 1;
  uint64_t* s_full = qk_full + 2;
  uint64_t* p_full = qk_full + 3;
  uint64_t* pv_done = qk_full + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(qk_full + 5);
  float* s_max = reinterpret_cast<float*>(qk_full + 8);  // [2][128]
  float* s_sum = s_max + 2 * BQ;                         // [2][128]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * BQ, h = blockIdx.y, b = blockIdx.z;
  const int klen = p.klen != nullptr ? max(1, min(__ldg(p.klen + b), p.Tk)) : p.Tk;
  const int nt_used = (klen + 63) >> 6;                  // 64-key tiles that hold visible keys
  const uint32_t tmem_cols = NT <= 3 ? 256u : 512u;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
    mbar_init(qk_full, 1); mbar_init(v_full, 1); mbar_init(s_full, 1); mbar_init(p_full, 8); mbar_init(pv_done, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch();
  griddep_wait();
  const uint32_t t_s = tmem_base + 64, t_o = tmem_base;  // O at columns 0..63, S at 64 .. 64 + 64 NT

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(qk_full, TILE_P + NT * TILE_KV);
      tma_load_3d(s_p, &tmQ, qk_full, h * kDh, q0, b);
      tma_load_3d(s_k, &tmK, qk_full, h * kDh, 0, b);
      mbar_expect_tx(v_full, NT * TILE_KV);
      tma_load_3d(s_v, &tmV, v_full, h * kDh, 0, b);
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
      const uint64_t dq = umma_desc_sw128(smem_u32(s_p)), dk = umma_desc_sw128(smem_u32(s_k));
      const uint32_t idesc_s = umma_idesc_bf16(BQ, 64 * NT);
      const uint32_t idesc_pv = umma_idesc_bf16(BQ, 64) | (1u << 16);          // B (= V) is MN-major
      mbar_wait(qk_full, 0);
      tc_fence_after();
#pragma unroll
      for (int k = 0; k < 4; ++k) tc_mma_bf16(t_s, dq + 2 * k, dk + 2 * k, idesc_s, k > 0 ? 1u : 0u);
      tc_commit(s_full);
      mbar_wait(v_full, 0);
      mbar_wait(p_full, 0);
      tc_fence_after();
      for (int t = 0; t < nt_used; ++t) {
        const uint64_t dp = umma_desc_sw128(smem_u32(s_p + t * TILE_P));
        const uint64_t dv = umma_desc_sw128(smem_u32(s_v + t * TILE_KV));
#pragma unroll
        for (int k = 0; k < 4; ++k)                      // 16 keys per step: P advances 32 B, V advances 16 rows
          tc_mma_bf16(t_o, dp + 2 * k, dv + 128 * k, idesc_pv, (t > 0 || k > 0) ? 1u : 0u);
      }
      tc_commit(pv_done);
    }
    __syncwarp();
  } else {
    const int hf = (warp - 2) >> 2;                      // chunk parity this thread owns
    const int q = warp & 3;                              // TMEM lane quadrant
    const int row = q * 32 + lane;
    const int qi = q0 + row;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    const uint32_t sw = static_cast<uint32_t>(row & 7);
    const int nch = 2 * NT;

    mbar_wait(s_full, 0);
    tc_fence_after();
    // ---- pass 1: row maximum over this thread's chunks
    float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
    for (int c = hf; c < nch && c * 32 < klen; c += 2) {
      uint32_t r[32];
      tmem_ld32(t_s + c * 32 + lane_off, r);
      tmem_ld_wait();
      const int nvis = klen - c * 32;
      if (nvis < 32) {
#pragma unroll
        for (int e = 0; e < 32; ++e)
          if (e >= nvis) r[e] = 0xff800000u;
      }
#pragma unroll
      for (int e = 0; e < 32; e += 4) {
        mx4[0] = fmaxf(mx4[0], __uint_as_float(r[e])); mx4[1] = fmaxf(mx4[1], __uint_as_float(r[e + 1]));
        mx4[2] = fmaxf(mx4[2], __uint_as_float(r[e + 2])); mx4[3] = fmaxf(mx4[3], __uint_as_float(r[e + 3]));
      }
    }
    float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
    s_max[hf * BQ + row] = mx;
    named_bar_sync(1, 256);
    mx = fmaxf(mx, s_max[(hf ^ 1) * BQ + row]);          // key 0 is always visible: finite
    const float nm = -mx * kLog2e;

    // ---- pass 2: p = exp2(s log2e - m); bf16 P tile for the MMA, fp32 p parked back in TMEM for the probabilities
    float sm4[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c = hf; c < nch; c += 2) {
      uint8_t* p_row = s_p + (c >> 1) * TILE_P + row * 128;
      const uint32_t cb = static_cast<uint32_t>(c & 1) * 4;
      if (c * 32 < klen) {
        uint32_t r[32];
        tmem_ld32(t_s + c * 32 + lane_off, r);
        tmem_ld_wait();
        const int nvis = klen - c * 32;
        if (nvis < 32) {
#pragma unroll
          for (int e = 0; e < 32; ++e)
            if (e >= nvis) r[e] = 0xff800000u;
        }
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const float pe = fast_exp2(fmaf(__uint_as_float(r[e]), kLog2e, nm));
          sm4[e & 3] += pe;
          r[e] = __float_as_uint(pe);
        }
        if (p.probs != nullptr) tmem_st32(t_s + c * 32 + lane_off, r);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 u;
          u.x = pack_bf16(__uint_as_float(r[j * 8 + 0]), __uint_as_float(r[j * 8 + 1]));
          u.y = pack_bf16(__uint_as_float(r[j * 8 + 2]), __uint_as_float(r[j * 8 + 3]));
          u.z = pack_bf16(__uint_as_float(r[j * 8 + 4]), __uint_as_float(r[j * 8 + 5]));
          u.w = pack_bf16(__uint_as_float(r[j * 8 + 6]), __uint_as_float(r[j * 8 + 7]));
          *reinterpret_cast<uint4*>(p_row + (((cb + j) ^ sw) << 4)) = u;
        }
      } else if ((c >> 1) < nt_used) {                   // masked half of a tile the MMA still reads
#pragma unroll
        for (int j = 0; j < 4; ++j)
          *reinterpret_cast<uint4*>(p_row + (((cb + j) ^ sw) << 4)) = make_uint4(0u, 0u, 0u, 0u);
      }
    }
    if (p.probs != nullptr) tmem_st_wait();
    fence_proxy_async();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(p_full);
    float l = (sm4[0] + sm4[1]) + (sm4[2] + sm4[3]);
    s_sum[hf * BQ + row] = l;
    named_bar_sync(1, 256);
    l += s_sum[(hf ^ 1) * BQ + row];
    const float inv = 1.f / l;

    // ---- O / l -> bf16 -> global: this thread stores 32 of the row's 64 head dims
    mbar_wait(pv_done, 0);
    tc_fence_after();
    {
      uint32_t o[32];
      tmem_ld32(t_o + hf * 32 + lane_off, o);
      tmem_ld_wait();
      if (qi < p.Tq) {
        bf16* og = p.O + ((size_t)b * p.Tq + qi) * p.ldo + h * kDh + hf * 32;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint4 u;
          u.x = pack_bf16(__uint_as_float(o[c * 8 + 0]) * inv, __uint_as_float(o[c * 8 + 1]) * inv);
          u.y = pack_bf16(__uint_as_float(o[c * 8 + 2]) * inv, __uint_as_float(o[c * 8 + 3]) * inv);
          u.z = pack_bf16(__uint_as_float(o[c * 8 + 4]) * inv, __uint_as_float(o[c * 8 + 5]) * inv);
          u.w = pack_bf16(__uint_as_float(o[c * 8 + 6]) * inv, __uint_as_float(o[c * 8 + 7]) * inv);
          *reinterpret_cast<uint4*>(og + c * 8) = u;
        }
      }
    }

    // ---- probabilities: stage [128][Tk] fp32 in its final layout over the dead operand buffers, bulk-store 32 rows
    if (p.probs != nullptr) {
      const int Tk = p.Tk;
      float* st_row = reinterpret_cast<float*>(smem) + (size_t)row * Tk;
      const int width = (Tk & 3) == 0 ? 4 : ((Tk & 1) == 0 ? 2 : 1);
      for (int c = hf; c * 32 < Tk; c += 2) {
        uint32_t r[32];
        if (c * 32 < klen) {
          tmem_ld32(t_s + c * 32 + lane_off, r);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 32; ++e) r[e] = __float_as_uint(__uint_as_float(r[e]) * inv);
        } else {
#pragma unroll
          for (int e = 0; e < 32; ++e) r[e] = 0u;
        }
        float* d = st_row + c * 32;
        const int nv = Tk - c * 32;                      // columns of this chunk that exist
        if (width == 4) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (j * 4 < nv) *reinterpret_cast<uint4*>(d + j * 4) = make_uint4(r[j * 4], r[j * 4 + 1], r[j * 4 + 2], r[j * 4 + 3]);
        } else if (width == 2) {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (j * 2 < nv) *reinterpret_cast<uint2*>(d + j * 2) = make_uint2(r[j * 2], r[j * 2 + 1]);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < nv) d[j] = __uint_as_float(r[j]);
        }
      }
      fence_proxy_async();
      named_bar_sync(2 + q, 64);                         // both chunk parities of this 32-row quadrant are staged
      if (hf == 0 && lane == 0) {
        const int rows = min(32, p.Tq - (q0 + q * 32));
        if (rows > 0) {
          float* g = p.probs + ((size_t)(b * kH + h) * p.Tq + q0 + q * 32) * Tk;
          bulk_store_1d(g, reinterpret_cast<float*>(smem) + (size_t)q * 32 * Tk, (uint32_t)(rows * Tk * 4));
          bulk_commit();
          bulk_wait_read<0>();
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

// ------------------------------------------------------------------ host side
bool attn3_supports(const AttnArgs& a) {
  auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  if (a.causal || a.Tk > 256 || a.Tk <= 0) return false;
  if (!(al(a.Q) && al(a.K) && al(a.V) && al(a.O) && a.ldq % 8 == 0 && a.ldk % 8 == 0 && a.ldv % 8 == 0 && a.ldo % 8 == 0))
    return false;
  // bulk stores of whole row blocks: every block start and size must be a multiple of 16 bytes
  if (a.probs != nullptr && (!al(a.probs) || ((long long)a.Tq * a.Tk) % 4 != 0)) return false;
  return true;
}

const char* attn3_launch(const AttnArgs& a, cudaStream_t s) {
  const int NT = (a.Tk + 63) / 64;
  static PerDeviceOnce once;
  DevInfo di;
  if (const char* e = device_info(&di)) return e;
  if (const char* e = ensure_smem(attn3_kernel, smem_total(4), once, di.dev, "cudaFuncSetAttribute(attention3) failed")) return e;
  const int bytes = smem_total(NT);
  CUtensorMap tmQ, tmK, tmV;
  const char* e = make_tmap_3d(&tmQ, a.Q, (uint64_t)kD, (uint64_t)a.Tq, (uint64_t)a.B, (uint64_t)a.ldq * 2,
                               (uint64_t)a.Tq * a.ldq * 2, 64, BQ);
  if (e) return e;
  e = make_tmap_3d(&tmK, a.K, (uint64_t)kD, (uint64_t)a.Tk, (uint64_t)a.B, (uint64_t)a.ldk * 2,
                   (uint64_t)a.Tk * a.ldk * 2, 64, (uint32_t)(64 * NT));
  if (e) return e;
  e = make_tmap_3d(&tmV, a.V, (uint64_t)kD, (uint64_t)a.Tk, (uint64_t)a.B, (uint64_t)a.ldv * 2,
                   (uint64_t)a.Tk * a.ldv * 2, 64, (uint32_t)(64 * NT));
  if (e) return e;
  Dev p;
  p.O = a.O; p.ldo = a.ldo; p.probs = a.probs; p.klen = a.klen;
  p.Tq = a.Tq; p.Tk = a.Tk; p.NT = NT;
  dim3 grid((a.Tq + BQ - 1) / BQ, kH, a.B);
  const cudaError_t err = launch_pdl(attn3_kernel, grid, dim3(THREADS), (size_t)bytes, s, tmQ, tmK, tmV, p);
  note_launch();
  return err == cudaSuccess ? nullptr : "attention3 kernel launch failed";
}

}  // namespace tts
