// Single-pass masked-softmax attention with the probability store, for short key sequences (Tk <= 184, non-causal)
// on tcgen05 / TMEM / TMA.
//
// This is the cross-attention of TransformerDecoderLayer._mha_block (model/layers.py:54-74 ->
// torch/nn/functional.py:6620-6665) -- the one whose per-head probabilities the reference returns as `alignments`.
// Phoneme sequences are short (Tp ~ 150), so a whole score row fits in TMEM: S = Q K^T is produced once (one
// 128 x 64*NT MMA group), the softmax is exact (no online rescale), P V follows, and the normalised probabilities
// leave the SM as contiguous blocks: rows q0..q0+127 of probs[b][h] are adjacent in memory, so the [128][Tk] fp32 tile
// is staged in shared memory in its final layout and written with cp.async.bulk (fully coalesced, async).
//
// The kernel is bound by that store (123 MB per decoder layer at cfg 2; pure writes reach ~3.9 TB/s on this part), so
// it is PERSISTENT -- one CTA per SM walks a contiguous range of (utterance, head, 128-query tile) items -- and
// pipelined so that the store of item i drains while item i+1 is loaded, multiplied and exponentiated:
//   warp 0      TMA producer : Q tiles into two buffers; K / V rows [0, 64*NT) once per (utterance, head)
//   warp 1      MMA issuer   : S(i) -> TMEM buffer i & 1, then O(i) = P V (V consumed MN-major); S(i+1) is issued
//                              right behind P V(i), i.e. while the softmax warps stage item i
//   warps 2..9  softmax      : thread == (query row, parity of the 32-column chunk); row max and row sum are
//                              exchanged through shared memory; the score chunks are read from TMEM once and stay in
//                              registers: exp2 -> bf16 P tiles in smem (A operand of P V) -> normalised fp32 into a
//                              DEDICATED staging buffer while P V runs; each 32-row quadrant leaves with one bulk
//                              store, waited for only when the quadrant stages its next item
// Shared memory (NT = 3, Tk = 150: 203 KB): [Q x2][K][V][P tiles][staging 128 x Tk fp32]; TMEM: two buffers of
// O (64 columns) + S (192 columns).
#include <stdlib.h>

#include "kernels.h"
#include "ptx.cuh"

namespace tts {

namespace a3 {
constexpr int BQ = 128;
constexpr int THREADS = 320;           // producer warp, MMA warp, 8 softmax warps (2 per TMEM lane quadrant)
constexpr int TILE_P = BQ * 128;       // 16 KB: 128 rows x 64 bf16 (Q, or one 64-key tile of P)
constexpr int TILE_KV = 64 * 128;      // 8 KB : 64 keys x 64 dims
constexpr int MAX_TK = 184;            // staging of 128 x Tk fp32 next to the operand buffers must fit in 227 KB
constexpr float kLog2e = 1.4426950408889634f;

struct Dev {
  bf16* O; int ldo;
  float* probs;
  const int* klen;
  int Tq, Tk, NT;
  int n_qt, total;                     // q-tiles per (b, h); items
};

__host__ __device__ inline int off_k(int NT) { return 2 * TILE_P; }
__host__ __device__ inline int off_v(int NT) { return off_k(NT) + NT * TILE_KV; }
__host__ __device__ inline int off_p(int NT) { return off_v(NT) + NT * TILE_KV; }
__host__ __device__ inline int off_stage(int NT) { return off_p(NT) + NT * TILE_P; }
__host__ __device__ inline int off_bar(int NT, int Tk) { return off_stage(NT) + ((BQ * Tk * 4 + 15) & ~15); }
__host__ __device__ inline int smem_total(int NT, int Tk) { return off_bar(NT, Tk) + 256 + 4 * BQ * 4 + 1024; }
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

__global__ void __launch_bounds__(THREADS, 1)
attn3_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
             const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO, const Dev p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  const int NT = p.NT;
  uint8_t* s_q = smem;                                   // [2] Q tiles
  uint8_t* s_k = smem + off_k(NT);
  uint8_t* s_v = smem + off_v(NT);
  uint8_t* s_p = smem + off_p(NT);                       // NT bf16 P tiles
  float* s_stage = reinterpret_cast<float*>(smem + off_stage(NT));   // [128][Tk] fp32, final layout
  uint64_t* q_full = reinterpret_cast<uint64_t*>(smem + off_bar(NT, p.Tk));   // [2]
  uint64_t* q_empty = q_full + 2;                        // [2]  S(i) has consumed Q buffer
  uint64_t* kv_full = q_full + 4;                        // [1]
  uint64_t* s_full = q_full + 5;                         // [2]
  uint64_t* s_free = q_full + 7;                         // [2]  softmax warps are done with TMEM buffer (8 arrivals)
  uint64_t* p_full = q_full + 9;                         // [1]  bf16 P tiles written (8 arrivals)
  uint64_t* pv_done = q_full + 10;                       // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(q_full + 12);
  float* s_max = reinterpret_cast<float*>(q_full + 32);  // [2][128]
  float* s_sum = s_max + 2 * BQ;                         // [2][128]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // contiguous item ranges, sizes differ by at most one (the first `rem` CTAs take one more)
  const int base_n = p.total / (int)gridDim.x, rem = p.total - base_n * (int)gridDim.x;
  const int it0 = (int)blockIdx.x * base_n + min((int)blockIdx.x, rem);
  const int n_items = base_n + ((int)blockIdx.x < rem ? 1 : 0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmO);
    for (int i = 0; i < 2; ++i) {
      mbar_init(q_full + i, 1); mbar_init(q_empty + i, 1); mbar_init(s_full + i, 1); mbar_init(s_free + i, 8);
      mbar_init(pv_done + i, 1);
    }
    mbar_init(kv_full, 1); mbar_init(p_full, 8);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch();
  griddep_wait();

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int prev_bh = -1, kv_loads = 0;
      for (int it = 0; it < n_items; ++it) {
        const int item = it0 + it, bh = item / p.n_qt, qt = item - bh * p.n_qt;
        const int b = bh / kH, h = bh - b * kH, j = it & 1;
        mbar_wait(q_empty + j, ((it >> 1) & 1) ^ 1);
        mbar_expect_tx(q_full + j, TILE_P);
        tma_load_3d(s_q + j * TILE_P, &tmQ, q_full + j, h * kDh, qt * BQ, b);
        if (bh != prev_bh) {
          // the P V product of the previous item was the last reader of the old K / V rows
          if (it > 0) mbar_wait(pv_done + ((it - 1) & 1), ((it - 1) >> 1) & 1);
          mbar_expect_tx(kv_full, 2 * NT * TILE_KV);
          tma_load_3d(s_k, &tmK, kv_full, h * kDh, 0, b);
          tma_load_3d(s_v, &tmV, kv_full, h * kDh, 0, b);
          prev_bh = bh; ++kv_loads;
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (one thread)
    if (lane == 0) {
      const uint64_t dk = umma_desc_sw128(smem_u32(s_k));
      const uint32_t idesc_s = umma_idesc_bf16(BQ, 64 * NT);
      const uint32_t idesc_pv = umma_idesc_bf16(BQ, 64) | (1u << 16);          // B (= V) is MN-major
      int prev_bh = -1, kv_loads = 0, klen_b = -1, klen = p.Tk;
      for (int it = 0; it < n_items; ++it) {
        const int item = it0 + it, bh = item / p.n_qt;
        const int b = bh / kH, j = it & 1;
        const uint32_t u = static_cast<uint32_t>(it >> 1);
        if (b != klen_b) { klen_b = b; klen = p.klen != nullptr ? max(1, min(__ldg(p.klen + b), p.Tk)) : p.Tk; }
        const int nt_used = (klen + 63) >> 6;              // 64-key tiles that hold visible keys
        const uint32_t t_o = tmem_base + j * 256, t_s = t_o + 64;
        mbar_wait(q_full + j, u & 1);
        if (bh != prev_bh) { mbar_wait(kv_full, kv_loads & 1); prev_bh = bh; ++kv_loads; }
        mbar_wait(s_free + j, (u & 1) ^ 1);                // TMEM buffer j: the staging of item it-2 has read it out
        tc_fence_after();
        const uint64_t dq = umma_desc_sw128(smem_u32(s_q + j * TILE_P));
#pragma unroll
        for (int k = 0; k < 4; ++k) tc_mma_bf16(t_s, dq + 2 * k, dk + 2 * k, idesc_s, k > 0 ? 1u : 0u);
        tc_commit(s_full + j);
        tc_commit(q_empty + j);
        mbar_wait(p_full, it & 1);
        tc_fence_after();
        for (int t = 0; t < nt_used; ++t) {
          const uint64_t dp = umma_desc_sw128(smem_u32(s_p + t * TILE_P));
          const uint64_t dv = umma_desc_sw128(smem_u32(s_v + t * TILE_KV));
#pragma unroll
          for (int k = 0; k < 4; ++k)                      // 16 keys per step: P advances 32 B, V advances 16 rows
            tc_mma_bf16(t_o, dp + 2 * k, dv + 128 * k, idesc_pv, (t > 0 || k > 0) ? 1u : 0u);
        }
        tc_commit(pv_done + j);
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ softmax / store warps
    const int hf = (warp - 2) >> 2;                      // chunk parity this thread owns (chunks hf, hf + 2, hf + 4)
    const int q = warp & 3;                              // TMEM lane quadrant
    const int row = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    const uint32_t sw = static_cast<uint32_t>(row & 7);
    const int nch = 2 * NT;
    const int Tk = p.Tk;
    const bool leader = hf == 0 && lane == 0;            // issues (and waits for) the quadrant's probability stores
    const bool leader_o = hf == 1 && lane == 0;          // issues (and waits for) the quadrant's output stores
    auto row_exchange = [&](float* buf, float mine, bool is_max) {   // combine the two chunk owners of a row
      buf[hf * BQ + row] = mine;
      named_bar_sync(1, 256);
      const float other = buf[(hf ^ 1) * BQ + row];
      return is_max ? fmaxf(mine, other) : mine + other;
    };
    uint8_t* s_oq = s_p + q * 4096;                      // output staging of this quadrant: its 32 rows of P tile 0
    float* st_row = s_stage + (size_t)row * Tk;
    const int width = (Tk & 3) == 0 ? 4 : ((Tk & 1) == 0 ? 2 : 1);
    int klen_b = -1, klen = Tk;                          // key length of utterance klen_b (one global load per utterance)

    for (int it = 0; it < n_items; ++it) {
      const int item = it0 + it, bh = item / p.n_qt, qt = item - bh * p.n_qt;
      const int b = bh / kH, h = bh - b * kH, j = it & 1;
      const uint32_t u = static_cast<uint32_t>(it >> 1);
      const int q0 = qt * BQ, qi = q0 + row;
      if (b != klen_b) { klen_b = b; klen = p.klen != nullptr ? max(1, min(__ldg(p.klen + b), Tk)) : Tk; }
      const int nt_used = (klen + 63) >> 6;
      const uint32_t t_o = tmem_base + j * 256, t_s = t_o + 64;

      mbar_wait(s_full + j, u & 1);
      tc_fence_after();
      // ---- this thread's (up to) three 32-column chunks of the score row, read from TMEM ONCE and kept in registers
      // through max, exp and normalisation
      uint32_t r[3][32];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const int c = hf + 2 * k;
        if (c < nch && c * 32 < klen) tmem_ld32(t_s + c * 32 + lane_off, r[k]);
      }
      tmem_ld_wait();
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const int c = hf + 2 * k;
        if (c < nch && c * 32 < klen) {
          const int nvis = klen - c * 32;
          if (nvis < 32) {
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (e >= nvis) r[k][e] = 0xff800000u;
          }
#pragma unroll
          for (int e = 0; e < 32; e += 4) {
            mx4[0] = fmaxf(mx4[0], __uint_as_float(r[k][e])); mx4[1] = fmaxf(mx4[1], __uint_as_float(r[k][e + 1]));
            mx4[2] = fmaxf(mx4[2], __uint_as_float(r[k][e + 2])); mx4[3] = fmaxf(mx4[3], __uint_as_float(r[k][e + 3]));
          }
        }
      }
      const float mx = row_exchange(s_max, fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])), true);   // key 0 is always visible: finite
      const float nm = -mx * kLog2e;

      // ---- p = exp2(s log2e - m) in place; bf16 P tiles for the MMA (the P tiles are free: these warps have seen
      // pv_done of the previous item, and the quadrant's previous output store, staged in its rows of tile 0, is read out)
      if (leader_o) bulk_wait_read<0>();
      named_bar_sync(2 + q, 64);
      float sm4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const int c = hf + 2 * k;
        if (c >= nch) continue;
        uint8_t* p_row = s_p + (c >> 1) * TILE_P + row * 128;
        const uint32_t cb = static_cast<uint32_t>(c & 1) * 4;
        if (c * 32 < klen) {
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            const float pe = fast_exp2(fmaf(__uint_as_float(r[k][e]), kLog2e, nm));
            sm4[e & 3] += pe;
            r[k][e] = __float_as_uint(pe);
          }
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            uint4 v;
            v.x = pack_bf16(__uint_as_float(r[k][jj * 8 + 0]), __uint_as_float(r[k][jj * 8 + 1]));
            v.y = pack_bf16(__uint_as_float(r[k][jj * 8 + 2]), __uint_as_float(r[k][jj * 8 + 3]));
            v.z = pack_bf16(__uint_as_float(r[k][jj * 8 + 4]), __uint_as_float(r[k][jj * 8 + 5]));
            v.w = pack_bf16(__uint_as_float(r[k][jj * 8 + 6]), __uint_as_float(r[k][jj * 8 + 7]));
            *reinterpret_cast<uint4*>(p_row + (((cb + jj) ^ sw) << 4)) = v;
          }
        } else if ((c >> 1) < nt_used) {                   // masked half of a tile the MMA still reads
#pragma unroll
          for (int jj = 0; jj < 4; ++jj)
            *reinterpret_cast<uint4*>(p_row + (((cb + jj) ^ sw) << 4)) = make_uint4(0u, 0u, 0u, 0u);
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
      const float inv = 1.f / row_exchange(s_sum, (sm4[0] + sm4[1]) + (sm4[2] + sm4[3]), false);

      // ---- probabilities (while P V runs on the tensor pipe): normalise from registers into the quadrant's 32 staging
      // rows (final layout) once the quadrant's previous store has read them out, then one bulk store; it drains while
      // the next item is computed
      if (leader) bulk_wait_read<0>();
      named_bar_sync(2 + q, 64);
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const int c = hf + 2 * k;
        if (c * 32 >= Tk) continue;
        if (c * 32 < klen) {
#pragma unroll
          for (int e = 0; e < 32; ++e) r[k][e] = __float_as_uint(__uint_as_float(r[k][e]) * inv);
        } else {
#pragma unroll
          for (int e = 0; e < 32; ++e) r[k][e] = 0u;
        }
        float* d = st_row + c * 32;
        const int nv = Tk - c * 32;                        // columns of this chunk that exist
        if (width == 4) {
#pragma unroll
          for (int jj = 0; jj < 8; ++jj)
            if (jj * 4 < nv) *reinterpret_cast<uint4*>(d + jj * 4) = make_uint4(r[k][jj * 4], r[k][jj * 4 + 1], r[k][jj * 4 + 2], r[k][jj * 4 + 3]);
        } else if (width == 2) {
#pragma unroll
          for (int jj = 0; jj < 16; ++jj)
            if (jj * 2 < nv) *reinterpret_cast<uint2*>(d + jj * 2) = make_uint2(r[k][jj * 2], r[k][jj * 2 + 1]);
        } else {
#pragma unroll
          for (int jj = 0; jj < 32; ++jj)
            if (jj < nv) d[jj] = __uint_as_float(r[k][jj]);
        }
      }
      fence_proxy_async();
      named_bar_sync(2 + q, 64);                          // both chunk owners of this 32-row quadrant have staged
      if (leader) {
        const int rows = min(32, p.Tq - (q0 + q * 32));
        if (rows > 0) {
          float* g = p.probs + ((size_t)(b * kH + h) * p.Tq + q0 + q * 32) * Tk;
          bulk_store_1d(g, s_stage + (size_t)q * 32 * Tk, (uint32_t)(rows * Tk * 4));
        }
        bulk_commit();
      }

      // ---- O / l -> bf16 -> the quadrant's rows of P tile 0 (dead after P V; 128-byte rows, SWIZZLE_128B) -> one TMA
      // store per quadrant (16-byte global stores from registers cost ~0.9 us per item behind the probability stores)
      mbar_wait(pv_done + j, u & 1);
      tc_fence_after();
      {
        uint32_t o[32];                                    // 32 of the row's 64 head dims
        tmem_ld32(t_o + hf * 32 + lane_off, o);
        tmem_ld_wait();
        // TMEM buffer j is read out: the S product of item it+2 may overwrite it
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s_free + j);
        uint8_t* o_row = s_oq + lane * 128;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint4 v;
          v.x = pack_bf16(__uint_as_float(o[c * 8 + 0]) * inv, __uint_as_float(o[c * 8 + 1]) * inv);
          v.y = pack_bf16(__uint_as_float(o[c * 8 + 2]) * inv, __uint_as_float(o[c * 8 + 3]) * inv);
          v.z = pack_bf16(__uint_as_float(o[c * 8 + 4]) * inv, __uint_as_float(o[c * 8 + 5]) * inv);
          v.w = pack_bf16(__uint_as_float(o[c * 8 + 6]) * inv, __uint_as_float(o[c * 8 + 7]) * inv);
          *reinterpret_cast<uint4*>(o_row + (((static_cast<uint32_t>(hf) * 4 + c) ^ sw) << 4)) = v;
        }
        fence_proxy_async();
        named_bar_sync(2 + q, 64);
        if (leader_o) {
          tma_store_3d(&tmO, s_oq, h * kDh, q0 + q * 32, b);       // rows >= Tq are clipped by the tensor map
          bulk_commit();
        }
      }
    }
    if (leader || leader_o) bulk_wait<0>();              // every store has landed before the CTA retires
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------ host side
bool attn3_supports(const AttnArgs& a) {
  auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  if (a.causal || a.Tk > MAX_TK || a.Tk <= 0 || a.probs == nullptr) return false;
  if (!(al(a.Q) && al(a.K) && al(a.V) && al(a.O) && a.ldq % 8 == 0 && a.ldk % 8 == 0 && a.ldv % 8 == 0 && a.ldo % 8 == 0))
    return false;
  // bulk stores of whole row blocks: every block start and size must be a multiple of 16 bytes
  if (!al(a.probs) || ((long long)a.Tq * a.Tk) % 4 != 0) return false;
  return true;
}

const char* attn3_launch(const AttnArgs& a, cudaStream_t s) {
  const int NT = (a.Tk + 63) / 64;
  static PerDeviceOnce once;
  DevInfo di;
  if (const char* e = device_info(&di)) return e;
  if (const char* e = ensure_smem(attn3_kernel, smem_total(3, MAX_TK), once, di.dev, "cudaFuncSetAttribute(attention3) failed")) return e;
  const int bytes = smem_total(NT, a.Tk);
  CUtensorMap tmQ, tmK, tmV, tmO;
  const char* e = make_tmap_3d(&tmQ, a.Q, (uint64_t)kD, (uint64_t)a.Tq, (uint64_t)a.B, (uint64_t)a.ldq * 2,
                               (uint64_t)a.Tq * a.ldq * 2, 64, BQ);
  if (e) return e;
  e = make_tmap_3d(&tmO, a.O, (uint64_t)kD, (uint64_t)a.Tq, (uint64_t)a.B, (uint64_t)a.ldo * 2,
                   (uint64_t)a.Tq * a.ldo * 2, 64, 32);
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
  p.n_qt = (a.Tq + BQ - 1) / BQ;
  p.total = p.n_qt * kH * a.B;
  const int grid = p.total < di.sm_count ? p.total : di.sm_count;
  const cudaError_t err = launch_pdl(attn3_kernel, dim3(grid), dim3(THREADS), (size_t)bytes, s, tmQ, tmK, tmV, tmO, p);
  note_launch();
  return err == cudaSuccess ? nullptr : "attention3 kernel launch failed";
}

}  // namespace tts
