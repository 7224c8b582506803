// Fused masked-softmax attention on tcgen05 / TMEM / TMA (head dim 64, sm_100a), streaming over the keys.
//
// Replaces nn.MultiheadAttention's core as the reference calls it: scores = (q/8) k^T + (-inf masks),
// softmax over keys, P v (torch/nn/functional.py:6620-6665), masks of TransformerTTS._get_mask
// (model/model.py:229-257) derived from per-utterance lengths, and the per-head cross-attention
// probabilities of TransformerDecoderLayer._mha_block (model/layers.py:54-74) stored fp32 [B,H,Tq,Tk]
// only when asked for (long key sequences; short ones take attention3.cu).
//
// Work item = 128 queries of one (utterance, head); keys stream in 64-row tiles.  The kernel is PERSISTENT: two
// CTAs per SM walk the item list (heaviest causal query tiles first, snake order over the grid), because a
// one-item CTA spends ~5 us on launch, barrier / TMEM setup, the first Q/K/V round trip and the drain -- as much as
// its ~8 key tiles of work (measured: time vs. key tiles per CTA is 5.2 us + 0.8 us per tile).  Across items the
// three roles keep streaming: the producer prefetches the next item's Q and K/V while the current one drains, the
// MMA warp runs two S tiles ahead of the softmax, also across the item boundary.
//   warp 0      TMA producer : Q tile per item (two buffers), K/V tiles into a 3-stage ring (SWIZZLE_128B)
//   warp 1      MMA issuer   : S = Q K^T (128x64x64) into one of two TMEM buffers, O += P V (128x64x64,
//                              V consumed MN-major straight from its [key][dim] tile), fp32 accumulators in TMEM
//   warps 2..5  softmax      : thread == query row: tcgen05.ld S -> masks -> online max/sum (exp2) -> bf16 P stored
//                              back over the first 32 of the 64 S columns it came from (tcgen05.st): P V takes its A
//                              operand from tensor memory, so P never touches shared memory and tile j+1's P does not
//                              wait for P V of tile j; O is rescaled in TMEM only when the running max grew by more
//                              than 2^8 (lazy rescale), normalised once at the end
// With probs != nullptr a second sweep recomputes S per tile and stores exp2(s - m) / l through a padded smem
// transpose so every row leaves as 128-byte coalesced segments.
#include "kernels.h"
#include "ptx.cuh"

namespace tts {

namespace a2 {
constexpr int BQ = 128, BKV = 64, ST = 3;
constexpr int THREADS = 192;
constexpr int Q_BYTES = BQ * 64 * 2;            // 16 KB
constexpr int KV_TILE = BKV * 64 * 2;           // 8 KB  (K or V)
constexpr int STAGE_BYTES = 2 * KV_TILE;        // 16 KB
constexpr int Q_OFF = 0;                        // two Q buffers
constexpr int KV_OFF = 2 * Q_BYTES;
constexpr int P_OFF = KV_OFF + ST * STAGE_BYTES;            // 80 KB
constexpr int P_BYTES = 17 * 1024;                          // 16 KB P tile; 16.5 KB as fp32 [128][33] staging
constexpr int BAR_OFF = P_OFF + P_BYTES;
constexpr int TOTAL = BAR_OFF + 256 + 1024;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLazy = 8.f;                    // rescale O only when the max grows by more than 2^8

struct Dev {
  bf16* O; int ldo;
  float* probs;
  const int* klen;
  int Tq, Tk, causal;
  int B, n_qt, total;                           // work items: n_qt query tiles x 4 heads x B utterances
};

struct Item { int b, h, q0, klen, n1, n_all; };

// The CTA's walk over the item list.  Items are ordered by query tile from the last (most causal key tiles) to the
// first; round k of the grid takes items [k * grid, (k + 1) * grid), odd rounds in reverse CTA order, so every CTA
// gets a similar mix of heavy and light tiles.  All roles of a CTA run the same walk.
struct Walker {
  int k = 0;
  __device__ bool next(const Dev& p, Item& it) {
    const int g = static_cast<int>(gridDim.x), c = static_cast<int>(blockIdx.x);
    const int idx = k * g + ((k & 1) ? g - 1 - c : c);
    if (idx >= p.total) return false;
    ++k;
    const int per = kH * p.B;
    const int level = idx / per, rem = idx - level * per;
    it.h = rem % kH;
    it.b = rem / kH;
    it.q0 = (p.n_qt - 1 - level) * BQ;
    it.klen = p.klen != nullptr ? min(__ldg(p.klen + it.b), p.Tk) : p.Tk;
    int kv_end = it.klen;
    if (p.causal) kv_end = min(kv_end, it.q0 + BQ);
    it.n1 = (kv_end + BKV - 1) / BKV;                                   // sweep 1: S, softmax, P V
    it.n_all = it.n1 + (p.probs != nullptr ? (p.Tk + BKV - 1) / BKV : 0);   // sweep 2: S again, probabilities out
    return true;
  }
};
}  // namespace a2

using namespace a2;

__device__ __forceinline__ uint32_t idesc_s() { return umma_idesc_bf16(BQ, BKV); }
__device__ __forceinline__ uint32_t idesc_pv() { return umma_idesc_bf16(BQ, 64) | (1u << 16); }   // B is MN-major

__global__ void __launch_bounds__(THREADS, 2)
attn2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
             const __grid_constant__ CUtensorMap tmV, const Dev p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint64_t* q_full = reinterpret_cast<uint64_t*>(smem + BAR_OFF);   // [2]
  uint64_t* q_empty = q_full + 2;        // [2]  every S product of the item that used this Q buffer has completed
  uint64_t* kv_full = q_empty + 2;       // [ST]
  uint64_t* kv_empty = kv_full + ST;     // [ST]
  uint64_t* s_full = kv_empty + ST;      // [2]
  uint64_t* p_full = s_full + 2;         // [2]  softmax is done with S buffer (and P / O are ready)
  uint64_t* pv_done = p_full + 2;        // [2]  P V number g (counted over the CTA's whole walk) commits to pv_done[g & 1]:
                                         //      the softmax warps skip most of these waits, and a parity wait is only
                                         //      meaningful while the barrier is at most one phase away from the waiter
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pv_done + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
    for (int i = 0; i < 2; ++i) { mbar_init(q_full + i, 1); mbar_init(q_empty + i, 1); }
    for (int s = 0; s < ST; ++s) { mbar_init(kv_full + s, 1); mbar_init(kv_empty + s, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(s_full + i, 1); mbar_init(p_full + i, 4); }
    mbar_init(pv_done, 1); mbar_init(pv_done + 1, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch();
  griddep_wait();
  const uint32_t t_s0 = tmem_base, t_o = tmem_base + 128;          // S buffers at +0 / +64, O at +128

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      Walker w; Item it;
      uint32_t G = 0;                                   // tiles issued so far (all items): ring position
      for (int itc = 0; w.next(p, it); ++itc) {
        const int qb = itc & 1;
        mbar_wait(q_empty + qb, ((itc >> 1) & 1) ^ 1);
        mbar_expect_tx(q_full + qb, Q_BYTES);
        tma_load_3d(smem + Q_OFF + qb * Q_BYTES, &tmQ, q_full + qb, it.h * kDh, it.q0, it.b);
        for (int j = 0; j < it.n_all; ++j, ++G) {
          const bool sweep2 = j >= it.n1;
          const int k0 = (sweep2 ? j - it.n1 : j) * BKV;
          const int s = G % ST;
          mbar_wait(kv_empty + s, ((G / ST) & 1) ^ 1);
          uint8_t* st = smem + KV_OFF + s * STAGE_BYTES;
          mbar_expect_tx(kv_full + s, sweep2 ? KV_TILE : STAGE_BYTES);
          tma_load_3d(st, &tmK, kv_full + s, it.h * kDh, k0, it.b);
          if (!sweep2) tma_load_3d(st + KV_TILE, &tmV, kv_full + s, it.h * kDh, k0, it.b);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (one thread)
    if (lane == 0) {
      // S cursor: runs two tiles ahead of the P V cursor through the same flattened (item, tile) stream
      Walker wS; Item iS;
      bool haveS = wS.next(p, iS);
      int jS = 0, itcS = 0;
      uint32_t GS = 0;
      auto issue_next_s = [&]() {
        if (!haveS) return;
        const int qb = itcS & 1;
        if (jS == 0) mbar_wait(q_full + qb, (itcS >> 1) & 1);
        const int s = GS % ST;
        mbar_wait(kv_full + s, (GS / ST) & 1);
        tc_fence_after();
        const uint64_t dq = umma_desc_sw128(smem_u32(smem + Q_OFF + qb * Q_BYTES));
        const uint64_t dk = umma_desc_sw128(smem_u32(smem + KV_OFF + s * STAGE_BYTES));
#pragma unroll
        for (int k = 0; k < 4; ++k) tc_mma_bf16(t_s0 + (GS & 1) * 64, dq + 2 * k, dk + 2 * k, idesc_s(), k > 0 ? 1u : 0u);
        tc_commit(s_full + (GS & 1));
        if (jS >= iS.n1) tc_commit(kv_empty + s);       // sweep 2 holds no V: the stage is free once S is done
        ++GS;
        if (++jS == iS.n_all) {                         // last S of the item: its Q buffer may be refilled
          tc_commit(q_empty + qb);
          ++itcS; jS = 0;
          haveS = wS.next(p, iS);
        }
      };
      issue_next_s();
      issue_next_s();
      Walker wP; Item iP;
      uint32_t GP = 0;
      uint32_t NP = 0;                                  // P V products issued so far
      while (wP.next(p, iP)) {
        for (int j = 0; j < iP.n_all; ++j, ++GP) {
          mbar_wait(p_full + (GP & 1), (GP >> 1) & 1);   // softmax consumed this S (and wrote P, rescaled O)
          tc_fence_after();
          if (j < iP.n1) {
            const int s = GP % ST;
            const uint64_t dv = umma_desc_sw128(smem_u32(smem + KV_OFF + s * STAGE_BYTES + KV_TILE));
#pragma unroll
            for (int k = 0; k < 4; ++k)                  // 16 keys per step: P advances 8 TMEM columns, V 16 rows
              tc_mma_bf16_ts(t_o, t_s0 + (GP & 1) * 64 + 8 * k, dv + 128 * k, idesc_pv(), (j > 0 || k > 0) ? 1u : 0u);
            tc_commit(kv_empty + s);
            tc_commit(pv_done + (NP & 1));
            ++NP;
          }
          issue_next_s();
        }
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ softmax: thread == query row
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    Walker w; Item it;
    uint32_t G = 0;                                      // tiles consumed so far (all items): S buffer / phase
    uint32_t np = 0;                                     // P V products completed before this item: pv_done phase
    while (w.next(p, it)) {
      const int qi = it.q0 + row;
      const int n1 = it.n1;
      float m_run = -INFINITY, m_used = -INFINITY, l = 0.f;
      const int row_lim = p.causal ? min(it.klen, qi + 1) : it.klen;   // keys [0, row_lim) are visible to this row

      for (int j = 0; j < n1; ++j, ++G) {
        mbar_wait(s_full + (G & 1), (G >> 1) & 1);
        tc_fence_after();
        uint32_t r[64];
        {
          uint32_t (&r0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&r[0]);
          uint32_t (&r1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&r[32]);
          tmem_ld32(t_s0 + (G & 1) * 64 + lane_off, r0);
          tmem_ld32(t_s0 + (G & 1) * 64 + 32 + lane_off, r1);
          tmem_ld_wait();
        }
        const int nvis = row_lim - j * BKV;                       // visible keys of this tile: [0, nvis)
        // raw scores stay in r[]; the exponent is one FFMA: exp2(s * log2e - m).  Only tiles that cross a mask
        // boundary for some row of this warp pay for the per-element select.
        if (!__all_sync(0xffffffffu, nvis >= BKV)) {
#pragma unroll
          for (int e = 0; e < 64; ++e)
            if (e >= nvis) r[e] = 0xff800000u;                    // -inf
        }
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};   // four chains: ILP, only 2 warps per scheduler
#pragma unroll
        for (int e = 0; e < 64; e += 8) {                             // FMNMX3: two scores per instruction
          mx[0] = fmax3(mx[0], __uint_as_float(r[e]), __uint_as_float(r[e + 1]));
          mx[1] = fmax3(mx[1], __uint_as_float(r[e + 2]), __uint_as_float(r[e + 3]));
          mx[2] = fmax3(mx[2], __uint_as_float(r[e + 4]), __uint_as_float(r[e + 5]));
          mx[3] = fmax3(mx[3], __uint_as_float(r[e + 6]), __uint_as_float(r[e + 7]));
        }
        const float smax = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
        m_run = fmaxf(m_run, smax * kLog2e);
        float scale = 1.f;
        bool grew = false;
        if (m_run > m_used + kLazy || (m_used == -INFINITY && m_run > -INFINITY)) {
          scale = m_used == -INFINITY ? 0.f : fast_exp2(m_used - m_run);
          m_used = m_run;
          l *= scale;
          grew = true;
        }
        const float nmref = m_used == -INFINITY ? 0.f : -m_used;
        float sm[4] = {0.f, 0.f, 0.f, 0.f};                          // two FADD2 chains
        uint32_t pk[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          float x0, x1;                                               // FFMA2: both exponents in one issue slot
          ffma2(x0, x1, __uint_as_float(r[2 * e]), __uint_as_float(r[2 * e + 1]), kLog2e, kLog2e, nmref, nmref);
          const float p0 = fast_exp2(x0), p1 = fast_exp2(x1);
          fadd2(sm[2 * (e & 1)], sm[2 * (e & 1) + 1], sm[2 * (e & 1)], sm[2 * (e & 1) + 1], p0, p1);
          pk[e] = pack_bf16(p0, p1);
        }
        l += (sm[0] + sm[1]) + (sm[2] + sm[3]);
        if (j > 0 && __any_sync(0xffffffffu, grew)) {             // rescale: P V of tile j-1 must be done with O
          const uint32_t g = np + j - 1;                          // (s_full of this tile implies P V of tile j-2,
          mbar_wait(pv_done + (g & 1), (g >> 1) & 1);             //  so pv_done[g & 1] is at most one phase behind)
          tc_fence_after();
          {
            uint32_t o[32];
#pragma unroll
            for (int hlf = 0; hlf < 2; ++hlf) {
              tmem_ld32(t_o + hlf * 32 + lane_off, o);
              tmem_ld_wait();
#pragma unroll
              for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * scale);
              tmem_st32(t_o + hlf * 32 + lane_off, o);
            }
            tmem_st_wait();
          }
        }
        // P(j) over S(j): this thread's 64 scores are in registers, the next S into this buffer (tile j+2) is issued
        // after P V of tile j and the tensor pipe runs in order
        tmem_st32(t_s0 + (G & 1) * 64 + lane_off, pk);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(p_full + (G & 1));
      }

      // ---- O = O / l -> bf16 -> global (one 128-byte row segment per thread)
      {
        const uint32_t g = np + n1 - 1;                           // last P V of the item (tile n1-3 is known complete)
        mbar_wait(pv_done + (g & 1), (g >> 1) & 1);
      }
      tc_fence_after();
      np += n1;
      const float inv = l > 0.f ? 1.f / l : 0.f;
      {
        uint32_t o[32];
        bf16* og = p.O + ((size_t)it.b * p.Tq + qi) * p.ldo + it.h * kDh;
#pragma unroll
        for (int hlf = 0; hlf < 2; ++hlf) {
          tmem_ld32(t_o + hlf * 32 + lane_off, o);
          tmem_ld_wait();
          if (qi < p.Tq) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              uint4 u;
              u.x = pack_bf16(__uint_as_float(o[c * 8 + 0]) * inv, __uint_as_float(o[c * 8 + 1]) * inv);
              u.y = pack_bf16(__uint_as_float(o[c * 8 + 2]) * inv, __uint_as_float(o[c * 8 + 3]) * inv);
              u.z = pack_bf16(__uint_as_float(o[c * 8 + 4]) * inv, __uint_as_float(o[c * 8 + 5]) * inv);
              u.w = pack_bf16(__uint_as_float(o[c * 8 + 6]) * inv, __uint_as_float(o[c * 8 + 7]) * inv);
              *reinterpret_cast<uint4*>(og + hlf * 32 + c * 8) = u;
            }
          }
        }
      }

      // ---- sweep 2: normalised probabilities, fp32 [B][H][Tq][Tk]
      if (it.n_all > n1) {
        // exp2(s*log2e - m) / l == exp2(s*log2e - (m + log2 l))
        const float nmref2 = (m_used == -INFINITY || !(l > 0.f)) ? -INFINITY : -(m_used + log2f(l));
        float* stage = reinterpret_cast<float*>(smem + P_OFF) + q * 32 * 33;      // this warp: [32 rows][33]
        float* pg = p.probs + ((size_t)(it.b * kH + it.h) * p.Tq + it.q0 + q * 32) * p.Tk;
        const int rows_here = min(32, p.Tq - (it.q0 + q * 32));
        for (int j = n1; j < it.n_all; ++j, ++G) {
          mbar_wait(s_full + (G & 1), (G >> 1) & 1);
          tc_fence_after();
          const int k0 = (j - n1) * BKV;
          const int nvis = row_lim - k0;
#pragma unroll
          for (int hlf = 0; hlf < 2; ++hlf) {
            uint32_t r[32];
            tmem_ld32(t_s0 + (G & 1) * 64 + hlf * 32 + lane_off, r);
            tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 32; ++e)
              stage[lane * 33 + e] = (hlf * 32 + e) < nvis ? fast_exp2(fmaf(__uint_as_float(r[e]), kLog2e, nmref2)) : 0.f;
            __syncwarp();
            const int kc = k0 + hlf * 32 + lane;
            if (kc < p.Tk) {
#pragma unroll 8
              for (int rr = 0; rr < 32; ++rr)
                if (rr < rows_here) pg[(size_t)rr * p.Tk + kc] = stage[rr * 33 + lane];
            }
            __syncwarp();
          }
          // the staging bytes alias the P tile of the next item: release S (and with it the next P V) only now
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(p_full + (G & 1));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

// ------------------------------------------------------------------ host side
bool attn2_supports(const AttnArgs& a) {
  auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  return al(a.Q) && al(a.K) && al(a.V) && al(a.O) && a.ldq % 8 == 0 && a.ldk % 8 == 0 && a.ldv % 8 == 0 &&
         a.ldo % 8 == 0;
}

const char* attn2_launch(const AttnArgs& a, cudaStream_t s) {
  static PerDeviceOnce once;
  DevInfo di;
  if (const char* e = device_info(&di)) return e;
  if (const char* e = ensure_smem(attn2_kernel, TOTAL, once, di.dev, "cudaFuncSetAttribute(attention2) failed")) return e;
  const int sm_count = di.sm_count;
  CUtensorMap tmQ, tmK, tmV;
  // each map covers one head's 64 columns at a time: dims (row pitch in elements, rows, batch)
  const char* e = make_tmap_3d(&tmQ, a.Q, (uint64_t)kD, (uint64_t)a.Tq, (uint64_t)a.B, (uint64_t)a.ldq * 2,
                               (uint64_t)a.Tq * a.ldq * 2, 64, BQ);
  if (e) return e;
  e = make_tmap_3d(&tmK, a.K, (uint64_t)kD, (uint64_t)a.Tk, (uint64_t)a.B, (uint64_t)a.ldk * 2,
                   (uint64_t)a.Tk * a.ldk * 2, 64, BKV);
  if (e) return e;
  e = make_tmap_3d(&tmV, a.V, (uint64_t)kD, (uint64_t)a.Tk, (uint64_t)a.B, (uint64_t)a.ldv * 2,
                   (uint64_t)a.Tk * a.ldv * 2, 64, BKV);
  if (e) return e;
  Dev p;
  p.O = a.O; p.ldo = a.ldo; p.probs = a.probs; p.klen = a.klen;
  p.Tq = a.Tq; p.Tk = a.Tk; p.causal = a.causal;
  p.B = a.B; p.n_qt = (a.Tq + BQ - 1) / BQ; p.total = p.n_qt * kH * a.B;
  const int grid = p.total < 2 * sm_count ? p.total : 2 * sm_count;
  const cudaError_t err = launch_pdl(attn2_kernel, dim3(grid), dim3(THREADS), TOTAL, s, tmQ, tmK, tmV, p);
  note_launch();
  return err == cudaSuccess ? nullptr : "attention2 kernel launch failed";
}

}  // namespace tts
