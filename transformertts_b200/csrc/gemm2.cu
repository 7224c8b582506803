// Persistent tcgen05 / TMEM / TMA contraction, bf16 in -> bf16 out, N a multiple of 256 (sm_100a).
//
// The workhorse of the teacher-forced forward: QKV / out / FFN projections, the conv1d layers as implicit
// GEMM (tap-shifted TMA windows) and the post-norm residual + LayerNorm blocks.  One CTA per SM walks the
// (m, n) tile list with a fixed stride.  Roles:
//   warp 0      TMA producer  : A 128x64 and W 256x64 k-blocks (SWIZZLE_128B) into a 3-stage ring
//   warp 1      MMA issuer    : one thread, tcgen05.mma.cta_group::1.kind::f16 128x256x16, fp32 accumulators in
//                               TMEM, two 256-column buffers so tile i+1 is multiplied while tile i drains
//   warps 2..5  epilogue      : tcgen05.ld (thread == output row) -> +bias -> ReLU/tanh -> +alpha*PE ->
//                               [+residual -> LayerNorm(256) with the pre-norm row parked back in TMEM] ->
//                               bf16 -> 128B-swizzled shared slabs -> TMA tensor store (coalesced, async)
// The residual tile arrives by TMA into the same slabs the result leaves from (in-place per thread).
//
// Replaces: nn.Linear / LinearNorm (model/module.py:36-53), Conv1d+BatchNorm1d of ConvNormBN
// (model/module.py:18-19,28-33, BN folded), MultiheadAttention in/out projections (torch/nn/functional.py:5798,6665),
// FFN + residual + LayerNorm (torch/nn/modules/transformer.py:951-982), PositionalEncoding (model/model.py:95).
#include <stdlib.h>
#include <string.h>

#include "kernels.h"
#include "ptx.cuh"

namespace tts {

namespace g2 {
constexpr int BM = 128, BN = 256, BK = 64;
constexpr int THREADS = 64 + 256;                  // producer warp, MMA warp, 8 epilogue warps
constexpr int A_BYTES = BM * BK * 2;               // 16 KB
constexpr int B_BYTES = BN * BK * 2;               // 32 KB
constexpr int CHUNK_BYTES = BM * 32 * 2;           // 128 rows x 32 bf16 columns = 8 KB staging buffer
constexpr int MAX_N = 2048;

// Shared-memory plan.  WS ("W stationary", K <= 256, one tap): the CTA owns one 256-wide n-tile, keeps its whole
// W slice (up to 4 k-blocks = 128 KB) resident and streams only A tiles, which cuts the L2 -> SM traffic per
// output tile from 192 KB to 64 KB (the K = 256 GEMMs of this model are L2-bandwidth bound otherwise).
// F32 (fp32 heads / PostNet tail, N <= 128, never LN or WS): the first four epilogue warps stage 32-column fp32
// chunks (128-byte rows, SWIZZLE_128B) plus an optional bf16 copy and publish both with TMA stores.
constexpr int CHUNK32_BYTES = BM * 32 * 4;         // 16 KB
// PAIR: two CTAs of a cluster share one 256 x 256 output tile pair (tcgen05 cta_group::2, M = 256).  Each CTA loads
// its own 128 rows of A and only HALF of the W rows, so a streamed k-block is 32 KB instead of 48 KB and a resident W
// slice 64 KB instead of 128 KB; the bytes saved become ring stages (the kernels are operand-delivery bound).
// CHAIN (W-stationary CTA pairs only): two GEMMs per row tile in one launch,  out2 = epi2( epi1(A W1^T) W2^T ), N = K2 = 256.
// The first result (after bias / ReLU or residual + LayerNorm) is packed to bf16 into the tensor-memory columns its
// accumulator came from and feeds the second GEMM as the A operand (tcgen05.mma with A in TMEM); W1 and W2 are both
// resident.  Used for  out-projection + residual + LayerNorm -> cross-attention Q projection  and  prenet 1 -> prenet 2
// + PE: one launch, one prologue / tail and no round trip of the intermediate activation instead of two of each.
template <bool LN, bool WS, bool F32 = false, bool PAIR = false, bool CHAIN = false>
struct Plan {
  static_assert(!F32 || (!LN && !WS && !PAIR), "the fp32-output plan streams W, has no LayerNorm and no CTA pairs");
  static_assert(!CHAIN || (WS && PAIR && !F32), "chained GEMMs run on W-stationary CTA pairs");
  static constexpr int WB = PAIR ? B_BYTES / 2 : B_BYTES;            // one W k-block as this CTA holds it
  static constexpr int W_BYTES = WS ? (CHAIN ? 8 : 4) * WB : 0;      // CHAIN: W1 in k-block slots 0..3, W2 in 4..7
  static constexpr int ST = CHAIN ? 3 : (PAIR ? (WS ? (LN ? 7 : 8) : 5) : (WS ? (LN ? 3 : 4) : (F32 ? 5 : 3)));
  static constexpr int STAGE_BYTES = WS ? A_BYTES : A_BYTES + (F32 ? WB / 2 : WB);   // F32: N <= 128, half a W block
  static constexpr int RING_OFF = W_BYTES;
  static constexpr int NBUF_H = (WS || F32) ? 2 : 3;                // staging buffers per epilogue half
  static constexpr int SLAB_OFF = RING_OFF + ST * STAGE_BYTES;
  static constexpr int SLAB_BYTES = F32 ? NBUF_H * (CHUNK32_BYTES + CHUNK_BYTES) : 2 * NBUF_H * CHUNK_BYTES;
  static constexpr int ID_OFF = SLAB_OFF + SLAB_BYTES;               // 64x64 bf16 identity (residual via MMA)
  static constexpr int BAR_OFF = ID_OFF + (LN ? 8192 : 0);
  static constexpr int PAR_OFF = BAR_OFF + 256;
  static constexpr int NBIAS = (WS || LN) ? BN : MAX_N;
  static constexpr int BIAS2_OFF = PAR_OFF + (NBIAS + (LN ? 2 * BN : 0)) * 4 + (LN ? 4096 : 0);   // after the LN row partials
  static constexpr int TOTAL = BIAS2_OFF + (CHAIN ? BN * 4 : 0) + 1024;
  static_assert(TOTAL <= 232448, "shared memory plan exceeds 227 KB");
};

struct Dev {
  int T, N;
  int tap_off, kc_per_tap, num_kb;
  int tiles_per_utt, n_tiles, m_tiles, total_tiles;
  const float* bias;
  int act;
  const float* pe; float alpha; int pe_period;
  const float* ln_g; const float* ln_b;
  // fp32-output plan only
  int n_mma, w_stage_bytes, n_main, has_bf16;      // MMA N (multiple of 16), bytes of one W k-block, columns stored
  const float* res32; int ld_res32;                // fp32 residual [B*T][ld_res32] or null
  float* out_last;                                 // column N-1 -> out_last[row] (stop head) or null
  // chained plan only: second GEMM's bias / activation (pe / alpha / pe_period above apply to ITS epilogue), and whether
  // the first result is also stored (through tmO; the second leaves through tmO2)
  const float* bias2; int act2; int store1;
};
}  // namespace g2

using namespace g2;

// Four consecutive 32-column accumulator chunks of this thread's row, with the next tcgen05.ld in flight while
// f(k, chunk) runs on the current one.
template <class F>
__device__ __forceinline__ void for_chunks4(uint32_t taddr, F&& f) {
  uint32_t ra[32], rb[32];
  tmem_ld32(taddr, ra);
  tmem_ld_wait(); tmem_ld32(taddr + 32, rb); f(0, ra);
  tmem_ld_wait(); tmem_ld32(taddr + 64, ra); f(1, rb);
  tmem_ld_wait(); tmem_ld32(taddr + 96, rb); f(2, ra);
  tmem_ld_wait(); f(3, rb);
}

template <bool LN, bool WS, bool F32, bool PAIR, bool CHAIN>
__global__ void __launch_bounds__(THREADS, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
             const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmR,
             const __grid_constant__ CUtensorMap tmW2, const __grid_constant__ CUtensorMap tmO2, const Dev p) {
  // tmO: bf16 output (may be unused when F32 && !has_bf16); tmR: bf16 residual (LN) or the fp32 output (F32);
  // tmW2 / tmO2: second weight matrix / second output of the chained plan (copies of tmW / tmO otherwise)
  using P = Plan<LN, WS, F32, PAIR, CHAIN>;
  constexpr int ST = P::ST;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + P::BAR_OFF);
  uint64_t* empty = full + ST;
  uint64_t* acc_full = empty + ST;      // [2]
  uint64_t* acc_empty = acc_full + 2;   // [2]
  uint64_t* w_full = acc_empty + 2;     // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_full + 1);
  float* s_bias = reinterpret_cast<float*>(smem + P::PAR_OFF);   // [NBIAS]
  float* s_lng = s_bias + P::NBIAS;                              // [256]
  float* s_lnb = s_lng + BN;                                     // [256]
  float* s_bias2 = reinterpret_cast<float*>(smem + P::BIAS2_OFF);  // [256] (CHAIN)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // tile walk: WS -> fixed n-tile per CTA, m-tiles strided; otherwise flat (m, n) list strided by the grid.
  // PAIR: the walker is the cluster and a list entry is two consecutive m-tiles (rank 0 the even one, rank 1 the odd
  // one; an odd tile past the end loads zeros and stores nothing), so both CTAs run the same number of iterations.
  const int rank = PAIR ? (int)cluster_cta_rank() : 0;
  const int wid = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;           // walker id
  const int nwalk = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int m_units = PAIR ? (p.m_tiles + 1) >> 1 : p.m_tiles;
  const int ws_n = WS ? wid % p.n_tiles : 0;
  const int ws_m0 = WS ? wid / p.n_tiles : 0;
  const int ws_step = WS ? nwalk / p.n_tiles : 0;
  auto tile_at = [&](int i, int& m_tile, int& n_tile) -> bool {
    int mu; bool ok;
    if (WS) {
      mu = ws_m0 + i * ws_step; n_tile = ws_n;
      ok = mu < m_units;
    } else {
      const int unit = wid + i * nwalk;
      n_tile = unit % p.n_tiles; mu = unit / p.n_tiles;
      ok = unit < m_units * p.n_tiles;
    }
    m_tile = PAIR ? 2 * mu + rank : mu;
    return ok;
  };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmO);
    if (LN) tma_prefetch_desc(&tmR);
    if (CHAIN) { tma_prefetch_desc(&tmW2); tma_prefetch_desc(&tmO2); }
    for (int s = 0; s < ST; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    // PAIR: the epilogue warps hand the accumulator back on their own CTA's barrier (a cluster-scope arrive costs
    // ~1000 cycles); rank 1's otherwise idle warp 1 waits there and forwards one arrive to the issuer in rank 0
    for (int i = 0; i < 2; ++i) { mbar_init(acc_full + i, 1); mbar_init(acc_empty + i, (PAIR && rank == 0) ? 9 : 8); }
    mbar_init(w_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    if (PAIR) { tmem_alloc_cg2(tmem_slot, 512); tmem_relinquish_cg2(); }
    else { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
  }
  if (warp >= 2) {
    const int nb0 = WS ? ws_n * BN : 0;
    const int nb = WS ? BN : (F32 ? BN : p.N);
    for (int i = threadIdx.x - 64; i < nb; i += 256) s_bias[i] = (!F32 || i < p.N) ? __ldg(p.bias + nb0 + i) : 0.f;
    if (CHAIN)
      for (int i = threadIdx.x - 64; i < BN; i += 256) s_bias2[i] = __ldg(p.bias2 + i);
    if (LN) {
      for (int i = threadIdx.x - 64; i < BN; i += 256) { s_lng[i] = __ldg(p.ln_g + i); s_lnb[i] = __ldg(p.ln_b + i); }
      // 64x64 identity in the K-major SWIZZLE_128B operand layout: the residual tile is added to the accumulator
      // by four extra N=64 MMAs (x * I, exact in fp32), so it rides the prefetched A pipeline instead of the epilogue
      // (PAIR: B is split by rows over the two CTAs, this one holds identity rows 32*rank .. 32*rank+31)
      for (int ci = threadIdx.x - 64; ci < (PAIR ? 256 : 512); ci += 256) {
        const int nl = ci >> 3, c16 = ci & 7;
        const int n = nl + (PAIR ? 32 * rank : 0);
        const uint32_t one = (n >> 3) == c16 ? ((n & 1) ? 0x3F800000u : 0x00003F80u) : 0u;   // bf16 1.0 at k == n
        const int wi = (n & 7) >> 1;
        *reinterpret_cast<uint4*>(smem + P::ID_OFF + nl * 128 + ((c16 ^ (nl & 7)) << 4)) =
            make_uint4(wi == 0 ? one : 0u, wi == 1 ? one : 0u, wi == 2 ? one : 0u, wi == 3 ? one : 0u);
      }
      fence_proxy_async();
    }
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all(); else __syncthreads();      // PAIR: the peer's barriers are initialised before anyone signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch();
  // Programmatic dependent launch: only the producer thread reads what the previous kernel wrote (activation and
  // residual tiles), so only it waits for that grid -- AFTER it has issued the loads of the resident W slice, which no
  // kernel writes: the 128 KB (1.2 us) of weights arrive while the tail of the previous grid drains.  Everything the
  // other roles touch is ordered behind those loads through the full / acc_full barriers.

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      // PAIR: both CTAs load their own halves; every byte is counted on rank 0's barrier (the MMA issuer waits there),
      // which rank 0 arms with the bytes of both
      const int nload = PAIR ? 2 : 1;
      auto load = [&](void* dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1, int c2) {
        if (PAIR) tma_load_3d_cg2(dst, tm, mapa_u32(smem_u32(bar), 0), c0, c1, c2);
        else tma_load_3d(dst, tm, bar, c0, c1, c2);
      };
      if (WS) {
        if (rank == 0) mbar_expect_tx(w_full, (p.num_kb + (CHAIN ? 4 : 0)) * P::WB * nload);
        for (int kb = 0; kb < p.num_kb; ++kb)
          load(smem + kb * P::WB, &tmW, w_full, kb * BK, ws_n * BN + rank * (BN / 2), 0);
        if (CHAIN)
          for (int kb = 0; kb < 4; ++kb)
            load(smem + (4 + kb) * P::WB, &tmW2, w_full, kb * BK, rank * (BN / 2), 0);
      }
      griddep_wait();                   // activations of the previous kernel are complete and visible from here on
      int s = 0; uint32_t ph = 0;
      int m_tile, n_tile;
      for (int i = 0; tile_at(i, m_tile, n_tile); ++i) {
        const int b = m_tile / p.tiles_per_utt, t0 = (m_tile % p.tiles_per_utt) * BM, n0 = n_tile * BN;
        for (int kb = 0; kb < p.num_kb + (LN ? 4 : 0); ++kb) {
          mbar_wait(empty + s, ph ^ 1);
          uint8_t* sa = smem + P::RING_OFF + s * P::STAGE_BYTES;
          if (kb < p.num_kb) {
            const int tap = kb / p.kc_per_tap;
            const int kc = kb - tap * p.kc_per_tap;
            if (rank == 0) mbar_expect_tx(full + s, (F32 ? A_BYTES + p.w_stage_bytes : P::STAGE_BYTES) * nload);
            load(sa, &tmA, full + s, kc * BK, t0 + tap + p.tap_off, b);
            if (!WS) load(sa + A_BYTES, &tmW, full + s, kb * BK, n0 + rank * (BN / 2), 0);
          } else {                                   // residual k-block: 64 columns of the residual tile
            if (rank == 0) mbar_expect_tx(full + s, A_BYTES * nload);
            load(sa, &tmR, full + s, (kb - p.num_kb) * BK, t0, b);
          }
          if (++s == ST) { s = 0; ph ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (one thread)
    if (lane == 0 && rank == 0) {
      constexpr int MM = PAIR ? 2 * BM : BM;         // MMA rows: the pair's 256 rows, 128 in each CTA's TMEM
      const uint32_t idesc = F32 ? umma_idesc_bf16(BM, p.n_mma) : umma_idesc_bf16(MM, BN);
      auto mma = [&](uint32_t d, uint64_t da, uint64_t db, uint32_t id, uint32_t acc) {
        if (PAIR) tc_mma_bf16_cg2(d, da, db, id, acc); else tc_mma_bf16(d, da, db, id, acc);
      };
      auto commit = [&](uint64_t* bar) { if (PAIR) tc_commit_cg2(bar, 3); else tc_commit(bar); };
      int s = 0; uint32_t ph = 0;
      int m_tile, n_tile;
      if (WS) { mbar_wait(w_full, 0); tc_fence_after(); }
      for (int it = 0; tile_at(it, m_tile, n_tile); ++it) {
        const int buf = CHAIN ? 0 : (it & 1);
        const uint32_t use = static_cast<uint32_t>(it >> 1);
        // CHAIN: GEMM 1 always accumulates into buffer 0 and needs no wait: the bf16 A operand of the previous tile's
        // GEMM 2 sits in the same columns, but that GEMM 2 was issued before (the tensor pipe runs in order), and it was
        // issued only after epilogue 1 of that tile had read the accumulator out
        if (!CHAIN) { mbar_wait(acc_empty + buf, (use & 1) ^ 1); tc_fence_after(); }   // epilogue has drained this TMEM buffer
        const uint32_t tacc = tmem_base + buf * BN;
        const int total_kb = p.num_kb + (LN ? 4 : 0);
        for (int kb = 0; kb < total_kb; ++kb) {
          mbar_wait(full + s, ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + P::RING_OFF + s * P::STAGE_BYTES);
          const uint64_t da = umma_desc_sw128(sa);
          if (kb < p.num_kb) {
            const uint64_t db = umma_desc_sw128(WS ? smem_u32(smem + kb * P::WB) : sa + A_BYTES);
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              mma(tacc, da + 2 * k, db + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          } else {                                   // acc[:, 64r .. 64r+63] += residual k-block * I
            constexpr uint32_t idesc64 = umma_idesc_bf16(MM, 64);
            const uint64_t di = umma_desc_sw128(smem_u32(smem + P::ID_OFF));
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              mma(tacc + (kb - p.num_kb) * 64, da + 2 * k, di + 2 * k, idesc64, 1u);
          }
          commit(empty + s);
          if (kb == total_kb - 1) commit(acc_full + buf);
          if (++s == ST) { s = 0; ph ^= 1; }
        }
        if constexpr (CHAIN) {
          // GEMM 2: A = the tile's first result as bf16 in TMEM columns [0, 128) (two K neighbours per column),
          // B = the resident W2, accumulator = buffer 1
          mbar_wait(acc_empty, it & 1);                      // epilogue 1 of this tile has stored the A operand
          mbar_wait(acc_empty + 1, (it & 1) ^ 1);            // epilogue 2 of the previous tile has drained buffer 1
          tc_fence_after();
          const uint32_t idesc2 = umma_idesc_bf16(2 * BM, BN);
          for (int kb = 0; kb < 4; ++kb) {
            const uint64_t db = umma_desc_sw128(smem_u32(smem + (4 + kb) * P::WB));
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              tc_mma_bf16_cg2_ts(tmem_base + BN, tmem_base + kb * 32 + k * 8, db + 2 * k, idesc2, (kb > 0 || k > 0) ? 1u : 0u);
          }
          commit(acc_full + 1);
        }
      }
    }
    if (PAIR && lane == 0 && rank == 1) {
      // relay: forward rank 1's accumulator hand-backs to the issuer
      const uint32_t remote = mapa_u32(smem_u32(acc_empty), 0);
      int m_tile, n_tile;
      for (int it = 0; tile_at(it, m_tile, n_tile); ++it) {
        if constexpr (CHAIN) {                               // both buffers are handed back once per tile
          mbar_wait(acc_empty, it & 1);
          mbar_arrive_cluster(remote);
          mbar_wait(acc_empty + 1, it & 1);
          mbar_arrive_cluster(remote + 8);
        } else {
          const int buf = it & 1;
          mbar_wait(acc_empty + buf, (it >> 1) & 1);
          mbar_arrive_cluster(remote + buf * 8);
        }
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ epilogue: 8 warps, thread == (output row, column half)
    // Warps 2..5 own columns 0..127 of the tile, warps 6..9 columns 128..255 (two warps per TMEM lane quadrant, two
    // per scheduler).  Each half walks its four 32-column chunks with the next tcgen05.ld in flight while the current
    // chunk is processed, stages bf16 results in its own ring of 8 KB buffers (64-byte rows, SWIZZLE_64B) and
    // publishes every chunk with one TMA store.
    const int ew = warp - 2;
    const int half = ew >> 2;
    const int q = warp & 3;                        // TMEM lane quadrant of this warp
    const int row = q * 32 + lane;
    const bool leader = (ew & 3) == 0 && lane == 0;
    constexpr int NB = P::NBUF_H;
    if constexpr (F32) {
      // ---- fp32-output plan: N <= 128, so only the first column half has work; the other four warps just hand
      // the accumulator buffer back.  thread == output row; every 32-column chunk leaves as one fp32 TMA store
      // (columns >= n_main and rows >= T are clipped by the tensor map) plus an optional bf16 copy.
      uint8_t* fbufs = smem + P::SLAB_OFF;                         // NB x 16 KB, 128-byte rows, SWIZZLE_128B
      uint8_t* hbufs = fbufs + NB * CHUNK32_BYTES;                 // NB x  8 KB,  64-byte rows, SWIZZLE_64B
      const int n_chunks = (p.N + 31) >> 5;
      const int last_chunk = (p.N - 1) >> 5, last_col = (p.N - 1) & 31;
      const uint32_t sw8 = static_cast<uint32_t>(row & 7), sw4 = static_cast<uint32_t>((row >> 1) & 3);
      int kc = 0;
      int m_tile, n_tile;
      for (int it = 0; tile_at(it, m_tile, n_tile); ++it) {
        const int b = m_tile / p.tiles_per_utt, t0 = (m_tile % p.tiles_per_utt) * BM;
        const int buf = it & 1;
        const uint32_t use = static_cast<uint32_t>(it >> 1);
        mbar_wait(acc_full + buf, use & 1);
        tc_fence_after();
        if (half == 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(acc_empty + buf);
          continue;
        }
        const uint32_t tacc = tmem_base + buf * BN + (static_cast<uint32_t>(q * 32) << 16);
        const int t_row = t0 + row;
        const bool valid = t_row < p.T;
        const size_t grow = (size_t)b * p.T + (valid ? t_row : 0);
        for (int k = 0; k < n_chunks; ++k) {
          uint32_t r[32];
          tmem_ld32(tacc + k * 32, r);
          tmem_ld_wait();
          if (k == n_chunks - 1) {                   // last read of this TMEM buffer
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_empty + buf);
          }
          float v[32];
          const float4* sb = reinterpret_cast<const float4*>(s_bias + k * 32);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b4 = sb[j];
            v[j * 4 + 0] = __uint_as_float(r[j * 4 + 0]) + b4.x; v[j * 4 + 1] = __uint_as_float(r[j * 4 + 1]) + b4.y;
            v[j * 4 + 2] = __uint_as_float(r[j * 4 + 2]) + b4.z; v[j * 4 + 3] = __uint_as_float(r[j * 4 + 3]) + b4.w;
          }
          if (p.act == ACT_RELU) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
          } else if (p.act == ACT_TANH) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fast_tanh(v[j]);
          }
          if (p.res32 != nullptr && valid) {
            const float4* rr = reinterpret_cast<const float4*>(p.res32 + grow * p.ld_res32 + k * 32);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (k * 32 + j * 4 < p.n_main) {       // n_main % 4 == 0 when a residual is present
                const float4 x = __ldg(rr + j);
                v[j * 4 + 0] += x.x; v[j * 4 + 1] += x.y; v[j * 4 + 2] += x.z; v[j * 4 + 3] += x.w;
              }
            }
          }
          if (p.out_last != nullptr && k == last_chunk && valid) {
            float z = 0.f;
#pragma unroll
            for (int j = 0; j < 32; ++j) z = j == last_col ? v[j] : z;
            p.out_last[grow] = z;
          }
          const int bi = kc % NB;
          uint8_t* wf = fbufs + bi * CHUNK32_BYTES + row * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(wf + ((static_cast<uint32_t>(j) ^ sw8) << 4)) =
                make_float4(v[j * 4 + 0], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
          if (p.has_bf16) {
            uint8_t* wh = hbufs + bi * CHUNK_BYTES + row * 64;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint4 u;
              u.x = pack_bf16(v[j * 8 + 0], v[j * 8 + 1]); u.y = pack_bf16(v[j * 8 + 2], v[j * 8 + 3]);
              u.z = pack_bf16(v[j * 8 + 4], v[j * 8 + 5]); u.w = pack_bf16(v[j * 8 + 6], v[j * 8 + 7]);
              *reinterpret_cast<uint4*>(wh + ((static_cast<uint32_t>(j) ^ sw4) << 4)) = u;
            }
          }
          fence_proxy_async();
          if (leader) bulk_wait_read<NB - 2>();      // the buffer the next chunk goes into has been read out
          named_bar_sync(1, 128);
          if (leader) {
            if (k * 32 < p.n_main) {
              tma_store_3d(&tmR, fbufs + bi * CHUNK32_BYTES, k * 32, t0, b);
              if (p.has_bf16) tma_store_3d(&tmO, hbufs + bi * CHUNK_BYTES, k * 32, t0, b);
            }
            bulk_commit();
          }
          ++kc;
        }
      }
      if (leader) bulk_wait<0>();
    } else if constexpr (CHAIN) {
      // ---- chained plan: epilogue 1 (bias + act, or residual + LayerNorm) -> bf16 -> [TMA store] + A operand of GEMM 2
      // in TMEM -> epilogue 2 (bias + act + PE) -> bf16 -> TMA store.  thread == (output row, column half), as below.
      uint8_t* bufs = smem + P::SLAB_OFF + half * (NB * CHUNK_BYTES);
      const uint32_t sw = static_cast<uint32_t>((row >> 1) & 3);
      uint8_t* my_row = bufs + row * 64;
      float2* s_stat = reinterpret_cast<float2*>(s_lnb + BN);
      const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
      int kc = 0;
      auto stage_and_store = [&](const CUtensorMap* tm, const uint32_t (&pk)[16], int c, int t0, int b) {
        const int bi = kc % NB;
        uint8_t* ws = my_row + bi * CHUNK_BYTES;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          *reinterpret_cast<uint4*>(ws + ((static_cast<uint32_t>(j) ^ sw) << 4)) =
              make_uint4(pk[j * 4], pk[j * 4 + 1], pk[j * 4 + 2], pk[j * 4 + 3]);
        fence_proxy_async();
        if (leader) bulk_wait_read<NB - 2>();
        named_bar_sync(1 + half, 128);
        if (leader) {
          tma_store_3d(tm, bufs + bi * CHUNK_BYTES, c * 32, t0, b);
          bulk_commit();
        }
        ++kc;
      };
      int m_tile, n_tile;
      for (int it = 0; tile_at(it, m_tile, n_tile); ++it) {
        const int b = m_tile / p.tiles_per_utt, t0 = (m_tile % p.tiles_per_utt) * BM;
        const uint32_t tacc = tmem_base + half * 128 + lane_off;
        const float* sbias_h = s_bias + half * 128;
        mbar_wait(acc_full, it & 1);
        tc_fence_after();
        float mean = 0.f, rstd = 1.f;
        if (LN) {
          float sum = 0.f, sq = 0.f;
          for_chunks4(tacc, [&](int k, const uint32_t (&r)[32]) {
            const float4* sb = reinterpret_cast<const float4*>(sbias_h + k * 32);
            float s4[4] = {0.f, 0.f, 0.f, 0.f}, q4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 b4 = sb[j];
              const float v0 = __uint_as_float(r[j * 4 + 0]) + b4.x, v1 = __uint_as_float(r[j * 4 + 1]) + b4.y;
              const float v2 = __uint_as_float(r[j * 4 + 2]) + b4.z, v3 = __uint_as_float(r[j * 4 + 3]) + b4.w;
              s4[0] += v0; s4[1] += v1; s4[2] += v2; s4[3] += v3;
              q4[0] = fmaf(v0, v0, q4[0]); q4[1] = fmaf(v1, v1, q4[1]); q4[2] = fmaf(v2, v2, q4[2]); q4[3] = fmaf(v3, v3, q4[3]);
            }
            sum += (s4[0] + s4[1]) + (s4[2] + s4[3]);
            sq += (q4[0] + q4[1]) + (q4[2] + q4[3]);
          });
          float2* st = s_stat + (it & 1) * 256;
          st[half * 128 + row] = make_float2(sum, sq);
          named_bar_sync(3, 256);
          const float2 o = st[(half ^ 1) * 128 + row];
          sum += o.x; sq += o.y;
          mean = sum * (1.f / BN);
          rstd = rsqrtf(fmaxf(sq * (1.f / BN) - mean * mean, 0.f) + 1e-5f);
        }
        uint32_t a2[4][16];                          // this thread's 128 first-stage outputs as bf16 pairs
        for_chunks4(tacc, [&](int k, const uint32_t (&r)[32]) {
          const int c = half * 4 + k;
          const float4* sb = reinterpret_cast<const float4*>(sbias_h + k * 32);
          float v[32];
          if (LN) {
            const float4* g4 = reinterpret_cast<const float4*>(s_lng + c * 32);
            const float4* b4 = reinterpret_cast<const float4*>(s_lnb + c * 32);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 g = g4[j], bb = b4[j], bi = sb[j];
              v[j * 4 + 0] = (__uint_as_float(r[j * 4 + 0]) + bi.x - mean) * (rstd * g.x) + bb.x;
              v[j * 4 + 1] = (__uint_as_float(r[j * 4 + 1]) + bi.y - mean) * (rstd * g.y) + bb.y;
              v[j * 4 + 2] = (__uint_as_float(r[j * 4 + 2]) + bi.z - mean) * (rstd * g.z) + bb.z;
              v[j * 4 + 3] = (__uint_as_float(r[j * 4 + 3]) + bi.w - mean) * (rstd * g.w) + bb.w;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 b4 = sb[j];
              v[j * 4 + 0] = __uint_as_float(r[j * 4 + 0]) + b4.x; v[j * 4 + 1] = __uint_as_float(r[j * 4 + 1]) + b4.y;
              v[j * 4 + 2] = __uint_as_float(r[j * 4 + 2]) + b4.z; v[j * 4 + 3] = __uint_as_float(r[j * 4 + 3]) + b4.w;
            }
            if (p.act == ACT_RELU) {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
            }
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) a2[k][j] = pack_bf16(v[2 * j], v[2 * j + 1]);
          if (p.store1) stage_and_store(&tmO, a2[k], c, t0, b);
        });
        // every thread of both halves has read its accumulator columns: the bf16 A operand may overwrite them
        named_bar_sync(3, 256);
#pragma unroll
        for (int k = 0; k < 4; ++k) tmem_st16(tmem_base + half * 64 + k * 16 + lane_off, a2[k]);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_empty);

        // ---- epilogue 2
        mbar_wait(acc_full + 1, it & 1);
        tc_fence_after();
        const int t_row = t0 + row;
        const int t_pe = p.pe != nullptr ? ((b * p.T + min(t_row, p.T - 1)) % p.pe_period) : 0;
        for_chunks4(tmem_base + BN + half * 128 + lane_off, [&](int k, const uint32_t (&r)[32]) {
          const int c = half * 4 + k;
          if (k == 3) {                              // last read of buffer 1: hand it back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_empty + 1);
          }
          const float4* sb = reinterpret_cast<const float4*>(s_bias2 + c * 32);
          float v[32];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b4 = sb[j];
            v[j * 4 + 0] = __uint_as_float(r[j * 4 + 0]) + b4.x; v[j * 4 + 1] = __uint_as_float(r[j * 4 + 1]) + b4.y;
            v[j * 4 + 2] = __uint_as_float(r[j * 4 + 2]) + b4.z; v[j * 4 + 3] = __uint_as_float(r[j * 4 + 3]) + b4.w;
          }
          if (p.act2 == ACT_RELU) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
          }
          if (p.pe != nullptr) {
            const float4* pr = reinterpret_cast<const float4*>(p.pe + (size_t)t_pe * kD + c * 32);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 q4 = __ldg(pr + j);
              v[j * 4 + 0] = fmaf(p.alpha, q4.x, v[j * 4 + 0]); v[j * 4 + 1] = fmaf(p.alpha, q4.y, v[j * 4 + 1]);
              v[j * 4 + 2] = fmaf(p.alpha, q4.z, v[j * 4 + 2]); v[j * 4 + 3] = fmaf(p.alpha, q4.w, v[j * 4 + 3]);
            }
          }
          uint32_t pk[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) pk[j] = pack_bf16(v[2 * j], v[2 * j + 1]);
          stage_and_store(&tmO2, pk, c, t0, b);
        });
      }
      if (leader) bulk_wait<0>();
    } else {
    uint8_t* bufs = smem + P::SLAB_OFF + half * (NB * CHUNK_BYTES);
    const uint32_t sw = static_cast<uint32_t>((row >> 1) & 3);
    uint8_t* my_row = bufs + row * 64;             // + buffer * CHUNK_BYTES + ((chunk16 ^ sw) << 4)
    float2* s_stat = reinterpret_cast<float2*>(s_lnb + BN);     // [2 parities][2 halves][128 rows] (LN only)
    int kc = 0;                                    // chunks this half has staged so far (selects the buffer)
    int m_tile, n_tile;
    for (int it = 0; tile_at(it, m_tile, n_tile); ++it) {
      const int b = m_tile / p.tiles_per_utt, t0 = (m_tile % p.tiles_per_utt) * BM, n0 = n_tile * BN;
      const int buf = it & 1;
      const uint32_t use = static_cast<uint32_t>(it >> 1);
      const uint32_t tacc = tmem_base + buf * BN + half * 128 + (static_cast<uint32_t>(q * 32) << 16);
      const float* sbias_h = s_bias + (WS ? 0 : n0) + half * 128;
      mbar_wait(acc_full + buf, use & 1);
      tc_fence_after();

      float mean = 0.f, rstd = 1.f;
      if (LN) {
        // pass 1: row statistics of v = acc + bias (the residual is already in the accumulator)
        float sum = 0.f, sq = 0.f;
        for_chunks4(tacc, [&](int k, const uint32_t (&r)[32]) {
          const float4* sb = reinterpret_cast<const float4*>(sbias_h + k * 32);
          float s4[4] = {0.f, 0.f, 0.f, 0.f}, q4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b4 = sb[j];
            const float v0 = __uint_as_float(r[j * 4 + 0]) + b4.x, v1 = __uint_as_float(r[j * 4 + 1]) + b4.y;
            const float v2 = __uint_as_float(r[j * 4 + 2]) + b4.z, v3 = __uint_as_float(r[j * 4 + 3]) + b4.w;
            s4[0] += v0; s4[1] += v1; s4[2] += v2; s4[3] += v3;
            q4[0] = fmaf(v0, v0, q4[0]); q4[1] = fmaf(v1, v1, q4[1]); q4[2] = fmaf(v2, v2, q4[2]); q4[3] = fmaf(v3, v3, q4[3]);
          }
          sum += (s4[0] + s4[1]) + (s4[2] + s4[3]);
          sq += (q4[0] + q4[1]) + (q4[2] + q4[3]);
        });
        float2* st = s_stat + (it & 1) * 256;
        st[half * 128 + row] = make_float2(sum, sq);
        named_bar_sync(3, 256);                      // both halves of every row have published their partial sums
        const float2 o = st[(half ^ 1) * 128 + row];
        sum += o.x; sq += o.y;
        mean = sum * (1.f / BN);
        rstd = rsqrtf(fmaxf(sq * (1.f / BN) - mean * mean, 0.f) + 1e-5f);
      }

      const int t_row = t0 + row;
      const int t_pe = p.pe != nullptr ? ((b * p.T + min(t_row, p.T - 1)) % p.pe_period) : 0;
      for_chunks4(tacc, [&](int k, const uint32_t (&r)[32]) {
        const int c = half * 4 + k;                  // chunk index within the 256-column tile
        if (k == 3) {                                // last read of this TMEM buffer: hand it back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(acc_empty + buf);
        }
        float v[32];
        const float4* sb = reinterpret_cast<const float4*>(sbias_h + k * 32);
        if (LN) {
          const float4* g4 = reinterpret_cast<const float4*>(s_lng + c * 32);
          const float4* b4 = reinterpret_cast<const float4*>(s_lnb + c * 32);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 g = g4[j], bb = b4[j], bi = sb[j];
            v[j * 4 + 0] = (__uint_as_float(r[j * 4 + 0]) + bi.x - mean) * (rstd * g.x) + bb.x;
            v[j * 4 + 1] = (__uint_as_float(r[j * 4 + 1]) + bi.y - mean) * (rstd * g.y) + bb.y;
            v[j * 4 + 2] = (__uint_as_float(r[j * 4 + 2]) + bi.z - mean) * (rstd * g.z) + bb.z;
            v[j * 4 + 3] = (__uint_as_float(r[j * 4 + 3]) + bi.w - mean) * (rstd * g.w) + bb.w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b4 = sb[j];
            v[j * 4 + 0] = __uint_as_float(r[j * 4 + 0]) + b4.x; v[j * 4 + 1] = __uint_as_float(r[j * 4 + 1]) + b4.y;
            v[j * 4 + 2] = __uint_as_float(r[j * 4 + 2]) + b4.z; v[j * 4 + 3] = __uint_as_float(r[j * 4 + 3]) + b4.w;
          }
          if (p.act == ACT_RELU) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
          } else if (p.act == ACT_TANH) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fast_tanh(v[j]);
          }
          if (p.pe != nullptr) {
            const float4* pr = reinterpret_cast<const float4*>(p.pe + (size_t)t_pe * kD + n0 + c * 32);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 q4 = __ldg(pr + j);
              v[j * 4 + 0] = fmaf(p.alpha, q4.x, v[j * 4 + 0]); v[j * 4 + 1] = fmaf(p.alpha, q4.y, v[j * 4 + 1]);
              v[j * 4 + 2] = fmaf(p.alpha, q4.z, v[j * 4 + 2]); v[j * 4 + 3] = fmaf(p.alpha, q4.w, v[j * 4 + 3]);
            }
          }
        }
        const int bi = kc % NB;
        uint8_t* ws = my_row + bi * CHUNK_BYTES;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 u;
          u.x = pack_bf16(v[j * 8 + 0], v[j * 8 + 1]); u.y = pack_bf16(v[j * 8 + 2], v[j * 8 + 3]);
          u.z = pack_bf16(v[j * 8 + 4], v[j * 8 + 5]); u.w = pack_bf16(v[j * 8 + 6], v[j * 8 + 7]);
          *reinterpret_cast<uint4*>(ws + ((static_cast<uint32_t>(j) ^ sw) << 4)) = u;
        }
        fence_proxy_async();
        // the buffer the next chunk goes into must have been read out by the store that last used it
        if (leader) bulk_wait_read<NB - 2>();
        named_bar_sync(1 + half, 128);
        if (leader) {
          tma_store_3d(&tmO, bufs + bi * CHUNK_BYTES, n0 + c * 32, t0, b);
          bulk_commit();
        }
        ++kc;
      });
    }
    if (leader) bulk_wait<0>();                     // all stores complete before the CTA retires its smem
    }  // !F32
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all(); else __syncthreads();      // PAIR: neither CTA retires while the other may still signal it
  if (warp == 1) {
    if (PAIR) tmem_dealloc_cg2(tmem_base, 512); else tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------ host side
template <bool LN, bool WS, bool F32 = false, bool PAIR = false, bool CHAIN = false>
static const char* launch2(const CUtensorMap& tmA, const CUtensorMap& tmW, const CUtensorMap& tmO,
                           const CUtensorMap& tmR, const Dev& p, cudaStream_t s, const CUtensorMap* tmW2p = nullptr,
                           const CUtensorMap* tmO2p = nullptr) {
  using P = Plan<LN, WS, F32, PAIR, CHAIN>;
  const CUtensorMap& tmW2 = tmW2p != nullptr ? *tmW2p : tmW;
  const CUtensorMap& tmO2 = tmO2p != nullptr ? *tmO2p : tmO;
  static PerDeviceOnce once;
  DevInfo di;
  if (const char* e = device_info(&di)) return e;
  if (const char* e = ensure_smem(gemm2_kernel<LN, WS, F32, PAIR, CHAIN>, P::TOTAL, once, di.dev, "cudaFuncSetAttribute(gemm2) failed"))
    return e;
  const int g_sm_count = di.sm_count;
  cudaError_t err;
  if (PAIR) {
    const int units = ((p.m_tiles + 1) / 2) * p.n_tiles;   // list entries: pairs of m-tiles
    int pairs = units < g_sm_count / 2 ? units : g_sm_count / 2;
    if (WS) pairs = (pairs / p.n_tiles) * p.n_tiles;      // every pair owns one n-tile
    err = launch_pdl_pairs(gemm2_kernel<LN, WS, F32, PAIR, CHAIN>, dim3(2 * pairs), dim3(THREADS), P::TOTAL, s, tmA, tmW, tmO, tmR,
                           tmW2, tmO2, p);
  } else {
    int grid = p.total_tiles < g_sm_count ? p.total_tiles : g_sm_count;
    if (WS) grid = (grid / p.n_tiles) * p.n_tiles;        // every CTA owns one n-tile
    err = launch_pdl(gemm2_kernel<LN, WS, F32, PAIR, CHAIN>, dim3(grid), dim3(THREADS), P::TOTAL, s, tmA, tmW, tmO, tmR, tmW2, tmO2, p);
  }
  note_launch();
  return err == cudaSuccess ? nullptr : "gemm2 kernel launch failed";
}

// W-stationary pairs: every pair owns one n-tile, so there must be at least n_tiles pairs (74 on a B200)
static bool g_pairs_ok(int n_tiles) { return n_tiles <= 64; }

static bool aligned16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; }

// fp32-output plan: heads (N = 81: 80 mel columns + the stop logit) and the PostNet tail (N = 80, fp32 residual)
static bool gemm2_supports_f32(const GemmArgs& a) {
  if (a.out_f32 == nullptr || a.N > 128 || a.ln_g != nullptr || a.pe != nullptr || a.residual != nullptr) return false;
  const int n_main = a.out_last != nullptr ? a.N - 1 : a.N;
  if (n_main <= 0 || a.ld_f32 % 4 != 0 || !aligned16(a.out_f32)) return false;
  if (a.out_bf16 != nullptr && (a.ld_bf16 % 8 != 0 || !aligned16(a.out_bf16))) return false;
  if (a.residual_f32 != nullptr && (a.ld_res32 % 4 != 0 || n_main % 4 != 0 || !aligned16(a.residual_f32))) return false;
  return true;
}

bool gemm2_supports(const GemmArgs& a) {
  if (gemm2_supports_f32(a)) return true;
  const bool ln = a.ln_g != nullptr;
  if (a.N % BN != 0 || a.N > MAX_N) return false;
  if (a.out_bf16 == nullptr || a.out_f32 != nullptr || a.out_last != nullptr || a.residual_f32 != nullptr) return false;
  if (a.ld_bf16 % 8 != 0 || !aligned16(a.out_bf16)) return false;
  if (ln != (a.residual != nullptr)) return false;            // residual only together with LayerNorm
  if (ln && (a.N != BN || a.pe != nullptr || a.act != ACT_NONE)) return false;
  if (ln && (a.ld_res % 8 != 0 || !aligned16(a.residual))) return false;
  return true;
}

// Chained plan: out2 = epi2( epi1(A W^T) W2^T ); N = K2 = 256, K1 <= 256, one tap, CTA pairs.
bool gemm2_chain_supports(const GemmArgs& a) {
  if (a.W2 == nullptr || a.out2 == nullptr || a.bias2 == nullptr) return false;
  if (a.N != BN || a.taps != 1 || a.kc_per_tap > 4 || a.out_f32 != nullptr || a.out_last != nullptr || a.residual_f32 != nullptr)
    return false;
  if ((a.ln_g != nullptr) != (a.residual != nullptr)) return false;
  if (a.ln_g != nullptr && a.act != ACT_NONE) return false;
  if (a.act == ACT_TANH || a.act2 == ACT_TANH) return false;
  if (a.out_bf16 != nullptr && (a.ld_bf16 % 8 != 0 || !aligned16(a.out_bf16))) return false;
  if (a.ld_out2 % 8 != 0 || !aligned16(a.out2) || !aligned16(a.W2)) return false;
  if (a.residual != nullptr && (a.ld_res % 8 != 0 || !aligned16(a.residual))) return false;
  const int m_tiles = a.B * ((a.T + BM - 1) / BM);
  return m_tiles >= 2;
}

const char* gemm2_chain_launch(const GemmArgs& a, cudaStream_t s) {
  Dev p;
  memset(&p, 0, sizeof(p));
  p.T = a.T; p.N = a.N;
  p.tap_off = a.tap_off; p.kc_per_tap = a.kc_per_tap; p.num_kb = a.kc_per_tap;
  p.tiles_per_utt = (a.T + BM - 1) / BM;
  p.m_tiles = a.B * p.tiles_per_utt;
  p.n_tiles = 1; p.total_tiles = p.m_tiles;
  p.bias = a.bias; p.act = a.act;
  p.pe = a.pe; p.alpha = a.alpha; p.pe_period = a.pe != nullptr ? a.pe_period : 1;
  p.ln_g = a.ln_g; p.ln_b = a.ln_b;
  p.bias2 = a.bias2; p.act2 = a.act2; p.store1 = a.out_bf16 != nullptr;
  const int ktot = p.num_kb * BK;
  CUtensorMap tmA, tmW, tmO, tmR, tmW2, tmO2;
  const char* e = make_tmap_3d(&tmA, a.A, (uint64_t)a.C, (uint64_t)a.T, (uint64_t)a.B, (uint64_t)a.C * 2,
                               (uint64_t)a.T * a.C * 2, BK, BM);
  if (e) return e;
  e = make_tmap_3d(&tmW, a.W, (uint64_t)ktot, (uint64_t)BN, 1, (uint64_t)ktot * 2, (uint64_t)ktot * 2 * BN, BK, BN / 2);
  if (e) return e;
  e = make_tmap_3d(&tmW2, a.W2, (uint64_t)BN, (uint64_t)BN, 1, (uint64_t)BN * 2, (uint64_t)BN * 2 * BN, BK, BN / 2);
  if (e) return e;
  e = make_tmap_3d(&tmO2, a.out2, (uint64_t)BN, (uint64_t)a.T, (uint64_t)a.B, (uint64_t)a.ld_out2 * 2,
                   (uint64_t)a.T * a.ld_out2 * 2, 32, BM, 64);
  if (e) return e;
  if (a.out_bf16 != nullptr) {
    e = make_tmap_3d(&tmO, a.out_bf16, (uint64_t)BN, (uint64_t)a.T, (uint64_t)a.B, (uint64_t)a.ld_bf16 * 2,
                     (uint64_t)a.T * a.ld_bf16 * 2, 32, BM, 64);
    if (e) return e;
  } else {
    tmO = tmO2;
  }
  if (a.residual != nullptr) {
    e = make_tmap_3d(&tmR, a.residual, (uint64_t)BN, (uint64_t)a.T, (uint64_t)a.B, (uint64_t)a.ld_res * 2,
                     (uint64_t)a.T * a.ld_res * 2, 64, BM);
    if (e) return e;
    return launch2<true, true, false, true, true>(tmA, tmW, tmO, tmR, p, s, &tmW2, &tmO2);
  }
  return launch2<false, true, false, true, true>(tmA, tmW, tmO, tmO, p, s, &tmW2, &tmO2);
}

const char* gemm2_launch(const GemmArgs& a, cudaStream_t s) {
  Dev p;
  memset(&p, 0, sizeof(p));
  p.T = a.T; p.N = a.N;
  p.tap_off = a.tap_off; p.kc_per_tap = a.kc_per_tap; p.num_kb = a.taps * a.kc_per_tap;
  p.tiles_per_utt = (a.T + BM - 1) / BM;
  p.m_tiles = a.B * p.tiles_per_utt;
  p.bias = a.bias; p.act = a.act;
  p.pe = a.pe; p.alpha = a.alpha; p.pe_period = a.pe != nullptr ? a.pe_period : 1;
  p.ln_g = a.ln_g; p.ln_b = a.ln_b;
  const int ktot = p.num_kb * BK;
  CUtensorMap tmA, tmW, tmO, tmR;
  const char* e = make_tmap_3d(&tmA, a.A, (uint64_t)a.C, (uint64_t)a.T, (uint64_t)a.B, (uint64_t)a.C * 2,
                               (uint64_t)a.T * a.C * 2, BK, BM);
  if (e) return e;
  if (gemm2_supports_f32(a)) {
    p.n_tiles = 1; p.total_tiles = p.m_tiles;
    p.n_mma = (a.N + 15) & ~15;
    p.w_stage_bytes = p.n_mma * BK * 2;
    p.n_main = a.out_last != nullptr ? a.N - 1 : a.N;
    p.has_bf16 = a.out_bf16 != nullptr;
    p.res32 = a.residual_f32; p.ld_res32 = a.ld_res32;
    p.out_last = a.out_last;
    // W rows [N, n_mma) are outside the tensor: TMA zero-fills them
    e = make_tmap_3d(&tmW, a.W, (uint64_t)ktot, (uint64_t)a.N, 1, (uint64_t)ktot * 2, (uint64_t)ktot * 2 * a.N, BK,
                     (uint32_t)p.n_mma);
    if (e) return e;
    e = make_tmap_3d(&tmR, a.out_f32, (uint64_t)p.n_main, (uint64_t)a.T, (uint64_t)a.B, (uint64_t)a.ld_f32 * 4,
                     (uint64_t)a.T * a.ld_f32 * 4, 32, BM, 128, 4);
    if (e) return e;
    if (p.has_bf16) {
      e = make_tmap_3d(&tmO, a.out_bf16, (uint64_t)p.n_main, (uint64_t)a.T, (uint64_t)a.B, (uint64_t)a.ld_bf16 * 2,
                       (uint64_t)a.T * a.ld_bf16 * 2, 32, BM, 64);
      if (e) return e;
    } else {
      tmO = tmR;
    }
    return launch2<false, false, true>(tmA, tmW, tmO, tmR, p, s);
  }
  p.n_tiles = a.N / BN;
  p.total_tiles = p.m_tiles * p.n_tiles;
  static const bool no_ws = getenv("TTS_GEMM_NO_WS") != nullptr;
  const bool ws = !no_ws && a.taps == 1 && p.num_kb <= 4;
  // CTA pairs (tcgen05 cta_group::2): measured faster for the long-K streaming plans (FFN2 + LN -10 %, PostNet
  // conv -6 %) and for the W-stationary LayerNorm plan (out-projection + LN -5 %: its ring grows from 3 to 7 stages),
  // equal or slower for the other W-stationary shapes (HBM-write bound in the steady state), so the default takes
  // pairs when W streams with K >= 1024 or when the plan is W-stationary with LayerNorm.
  // TTS_GEMM_PAIR=0: never, =2: whenever the shape allows (keeps every pair plan under test).
  static const char* pair_env = getenv("TTS_GEMM_PAIR");
  const int pair_mode = pair_env ? atoi(pair_env) : 1;
  const bool pair_ok = p.m_tiles >= 2 * (ws ? p.n_tiles : 1) && (!ws || g_pairs_ok(p.n_tiles));
  const bool pair_pick = ws ? a.residual != nullptr : p.num_kb >= 16;
  const bool pair = pair_ok && (pair_mode >= 2 || (pair_mode == 1 && pair_pick));
  e = make_tmap_3d(&tmW, a.W, (uint64_t)ktot, (uint64_t)a.N, 1, (uint64_t)ktot * 2, (uint64_t)ktot * 2 * a.N, BK,
                   pair ? BN / 2 : BN);
  if (e) return e;
  e = make_tmap_3d(&tmO, a.out_bf16, (uint64_t)a.N, (uint64_t)a.T, (uint64_t)a.B, (uint64_t)a.ld_bf16 * 2,
                   (uint64_t)a.T * a.ld_bf16 * 2, 32, BM, 64);
  if (e) return e;
  if (a.residual != nullptr) {
    e = make_tmap_3d(&tmR, a.residual, (uint64_t)a.N, (uint64_t)a.T, (uint64_t)a.B, (uint64_t)a.ld_res * 2,
                     (uint64_t)a.T * a.ld_res * 2, 64, BM);
    if (e) return e;
    if (pair) return ws ? launch2<true, true, false, true>(tmA, tmW, tmO, tmR, p, s)
                        : launch2<true, false, false, true>(tmA, tmW, tmO, tmR, p, s);
    return ws ? launch2<true, true>(tmA, tmW, tmO, tmR, p, s) : launch2<true, false>(tmA, tmW, tmO, tmR, p, s);
  }
  if (pair) return ws ? launch2<false, true, false, true>(tmA, tmW, tmO, tmO, p, s)
                      : launch2<false, false, false, true>(tmA, tmW, tmO, tmO, p, s);
  return ws ? launch2<false, true>(tmA, tmW, tmO, tmO, p, s) : launch2<false, false>(tmA, tmW, tmO, tmO, p, s);
}

}  // namespace tts
