// Fused position-wise FFN block on tcgen05 / TMEM / TMA (sm_100a):
//     y = LayerNorm(x + W2 relu(W1 x + b1) + b2),   d_model 256, hidden 1024
// (torch/nn/modules/transformer.py:951-957, 980-982 as called by model/model.py:189-200 and model/layers.py:50).
//
// The two GEMMs of the block are otherwise two launches with the [rows][1024] bf16 hidden activation (105 MB at
// cfg 2) written to and read back from L2/HBM in between.  Here a CTA PAIR (cluster of two, tcgen05 cta_group::2)
// owns 256 rows of x, 128 per CTA, and walks the hidden dimension in 8 chunks of 128: S = x W1[chunk]^T lands in
// TMEM, the epilogue warps turn it into H = relu(S + b1) in bf16 and store it back over the S columns it came from
// (tcgen05.st), and Y += H W2[:, chunk]^T accumulates in TMEM with H as the tensor-memory A operand; the hidden
// activation never leaves the SM, not even into shared memory.  With 256-row MMAs each CTA
// holds only HALF of the rows of every W block, so the whole of W1/W2 (1 MB) is streamed once per 256 rows and each
// SM ingests 512 KB of it instead of 1 MB per 128 rows.  The residual is added by identity MMAs from the resident
// x tile (exact in fp32), LayerNorm runs in the final epilogue as in gemm2.cu.
//   warp 0      TMA producer : (both CTAs) own x tile (4 k-blocks, resident for the tile), then its halves of W1 / W2
//                              as a stream of 16 KB slots through a 7-slot ring, in the order the MMA warp eats them;
//                              all bytes are counted on rank 0's barriers
//   warp 1      MMA issuer   : (rank 0 only) GEMM1 of chunk c into S[c & 1] (256 x 128, K = 256), then GEMM2 of chunk
//                              c-1 into Y (256 x 256, K = 128), so the epilogue of chunk c-1 overlaps GEMM1 of chunk c;
//                              completions are multicast to the barriers of both CTAs
//   warps 2..9  epilogue     : (both CTAs) thread == (row, 64-column half of S): bias + ReLU -> bf16 -> TMEM;  at the end
//                              of the tile thread == (row, 128-column half of Y): + b2, LayerNorm, bf16, TMA store;
//                              hand-backs arrive on the CTA's own barriers, rank 1's are forwarded by its idle warp 1
// TMEM per CTA: S0 | S1 | Y = 128 + 128 + 256 columns; H(c) lives in columns 0..31 and 64..95 of S[c & 1] and is
// protected by issue order alone: GEMM1 of chunk c+2, which overwrites it, is issued after GEMM2 of chunk c.
// Shared memory per CTA: x 64 KB, store slabs 32 KB, W ring 112 KB, identity 4 KB, b1 4 KB.
#include <stdlib.h>
#include <string.h>

#include "kernels.h"
#include "ptx.cuh"

namespace tts {

namespace ffn {
constexpr int BM = 128, BK = 64, HC = 128, NHC = kF / HC;      // 8 hidden chunks
constexpr int THREADS = 64 + 256;
constexpr int BLK = 128 * BK * 2;                   // 16 KB: one [128 x 64] bf16 operand block
constexpr int NSLOT = 7;
constexpr int X_OFF = 0;                            // 4 blocks
constexpr int H_OFF = X_OFF + 4 * BLK;              // store slabs: 2 halves x 2 buffers x 8 KB
constexpr int RING_OFF = H_OFF + 2 * BLK;
constexpr int ID_OFF = RING_OFF + NSLOT * BLK;      // this CTA's 32 rows of the 64 x 64 identity, 4 KB
constexpr int B1_OFF = ID_OFF + 4096;               // b1[1024] fp32, 4 KB
constexpr int SLAB_OFF = H_OFF;
constexpr int CHUNK_BYTES = BM * 32 * 2;            // 8 KB
constexpr int BAR_OFF = B1_OFF + 4096;
constexpr int PAR_OFF = BAR_OFF + 256;              // b2[256] g[256] b[256] + LN partials
constexpr int TOTAL = PAR_OFF + 3 * kD * 4 + 4096 + 1024;
static_assert(TOTAL <= 232448, "shared memory plan exceeds 227 KB");

struct Dev {
  int M, m_tiles, n_pairs;                          // CTA pairs in the grid
  const float* b1; const float* b2; const float* ln_g; const float* ln_b;
};
}  // namespace ffn

using namespace ffn;

__global__ void __launch_bounds__(THREADS, 1)
ffn_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW1,
           const __grid_constant__ CUtensorMap tmW2, const __grid_constant__ CUtensorMap tmO, const Dev p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint64_t* x_full = reinterpret_cast<uint64_t*>(smem + BAR_OFF);
  uint64_t* x_empty = x_full + 1;
  uint64_t* w_full = x_empty + 1;        // [NSLOT]
  uint64_t* w_empty = w_full + NSLOT;    // [NSLOT]
  uint64_t* s_full = w_empty + NSLOT;    // [2]
  uint64_t* h_full = s_full + 2;         // [2]  every epilogue warp has stored its part of H(c) over S(c).  One barrier per S
                                         //      buffer: S(c+1) may be ready before the slowest warp has finished chunk c, and a
                                         //      fast warp's arrival for c+1 must not be counted into chunk c's phase
  uint64_t* y_full = h_full + 2;         // [1]
  uint64_t* y_empty = y_full + 1;        // [1]  count 8: Y drained
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(y_empty + 1);
  float* s_b2 = reinterpret_cast<float*>(smem + PAR_OFF);
  float* s_lng = s_b2 + kD;
  float* s_lnb = s_lng + kD;
  float2* s_stat = reinterpret_cast<float2*>(s_lnb + kD);      // [2 tile parities][2 halves][128 rows]
  float* s_b1 = reinterpret_cast<float*>(smem + B1_OFF);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t crank = cluster_cta_rank();

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX); tma_prefetch_desc(&tmW1); tma_prefetch_desc(&tmW2); tma_prefetch_desc(&tmO);
    mbar_init(x_full, 1); mbar_init(x_empty, 1);
    for (int s = 0; s < NSLOT; ++s) { mbar_init(w_full + s, 1); mbar_init(w_empty + s, 1); }
    // hand-backs to the issuer: the 8 epilogue warps arrive on their OWN CTA's barrier (a cluster-scope arrive costs
    // ~1000 cycles); rank 1's otherwise idle warp 1 waits there and forwards one arrive to rank 0 (hence 8 + 1)
    const uint32_t nb = crank == 0 ? 9 : 8;
    for (int i = 0; i < 2; ++i) mbar_init(s_full + i, 1);
    mbar_init(h_full, nb); mbar_init(h_full + 1, nb); mbar_init(y_full, 1); mbar_init(y_empty, nb);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc_cg2(tmem_slot, 512);
    tmem_relinquish_cg2();
  }
  if (warp >= 2) {
    const int t = threadIdx.x - 64;
    for (int i = t; i < kD; i += 256) { s_b2[i] = __ldg(p.b2 + i); s_lng[i] = __ldg(p.ln_g + i); s_lnb[i] = __ldg(p.ln_b + i); }
    for (int i = t; i < kF; i += 256) s_b1[i] = __ldg(p.b1 + i);
    // rows 32*rank .. 32*rank+31 of the 64x64 identity in the K-major SWIZZLE_128B operand layout (see gemm2.cu):
    // residual by MMA, the B operand of a pair MMA is split by rows over the two CTAs
    for (int ci = t; ci < 256; ci += 256) {
      const int nl = ci >> 3, c16 = ci & 7;
      const int n = nl + 32 * static_cast<int>(crank);
      const uint32_t one = (n >> 3) == c16 ? ((n & 1) ? 0x3F800000u : 0x00003F80u) : 0u;   // bf16 1.0 at k == n
      const int wi = (n & 7) >> 1;
      *reinterpret_cast<uint4*>(smem + ID_OFF + nl * 128 + ((c16 ^ (nl & 7)) << 4)) =
          make_uint4(wi == 0 ? one : 0u, wi == 1 ? one : 0u, wi == 2 ? one : 0u, wi == 3 ? one : 0u);
    }
    fence_proxy_async();
  }
  tc_fence_before();
  cluster_sync_all();                   // the peer's barriers are initialised before anything is signalled on them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch();                     // the producer thread alone waits for the previous grid (see below)
  const uint32_t t_y = tmem_base + 256;
  // The two CTAs of a cluster work on neighbouring row tiles in lock step: one 256-row MMA stream issued by rank 0.
  const int pair = static_cast<int>(blockIdx.x >> 1);
  auto tile_of = [&](int it) { return (it * p.n_pairs + pair) * 2 + static_cast<int>(crank); };
  auto more = [&](int it) { return (it * p.n_pairs + pair) * 2 < p.m_tiles; };   // same verdict in both CTAs

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t slot = 0, sph = 0;                       // ring position / phase
      const int r = static_cast<int>(crank);
      auto bar0 = [&](uint64_t* bar) { return mapa_u32(smem_u32(bar), 0); };     // the barrier of that name in rank 0
      // one ring slot = 16 KB of this CTA's half: two [64 x 64] blocks of W1 or one [128 x 64] block of W2
      auto next = [&]() -> uint8_t* {
        mbar_wait(w_empty + slot, sph ^ 1);             // the MMAs that read what the slot held have completed
        if (r == 0) mbar_expect_tx(w_full + slot, 2 * BLK);
        return smem + RING_OFF + slot * BLK;
      };
      auto advance = [&]() { if (++slot == NSLOT) { slot = 0; sph ^= 1; } };
      // The W blocks of a tile in the order the MMA thread consumes them: for hidden chunk hc = 0 .. NHC: two W1 blocks of
      // chunk hc (hc < NHC), then two W2 blocks of chunk hc - 1 (hc >= 1).
      int hc = 0, step = 0;                             // position in that sequence
      auto emit_next = [&]() -> bool {                  // issues the next block; false when the tile's sequence is over
        for (;;) {
          if (hc > NHC) return false;
          const int st = step;
          const int h = hc;
          if (++step == 4) { step = 0; ++hc; }
          if (st < 2 && h < NHC) {                      // W1 rows h*128 + 64 r .. + 63, k-blocks 2 st and 2 st + 1
            uint8_t* dst = next();
            tma_load_3d_cg2(dst, &tmW1, bar0(w_full + slot), (2 * st) * BK, h * HC + r * 64, 0);
            tma_load_3d_cg2(dst + BLK / 2, &tmW1, bar0(w_full + slot), (2 * st + 1) * BK, h * HC + r * 64, 0);
            advance();
            return true;
          }
          if (st >= 2 && h >= 1) {                      // W2 rows 128 r .. + 127, hidden columns of chunk h - 1
            uint8_t* dst = next();
            tma_load_3d_cg2(dst, &tmW2, bar0(w_full + slot), (h - 1) * HC + (st - 2) * BK, r * 128, 0);
            advance();
            return true;
          }
        }
      };
      // Programmatic dependent launch: the weights are written by no kernel, so the ring is filled BEFORE waiting for
      // the previous grid (whose output is this kernel's x); the other roles touch nothing of that grid except through x.
      if (more(0))
        for (int i = 0; i < NSLOT; ++i) emit_next();
      griddep_wait();
      for (int it = 0; more(it); ++it) {
        const int tile = tile_of(it);
        mbar_wait(x_empty, (it & 1) ^ 1);
        if (r == 0) mbar_expect_tx(x_full, 2 * 4 * BLK);
        for (int kb = 0; kb < 4; ++kb) tma_load_3d_cg2(smem + X_OFF + kb * BLK, &tmX, bar0(x_full), kb * BK, tile * BM, 0);
        while (emit_next()) {}
        hc = 0; step = 0;
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (one thread)
    if (lane == 0 && crank == 0) {
      constexpr uint32_t idesc1 = umma_idesc_bf16(2 * BM, 128);     // S chunk: 256 rows (128 per CTA) x 128
      constexpr uint32_t idesc2 = umma_idesc_bf16(2 * BM, 256);     // Y: 256 x 256
      constexpr uint32_t idesc64 = umma_idesc_bf16(2 * BM, 64);
      uint32_t slot = 0, sph = 0;
      uint32_t c = 0;                                   // hidden chunks issued so far (all tiles): S buffer / phases
      uint32_t g2 = 0;                                  // GEMM2 chunks issued so far: h_full phase
      auto next_slot = [&]() -> uint32_t {              // waits for the next W block, returns its smem address
        mbar_wait(w_full + slot, sph);
        tc_fence_after();
        const uint32_t a = smem_u32(smem + RING_OFF + slot * BLK);
        return a;
      };
      auto release_slot = [&]() {
        tc_commit_cg2(w_empty + slot, (uint16_t)3);     // the pair's reads of the slot are done: tell both producers
        if (++slot == NSLOT) { slot = 0; sph ^= 1; }
      };
      auto gemm2 = [&](int ch, int it) {               // Y += H(ch) W2[:, ch]^T
        if (ch == 0) {
          mbar_wait(y_empty, (it & 1) ^ 1);             // previous tile's Y has been drained
          tc_fence_after();
          // Y starts as the residual: Y[:, 64 kb .. 64 kb + 63] = x k-block * I (exact in fp32)
          const uint64_t di = umma_desc_sw128(smem_u32(smem + ID_OFF));
          for (int kb = 0; kb < 4; ++kb) {
            const uint64_t dx = umma_desc_sw128(smem_u32(smem + X_OFF + kb * BLK));
#pragma unroll
            for (int k = 0; k < 4; ++k) tc_mma_bf16_cg2(t_y + kb * 64, dx + 2 * k, di + 2 * k, idesc64, k > 0 ? 1u : 0u);
          }
        }
        mbar_wait(h_full + (g2 & 1), (g2 >> 1) & 1);
        tc_fence_after();
        const uint32_t t_h = tmem_base + (g2 & 1) * 128;            // H(ch) sits where S(ch) was: S buffer = chunk parity
        for (int kb2 = 0; kb2 < 2; ++kb2) {
          const uint64_t dw = umma_desc_sw128(next_slot());
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc_mma_bf16_cg2_ts(t_y, t_h + kb2 * 64 + k * 8, dw + 2 * k, idesc2, 1u);
          release_slot();
        }
        ++g2;
      };
      for (int it = 0; more(it); ++it) {
        mbar_wait(x_full, it & 1);
        tc_fence_after();
        for (int hc = 0; hc < NHC; ++hc, ++c) {
          const uint32_t sb = c & 1;             // S[sb] held H(c-2): its GEMM2 was issued before, the pipe runs in order
          for (int h2 = 0; h2 < 2; ++h2) {
            const uint32_t wa = next_slot();
            for (int kk = 0; kk < 2; ++kk) {
              const int kb = 2 * h2 + kk;
              const uint64_t dx = umma_desc_sw128(smem_u32(smem + X_OFF + kb * BLK));
              const uint64_t dw = umma_desc_sw128(wa + kk * (BLK / 2));
#pragma unroll
              for (int k = 0; k < 4; ++k)
                tc_mma_bf16_cg2(tmem_base + sb * 128, dx + 2 * k, dw + 2 * k, idesc1, (kb > 0 || k > 0) ? 1u : 0u);
            }
            release_slot();
          }
          tc_commit_cg2(s_full + sb, (uint16_t)3);
          if (hc == NHC - 1) tc_commit_cg2(x_empty, (uint16_t)3);   // last read of the x tile (the residual went in with chunk 0):
                                                                   // the next tile's x loads overlap the last two GEMM2s
          if (hc >= 1) gemm2(hc - 1, it);
        }
        gemm2(NHC - 1, it);
        tc_commit_cg2(y_full, (uint16_t)3);
      }
    }
    if (lane == 0 && crank == 1) {
      // ---------------------------------------------------------- relay: rank 1's hand-backs, forwarded in the order
      // the epilogue produces them (per chunk: H stored; per tile: Y drained)
      const uint32_t r_h_full = mapa_u32(smem_u32(h_full), 0), r_y_empty = mapa_u32(smem_u32(y_empty), 0);
      uint32_t c = 0;
      for (int it = 0; more(it); ++it) {
        for (int hc = 0; hc < NHC; ++hc, ++c) {
          mbar_wait(h_full + (c & 1), (c >> 1) & 1);
          mbar_arrive_cluster(r_h_full + (c & 1) * 8);
        }
        mbar_wait(y_empty, it & 1);
        mbar_arrive_cluster(r_y_empty);
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ epilogue warps
    const int ew = warp - 2;
    const int hf = ew >> 2;                             // column half: 64 of S's 128 columns, 128 of Y's 256
    const int q = warp & 3;                             // TMEM lane quadrant
    const int row = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    const bool leader = (ew & 3) == 0 && lane == 0;
    uint8_t* bufs = smem + SLAB_OFF + hf * (2 * CHUNK_BYTES);
    const uint32_t sw4 = static_cast<uint32_t>((row >> 1) & 3);
    uint8_t* my_row = bufs + row * 64;
    uint32_t c = 0;                                     // hidden chunks processed so far (all tiles)
    int kc = 0;                                         // output chunks staged so far (selects the slab buffer)
    for (int it = 0; more(it); ++it) {
      const int tile = tile_of(it);
      // ---- hidden chunks: S -> relu(S + b1) -> bf16 H
      for (int hc = 0; hc < NHC; ++hc, ++c) {
        const uint32_t sb = c & 1;
        mbar_wait(s_full + sb, (c >> 1) & 1);
        tc_fence_after();
        uint32_t r0[32], r1[32];
        tmem_ld32(tmem_base + sb * 128 + hf * 64 + lane_off, r0);
        tmem_ld_wait();
        tmem_ld32(tmem_base + sb * 128 + hf * 64 + 32 + lane_off, r1);     // in flight while the first half is processed
        const float4* b4 = reinterpret_cast<const float4*>(s_b1 + hc * HC + hf * 64);
        uint32_t pk[32];
        auto half = [&](const uint32_t (&r)[32], int o) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 bb = b4[o * 8 + j];
            const float v0 = fmaxf(__uint_as_float(r[j * 4 + 0]) + bb.x, 0.f), v1 = fmaxf(__uint_as_float(r[j * 4 + 1]) + bb.y, 0.f);
            const float v2 = fmaxf(__uint_as_float(r[j * 4 + 2]) + bb.z, 0.f), v3 = fmaxf(__uint_as_float(r[j * 4 + 3]) + bb.w, 0.f);
            pk[o * 16 + j * 2] = pack_bf16(v0, v1);
            pk[o * 16 + j * 2 + 1] = pack_bf16(v2, v3);
          }
        };
        half(r0, 0);
        tmem_ld_wait();
        half(r1, 1);
        // 64 bf16 of this row = 32 packed columns, stored over the first half of the 64 S columns they came from
        tmem_st32(tmem_base + sb * 128 + hf * 64 + lane_off, pk);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(h_full + sb);
      }

      // ---- Y = x + H W2^T (+ b2) -> LayerNorm -> bf16 -> TMA store; thread == (row, 128-column half)
      mbar_wait(y_full, it & 1);
      tc_fence_after();
      const uint32_t tacc = t_y + hf * 128 + lane_off;
      const float* sb2 = s_b2 + hf * 128;
      float sum = 0.f, sq = 0.f;
      for (int k = 0; k < 4; ++k) {
        uint32_t r[32];
        tmem_ld32(tacc + k * 32, r);
        tmem_ld_wait();
        const float4* b4 = reinterpret_cast<const float4*>(sb2 + k * 32);
        float s4[4] = {0.f, 0.f, 0.f, 0.f}, q4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 bb = b4[j];
          const float v0 = __uint_as_float(r[j * 4 + 0]) + bb.x, v1 = __uint_as_float(r[j * 4 + 1]) + bb.y;
          const float v2 = __uint_as_float(r[j * 4 + 2]) + bb.z, v3 = __uint_as_float(r[j * 4 + 3]) + bb.w;
          s4[0] += v0; s4[1] += v1; s4[2] += v2; s4[3] += v3;
          q4[0] = fmaf(v0, v0, q4[0]); q4[1] = fmaf(v1, v1, q4[1]); q4[2] = fmaf(v2, v2, q4[2]); q4[3] = fmaf(v3, v3, q4[3]);
        }
        sum += (s4[0] + s4[1]) + (s4[2] + s4[3]);
        sq += (q4[0] + q4[1]) + (q4[2] + q4[3]);
      }
      float2* st = s_stat + (it & 1) * 256;
      st[hf * 128 + row] = make_float2(sum, sq);
      named_bar_sync(3, 256);                            // both halves of every row have published their partial sums
      const float2 o = st[(hf ^ 1) * 128 + row];
      sum += o.x; sq += o.y;
      const float mean = sum * (1.f / kD);
      const float rstd = rsqrtf(fmaxf(sq * (1.f / kD) - mean * mean, 0.f) + 1e-5f);
      for (int k = 0; k < 4; ++k) {
        uint32_t r[32];
        tmem_ld32(tacc + k * 32, r);
        tmem_ld_wait();
        if (k == 3) {                                    // last read of Y: the next tile's GEMM2 may overwrite it
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(y_empty);
        }
        const int cc = hf * 4 + k;                       // 32-column chunk index within the 256 outputs
        const float4* b4 = reinterpret_cast<const float4*>(sb2 + k * 32);
        const float4* g4 = reinterpret_cast<const float4*>(s_lng + cc * 32);
        const float4* l4 = reinterpret_cast<const float4*>(s_lnb + cc * 32);
        float v[32];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 g = g4[j], bb = l4[j], bi = b4[j];
          v[j * 4 + 0] = (__uint_as_float(r[j * 4 + 0]) + bi.x - mean) * (rstd * g.x) + bb.x;
          v[j * 4 + 1] = (__uint_as_float(r[j * 4 + 1]) + bi.y - mean) * (rstd * g.y) + bb.y;
          v[j * 4 + 2] = (__uint_as_float(r[j * 4 + 2]) + bi.z - mean) * (rstd * g.z) + bb.z;
          v[j * 4 + 3] = (__uint_as_float(r[j * 4 + 3]) + bi.w - mean) * (rstd * g.w) + bb.w;
        }
        const int bi2 = kc & 1;
        uint8_t* ws = my_row + bi2 * CHUNK_BYTES;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 u;
          u.x = pack_bf16(v[j * 8 + 0], v[j * 8 + 1]); u.y = pack_bf16(v[j * 8 + 2], v[j * 8 + 3]);
          u.z = pack_bf16(v[j * 8 + 4], v[j * 8 + 5]); u.w = pack_bf16(v[j * 8 + 6], v[j * 8 + 7]);
          *reinterpret_cast<uint4*>(ws + ((static_cast<uint32_t>(j) ^ sw4) << 4)) = u;
        }
        fence_proxy_async();
        if (leader) bulk_wait_read<0>();                 // the other buffer (next chunk's) has been read out
        named_bar_sync(1 + hf, 128);
        if (leader) {
          if (tile < p.m_tiles) tma_store_3d(&tmO, bufs + bi2 * CHUNK_BYTES, cc * 32, tile * BM, 0);
          bulk_commit();
        }
        ++kc;
      }
    }
    if (leader) bulk_wait<0>();
  }
  tc_fence_before();
  cluster_sync_all();                   // no commit or remote arrive may land in a CTA that has exited
  if (warp == 1) tmem_dealloc_cg2(tmem_base, 512);
}

// ------------------------------------------------------------------ host side
bool ffn_supports(const FfnArgs& a) {
  auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  return a.M > 0 && al(a.x) && al(a.w1) && al(a.w2) && al(a.out) && a.x != a.out;
}

const char* ffn_launch(const FfnArgs& a, cudaStream_t s) {
  static PerDeviceOnce once;
  DevInfo di;
  if (const char* e = device_info(&di)) return e;
  if (const char* e = ensure_smem(ffn_kernel, TOTAL, once, di.dev, "cudaFuncSetAttribute(ffn) failed")) return e;
  const int sm_count = di.sm_count;
  CUtensorMap tmX, tmW1, tmW2, tmO;
  const char* e = make_tmap_3d(&tmX, a.x, kD, (uint64_t)a.M, 1, kD * 2, (uint64_t)a.M * kD * 2, BK, BM);
  if (e) return e;
  e = make_tmap_3d(&tmW1, a.w1, kD, kF, 1, kD * 2, (uint64_t)kF * kD * 2, BK, 64);     // this CTA's half of a hidden chunk
  if (e) return e;
  e = make_tmap_3d(&tmW2, a.w2, kF, kD, 1, kF * 2, (uint64_t)kF * kD * 2, BK, 128);
  if (e) return e;
  e = make_tmap_3d(&tmO, a.out, kD, (uint64_t)a.M, 1, kD * 2, (uint64_t)a.M * kD * 2, 32, BM, 64);
  if (e) return e;
  Dev p;
  p.M = a.M; p.m_tiles = (a.M + BM - 1) / BM;
  p.b1 = a.b1; p.b2 = a.b2; p.ln_g = a.ln_g; p.ln_b = a.ln_b;
  // clusters that can be resident at once (GPC boundaries may leave single SMs unpaired): the tile walk is sized to it
  static int max_pairs_dev[kMaxDevices] = {};        // benign race: every thread computes the same value
  int& max_pairs = max_pairs_dev[di.dev];
  if (max_pairs == 0) {
    cudaLaunchConfig_t q = {};
    q.gridDim = dim3(2 * (sm_count / 2)); q.blockDim = dim3(THREADS); q.dynamicSmemBytes = TOTAL;
    cudaLaunchAttribute qa[1];
    qa[0].id = cudaLaunchAttributeClusterDimension;
    qa[0].val.clusterDim.x = 2; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
    q.attrs = qa; q.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, ffn_kernel, &q) != cudaSuccess || n <= 0) { cudaGetLastError(); n = sm_count / 2; }
    max_pairs = n < sm_count / 2 ? n : sm_count / 2;
  }
  const int pairs_needed = (p.m_tiles + 1) / 2;
  p.n_pairs = pairs_needed < max_pairs ? pairs_needed : max_pairs;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * p.n_pairs); cfg.blockDim = dim3(THREADS); cfg.dynamicSmemBytes = TOTAL; cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 2 : 1;
  const cudaError_t err = cudaLaunchKernelEx(&cfg, ffn_kernel, tmX, tmW1, tmW2, tmO, p);
  note_launch();
  return err == cudaSuccess ? nullptr : "ffn kernel launch failed";
}

}  // namespace tts
