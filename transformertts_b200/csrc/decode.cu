// Persistent autoregressive greedy mel decode with a per-layer K/V cache (sm_100a).
//
// Replaces the loop of TransformerTTS.inference (model/model.py:354-384).  The reference re-runs
// prenet + PE + all decoder layers over the whole prefix every step and syncs with the host three
// times per step; here ONE kernel launch runs all steps.  A thread-block cluster of C = 4 G CTAs owns R utterances;
// CTA (head h = rank % 4, part g = rank / 4) keeps the R x 256 fp32 residual stream in shared memory and works
// Megatron-style, so that one exchange per sub-block is enough:
//   * attention sub-blocks: the CTA computes the query of ITS head in full and its 1/G piece of the step's key / value
//     (appended to the cache), attends over its 1/G share of the cached keys (16-byte coalesced K/V reads, online
//     softmax) and exchanges (max, sum, o[64], partial score and value piece of the step's own key) inside its head
//     group -- exchange 1; the out-projection is contracted K-SPLIT over the 64 / G head dimensions the CTA holds, and
//     the 256-wide partial sums are all-reduced over the cluster -- exchange 2 -- followed by residual + LayerNorm,
//     computed redundantly by every CTA;
//   * FFN: hidden slice 1024 / C (column-parallel, no exchange), second matrix K-split over that slice, all-reduce;
//   * prenet: first layer column-parallel, second K-split, all-reduce; heads: column-parallel, all-gather of the frame.
// 17 exchanges per step (5 per layer + prenet + heads) where splitting every projection by columns needs 33.
// Exchanges are st.async stores into the peers' shared memory (DSMEM) that complete bytes on an mbarrier of the
// destination CTA: a phase boundary is a local wait for the bytes of the phase (no release fence, no cluster
// rendezvous); one barrier.cluster per step remains (stop flag, K/V cache visibility).  The stop rule ("all rows >=
// threshold at the same step", model/model.py:382-384) is evaluated on the device through a per-step counter; no host
// round trip happens inside the loop.
#include <cooperative_groups.h>
#include <type_traits>

#include "decode.cuh"
#include "ptx.cuh"

namespace cg = cooperative_groups;

namespace tts {

constexpr int DEC_THREADS = 256;
constexpr int DEC_WARPS = DEC_THREADS / 32;
constexpr int PSTRIDE = 68;        // floats per attention partial: m, l, pad, pad, o[64]
constexpr int MAX_CLUSTER = 16;
constexpr float kLog2eD = 1.4426950408889634f;

constexpr int PMSG = 68 + 64;      // floats per attention message: m, l, d, pad, o[64], value piece (<= 64)

template <int R>
struct DecSmem {
  float x[R][kD];                  // residual stream (replicated in every CTA)
  float q[R][kDh];                 // this CTA's head: query of the step (pre-scaled)
  float kcur[R][kDh];              // this CTA's piece of the step's key / value, rounded to bf16 like the cache
  float vcur[R][kDh];
  float att[R][kDh];               // this CTA's head: normalised attention output
  float h1[R][kD];                 // prenet-1 / FFN hidden slice of this CTA (<= 256 columns)
  float melprev[R][kMelPad];       // previous frame, zero padded to 128 (all-gather target)
  float pout[2][R][kD];            // this CTA's partial outputs, staged for the all-reduce (alternating like `red`)
  float msg[R][PMSG];              // this CTA's attention message, staged for the head-group exchange
  float part[4][R][PMSG];          // messages of the head group (landing buffer)
  float wpart[DEC_WARPS][PSTRIDE]; // per-warp partials before the in-CTA combine
  float stat[R][DEC_WARPS][2];     // LayerNorm: per-warp (sum, sum of squares)
  int exit_flag;
  alignas(8) uint64_t bar[4];      // data-arrival barriers: all-reduce buffers 0 / 1, head-group messages, frame
};
// followed by the all-reduce landing buffers: float red[2][C][R][256]
template <int R>
static size_t dec_smem_bytes(int C) { return ((sizeof(DecSmem<R>) + 15) & ~size_t(15)) + (size_t)2 * C * R * kD * 4; }

// Remote (DSMEM) stores that complete bytes on an mbarrier of the DESTINATION CTA: the consumer waits on its own
// barrier for the bytes of the phase, so no release fence (MEMBAR.ALL.GPU), no cluster rendezvous and no L1
// invalidate is paid per phase (profiles/r01_v11_decode_b1_ncu.txt: 25 % + 19 % of the step were those).
__device__ __forceinline__ void st_async_f32(uint32_t raddr, float v, uint32_t rbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(raddr),
               "r"(__float_as_uint(v)), "r"(rbar)
               : "memory");
}
__device__ __forceinline__ void st_async_f32x4(uint32_t raddr, float a, float b, float c, float d, uint32_t rbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(raddr),
               "r"(__float_as_uint(a)), "r"(__float_as_uint(b)), "r"(__float_as_uint(c)), "r"(__float_as_uint(d)), "r"(rbar)
               : "memory");
}

__device__ __forceinline__ void st_async_f32x2(uint32_t raddr, float a, float b, uint32_t rbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b32 [%0], {%1, %2}, [%3];" ::"r"(raddr),
               "r"(__float_as_uint(a)), "r"(__float_as_uint(b)), "r"(rbar)
               : "memory");
}
template <int U>
__device__ __forceinline__ void st_async_vec(uint32_t raddr, const float (&v)[U], uint32_t rbar) {
  static_assert(U == 2 || U == 4, "vector width");
  if constexpr (U == 2) st_async_f32x2(raddr, v[0], v[1], rbar);
  else st_async_f32x4(raddr, v[0], v[1], v[2], v[3], rbar);
}

// One bulk copy from this CTA's shared memory into a peer's (DSMEM), completing `bytes` on an mbarrier of the
// destination CTA: ONE transaction on the destination barrier per block instead of one per 16 bytes.
__device__ __forceinline__ void bulk_copy_to_peer(uint32_t dst_cluster_addr, const void* src, uint32_t bytes,
                                                  uint32_t bar_cluster_addr) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_cluster_addr), "r"(smem_u32(src)), "r"(bytes), "r"(bar_cluster_addr)
               : "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y), c = unpack_bf16(u.z), d = unpack_bf16(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

// Reduce V = 16 per-lane partial sums over the 32 lanes of a warp with 16 shuffles instead of 16 x 5: every step
// exchanges one half of the remaining values with the lane `width` away and keeps the sums of the other half, so the
// number of live values halves with the lane distance (16, 8, 4, 2), and one butterfly step finishes the last value.
// Returns the total of value index  ((lane >> 1) & 15)  bit-reversed ... precisely: index bit 3 = lane bit 4, bit 2 =
// lane bit 3, bit 1 = lane bit 2, bit 0 = lane bit 1 (lanes 2i and 2i+1 hold the same total).
__device__ __forceinline__ float warp_reduce16(float (&v)[16], int lane) {
#pragma unroll
  for (int n = 16, width = 16; n > 1; n >>= 1, width >>= 1) {
    const bool upper = (lane & width) != 0;
#pragma unroll
    for (int j = 0; j < n / 2; ++j) {
      const float send = upper ? v[j] : v[j + n / 2];
      const float keep = upper ? v[j + n / 2] : v[j];
      v[j] = keep + __shfl_xor_sync(0xffffffffu, send, width);
    }
  }
  return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 1);
}
__device__ __forceinline__ int warp_reduce16_index(int lane) { return (lane >> 1) & 15; }

// Column-parallel GEMV: y[r][i] = sum_k in[r][k] * W[row_of(i)][k] + bias[row_of(i)] for i in [0, n_count): a warp
// takes U = 16 / R outputs per pass (16 (output, row) sums per lane), lanes stride K in 16-byte weight loads (K = 256:
// exactly one load per lane and output), and the 16 sums are reduced with warp_reduce16.  emit(i, n, r, value) is called
// by ONE lane per (output, row).
template <int R, class RowOf, class Emit>
__device__ __forceinline__ void gemv_rows(const bf16* __restrict__ W, int ldw, const float* __restrict__ bias, int K,
                                          int n_count, RowOf&& row_of, const float* in, int ld_in, Emit&& emit) {
  constexpr int U = 16 / R;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int ib = warp * U; ib < n_count; ib += DEC_WARPS * U) {
    float acc[16];                                 // [u][r]
    int nn[U];
#pragma unroll
    for (int u = 0; u < U; ++u) nn[u] = row_of(min(ib + u, n_count - 1));
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = 0.f;
    for (int k0 = lane * 8; k0 < K; k0 += 256) {
      uint4 w[U];
#pragma unroll
      for (int u = 0; u < U; ++u) w[u] = __ldg(reinterpret_cast<const uint4*>(W + (size_t)nn[u] * ldw + k0));
      float xin[R][8];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float4 a = *reinterpret_cast<const float4*>(in + r * ld_in + k0);
        const float4 b = *reinterpret_cast<const float4*>(in + r * ld_in + k0 + 4);
        xin[r][0] = a.x; xin[r][1] = a.y; xin[r][2] = a.z; xin[r][3] = a.w;
        xin[r][4] = b.x; xin[r][5] = b.y; xin[r][6] = b.z; xin[r][7] = b.w;
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        float wf[8];
        unpack8(w[u], wf);
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[u * R + r] = fmaf(wf[i], xin[r][i], acc[u * R + r]);
      }
    }
    const float total = warp_reduce16(acc, lane);
    const int idx = warp_reduce16_index(lane), u = idx / R, r = idx - u * R;
    if ((lane & 1) == 0 && ib + u < n_count) {
      const int n = row_of(ib + u);
      emit(ib + u, n, r, total + __ldg(bias + n));
    }
  }
}

// Same contraction with the result replicated in every lane (U = 1, full butterfly): the heads, whose emit publishes
// with lane p -> peer p.
template <int R, class RowOf, class Emit>
__device__ __forceinline__ void gemv_rows_bcast(const bf16* __restrict__ W, int ldw, const float* __restrict__ bias, int K,
                                                int n_count, RowOf&& row_of, const float* in, int ld_in, Emit&& emit) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = warp; i < n_count; i += DEC_WARPS) {
    const int n = row_of(i);
    float acc[R];
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = 0.f;
    for (int k0 = lane * 8; k0 < K; k0 += 256) {
      float wf[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(W + (size_t)n * ldw + k0)), wf);
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float4 a = *reinterpret_cast<const float4*>(in + r * ld_in + k0);
        const float4 b = *reinterpret_cast<const float4*>(in + r * ld_in + k0 + 4);
        acc[r] = fmaf(wf[0], a.x, acc[r]); acc[r] = fmaf(wf[1], a.y, acc[r]);
        acc[r] = fmaf(wf[2], a.z, acc[r]); acc[r] = fmaf(wf[3], a.w, acc[r]);
        acc[r] = fmaf(wf[4], b.x, acc[r]); acc[r] = fmaf(wf[5], b.y, acc[r]);
        acc[r] = fmaf(wf[6], b.z, acc[r]); acc[r] = fmaf(wf[7], b.w, acc[r]);
      }
    }
    const float bn = __ldg(bias + n);
#pragma unroll
    for (int r = 0; r < R; ++r) emit(i, n, r, warp_sum(acc[r]) + bn);
  }
}

// K-split GEMV: out[r][n] = sum_{k in slice} W[n][k] * in[r][k - k_begin] for all 256 outputs n = threadIdx.x, W given
// as its K-blocked transpose Wt[K/8][256][8] (16-byte words, consecutive threads -> consecutive words).  `in` is
// shared memory (broadcast reads).  No bias: the partial sums are all-reduced afterwards.
template <int R>
__device__ __forceinline__ void gemv_ksplit(const bf16* __restrict__ Wt, int kb_begin, int n_kb, const float* in, int ld_in,
                                            float (*out)[kD]) {
  const int n = threadIdx.x;
  const uint4* wp = reinterpret_cast<const uint4*>(Wt) + (size_t)kb_begin * kD + n;
  float acc[R];
#pragma unroll
  for (int r = 0; r < R; ++r) acc[r] = 0.f;
#pragma unroll 4
  for (int kb = 0; kb < n_kb; ++kb) {
    const uint4 w = __ldg(wp + (size_t)kb * kD);
    float wf[8];
    unpack8(w, wf);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float4 a = *reinterpret_cast<const float4*>(in + r * ld_in + kb * 8);
      const float4 b = *reinterpret_cast<const float4*>(in + r * ld_in + kb * 8 + 4);
      acc[r] = fmaf(wf[0], a.x, acc[r]); acc[r] = fmaf(wf[1], a.y, acc[r]);
      acc[r] = fmaf(wf[2], a.z, acc[r]); acc[r] = fmaf(wf[3], a.w, acc[r]);
      acc[r] = fmaf(wf[4], b.x, acc[r]); acc[r] = fmaf(wf[5], b.y, acc[r]);
      acc[r] = fmaf(wf[6], b.z, acc[r]); acc[r] = fmaf(wf[7], b.w, acc[r]);
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) out[r][n] = acc[r];
}

// x[r] = LayerNorm(z[r]) for the R rows (eps 1e-5, biased variance), z[r][tid] held by thread tid.
// R <= 2: column-parallel -- every warp reduces its 32 columns (sum and sum of squares), the eight partial pairs meet
// in shared memory and every thread normalises its own column; R > 2: one warp per row (two-pass variance).
template <int R>
__device__ __forceinline__ void layernorm_cols(float (*x)[kD], const float (&z)[R], float (*stat)[DEC_WARPS][2],
                                               const float* __restrict__ g, const float* __restrict__ b) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if constexpr (R <= 2) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float s = warp_sum(z[r]), q = warp_sum(z[r] * z[r]);
      if (lane == 0) { stat[r][warp][0] = s; stat[r][warp][1] = q; }
    }
    __syncthreads();
    const float gg = __ldg(g + tid), bb = __ldg(b + tid);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float s = 0.f, q = 0.f;
#pragma unroll
      for (int w = 0; w < DEC_WARPS; ++w) { s += stat[r][w][0]; q += stat[r][w][1]; }
      const float mean = s * (1.f / kD);
      const float rstd = rsqrtf(fmaxf(q * (1.f / kD) - mean * mean, 0.f) + 1e-5f);
      x[r][tid] = (z[r] - mean) * rstd * gg + bb;
    }
    __syncthreads();
  } else {
#pragma unroll
    for (int r = 0; r < R; ++r) x[r][tid] = z[r];
    __syncthreads();
    for (int r = warp; r < R; r += DEC_WARPS) {
      float v[8];
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) { v[i] = x[r][lane + 32 * i]; s += v[i]; }
      const float mean = warp_sum(s) * (1.f / kD);
      float q = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) { const float d = v[i] - mean; q += d * d; }
      const float rstd = rsqrtf(warp_sum(q) * (1.f / kD) + 1e-5f);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int c = lane + 32 * i;
        x[r][c] = (v[i] - mean) * rstd * __ldg(g + c) + __ldg(b + c);
      }
    }
    __syncthreads();
  }
}

struct Partial {      // one 8-lane group's running softmax state over its keys (8 head dims per lane)
  float m, l, o[8];
};

__device__ __forceinline__ void partial_update(Partial& p, float s, const float (&vf)[8]) {
  const float mn = fmaxf(p.m, s);
  const float corr = exp2f((p.m - mn) * kLog2eD);       // p.m == -inf -> 0
  const float e = exp2f((s - mn) * kLog2eD);
  p.l = p.l * corr + e;
#pragma unroll
  for (int i = 0; i < 8; ++i) p.o[i] = p.o[i] * corr + e * vf[i];
  p.m = mn;
}

// merge the state held by the lane `mask` away (same head-dim slice) into this lane's state
__device__ __forceinline__ void partial_merge_xor(Partial& p, int mask) {
  const float m2 = __shfl_xor_sync(0xffffffffu, p.m, mask);
  const float l2 = __shfl_xor_sync(0xffffffffu, p.l, mask);
  const float mn = fmaxf(p.m, m2);
  const float c1 = p.m == -INFINITY ? 0.f : exp2f((p.m - mn) * kLog2eD);
  const float c2 = m2 == -INFINITY ? 0.f : exp2f((m2 - mn) * kLog2eD);
  p.l = p.l * c1 + l2 * c2;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float o2 = __shfl_xor_sync(0xffffffffu, p.o[i], mask);
    p.o[i] = p.o[i] * c1 + o2 * c2;
  }
  p.m = mn;
}

// Attention of one (row, head) query over this warp's share of the cached keys.
//   kv        : first key row of this utterance/layer, bf16 [n_keys][512] (K at +0, V at +256)
//   q         : the head's 64 fp32 query values (already scaled)
//   unit/units: this warp's index / count among the warps sharing the (row, head) key range
__device__ __forceinline__ Partial attend_keys(const bf16* kv, int n_keys, const float* q, int head, int unit, int units) {
  const int lane = threadIdx.x & 31;
  const int grp = lane >> 3, dl = lane & 7;
  float qf[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) qf[i] = q[dl * 8 + i];
  Partial p;
  p.m = -INFINITY; p.l = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) p.o[i] = 0.f;
  const bf16* base = kv + head * kDh + dl * 8;
  const int n_groups = (n_keys + 3) >> 2;
  auto score = [&](const uint4& kq) {
    float kf[8];
    unpack8(kq, kf);
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s = fmaf(qf[i], kf[i], s);
    s += __shfl_xor_sync(0xffffffffu, s, 1); s += __shfl_xor_sync(0xffffffffu, s, 2); s += __shfl_xor_sync(0xffffffffu, s, 4);
    return s;
  };
  // NG independent 4-key groups per iteration keep 2 NG 16-byte loads in flight per lane: the per-warp key loop is bound
  // by the L2 / HBM round trip of its loads, not by their bytes.  Long key ranges per warp (many rows per cluster, late
  // steps) take four groups at a time, short ones two (measured: B = 128 +9 %, B = 1 -3 % with four everywhere).
  auto sweep = [&](auto ng_tag) {
    constexpr int NG = decltype(ng_tag)::value;
    for (int g4 = unit; g4 < n_groups; g4 += NG * units) {
      uint4 kq[NG], vq[NG];
      bool ok[NG];
#pragma unroll
      for (int u = 0; u < NG; ++u) {
        const int j = (g4 + u * units) * 4 + grp;
        ok[u] = j < n_keys;
        kq[u] = make_uint4(0, 0, 0, 0); vq[u] = kq[u];
        if (ok[u]) {
          kq[u] = __ldcg(reinterpret_cast<const uint4*>(base + (size_t)j * 512));
          vq[u] = __ldcg(reinterpret_cast<const uint4*>(base + (size_t)j * 512 + 256));
        }
      }
#pragma unroll
      for (int u = 0; u < NG; ++u) {
        const float s = score(kq[u]);
        float vf[8];
        unpack8(vq[u], vf);
        if (ok[u]) partial_update(p, s, vf);
      }
    }
  };
  if (n_groups > 4 * units) sweep(std::integral_constant<int, 4>{});
  else sweep(std::integral_constant<int, 2>{});
  partial_merge_xor(p, 8);
  partial_merge_xor(p, 16);
  return p;          // lanes 0..7 (and their mirrors) hold the warp's result for dims dl*8..dl*8+7
}

template <int R>
__global__ void __launch_bounds__(DEC_THREADS, R <= 2 ? 2 : 1) decode_kernel(const DecodeArgs a) {   // R <= 2: two CTAs per SM (<= 128 registers), which the one-wave plans count on
  cg::cluster_group cluster = cg::this_cluster();
  const int C = (int)cluster.num_blocks();
  const int rank = (int)cluster.block_rank();
  extern __shared__ __align__(16) uint8_t dec_smem_raw[];
  DecSmem<R>& sm = *reinterpret_cast<DecSmem<R>*>(dec_smem_raw);
  float* red = reinterpret_cast<float*>(dec_smem_raw + ((sizeof(DecSmem<R>) + 15) & ~size_t(15)));   // [2][C][R][256]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row0 = (blockIdx.x / C) * R;

  // work split: CTA -> (head, part of the head group); warp -> (row, sub-unit of the key range)
  const int head = rank & 3, g = rank >> 2, G = C >> 2;
  const int pk = kDh / G;                                        // head dimensions (key / value piece, K slice) of this CTA
  constexpr int WPR = DEC_WARPS / R;                             // warps per row
  const int arow = warp % R, wi = warp / R;
  const int unit = g * WPR + wi, units = G * WPR;

  for (int i = tid; i < R * kMelPad; i += DEC_THREADS) (&sm.melprev[0][0])[i] = 0.f;
  if (tid == 0) {
    sm.exit_flag = 0;
    for (int i = 0; i < 4; ++i) mbar_init(&sm.bar[i], 1);
    fence_mbar_init();
  }
  cluster.sync();

  // Phase protocol.  Every exchange publishes with st.async into the peers and waits on the local barrier of the
  // destination buffer: thread 0 arms it with the bytes the senders will deliver, everybody waits for the phase to
  // complete.  The all-reduce phases are all-to-all and alternate between two landing buffers: a CTA can reach
  // all-reduce p only after every peer has sent all-reduce p-1, i.e. after every peer has finished reading buffer
  // p-2 -- the one all-reduce p overwrites.  Two head-group exchanges are always separated by an all-reduce, which
  // covers the single message buffer the same way.  One cluster barrier per step remains (prenet): the stop flag, and
  // its release / acquire makes the K/V rows the peers appended to the cache in global memory visible.
  enum { BAR_R0 = 0, BAR_R1 = 1, BAR_P = 2, BAR_M = 3 };
  uint32_t par[4] = {0, 0, 0, 0};
  int red_sel = 0;                                               // landing buffer of the next all-reduce
  auto arm = [&](int d, uint32_t bytes) { if (tid == 0) mbar_expect_tx(&sm.bar[d], bytes); };
  auto arrived = [&](int d) {
    mbar_wait(&sm.bar[d], par[d]);
    par[d] ^= 1u;
    __syncthreads();
  };
  auto row_valid = [&](int r) { return row0 + r < a.B; };
  auto row_b = [&](int r) { return min(row0 + r, a.B - 1); };

  // ---- all-reduce of the partial sums staged in sm.pout: every CTA sends its [R][256] block to every peer (16-byte
  // st.async, 512 contiguous bytes per warp instruction) and then adds the C blocks that arrived in its own buffer
  auto all_reduce_send = [&]() {
    const int sel = red_sel;
    arm(sel, (uint32_t)(C * R * kD * 4));
    fence_proxy_async();                                         // sm.pout[sel] (generic stores) -> async proxy
    __syncthreads();                                             // sm.pout[sel] complete
    if (tid < C) {                                               // thread p: one bulk copy of the [R][256] block into CTA p
      const uint32_t dst = mapa_u32(smem_u32(red + ((size_t)(sel * C + rank) * R) * kD), (uint32_t)tid);
      bulk_copy_to_peer(dst, &sm.pout[sel][0][0], (uint32_t)(R * kD * 4), mapa_u32(smem_u32(&sm.bar[sel]), (uint32_t)tid));
    }
  };
  // sum of the C blocks for column n = tid of row r (call after all_reduce_wait)
  auto all_reduce_wait = [&]() -> const float* {
    const int sel = red_sel;
    arrived(sel);
    red_sel ^= 1;
    return red + (size_t)sel * C * R * kD;
  };
  auto reduced = [&](const float* buf, int r, int n) {
    float s = 0.f;
    for (int c = 0; c < C; ++c) s += buf[((size_t)c * R + r) * kD + n];
    return s;
  };
  // x = LayerNorm(x + bias + sum of the partial blocks)
  auto residual_layernorm = [&](const float* __restrict__ bias, const float* __restrict__ ln_g, const float* __restrict__ ln_b) {
    const float* buf = all_reduce_wait();
    const float bn = __ldg(bias + tid);
    float z[R];
#pragma unroll
    for (int r = 0; r < R; ++r) z[r] = sm.x[r][tid] + bn + reduced(buf, r, tid);
    layernorm_cols<R>(sm.x, z, sm.stat, ln_g, ln_b);
  };

  // ---- attention of this CTA's head over its share of the cached keys; the head group exchanges the partial softmax
  // states and, for self-attention, the pieces of the step's own key (partial score) and value; leaves the normalised
  // head output in sm.att.  `own`: the step's key / value piece is in sm.kcur / sm.vcur (self-attention).
  auto attention_phase = [&](const bf16* kv, int n_keys, bool own) {
    const int msg_f = 68 + (own ? pk : 0);                       // floats per message
    arm(BAR_P, (uint32_t)(G * R * msg_f * 4));
    float d_own = 0.f;                                           // partial score of the step's own key over this CTA's dims
    if (own && warp < R) {
      for (int i = lane; i < pk; i += 32) d_own += sm.q[warp][g * pk + i] * sm.kcur[warp][i];
      d_own = warp_sum(d_own);
    }
    Partial p = attend_keys(kv, n_keys, &sm.q[arow][0], head, unit, units);
    const int dl = lane & 7;
    if (lane < 8) {
      float* w = sm.wpart[warp];
      if (lane == 0) { w[0] = p.m; w[1] = p.l; }
#pragma unroll
      for (int i = 0; i < 8; ++i) w[4 + dl * 8 + i] = p.o[i];
    }
    __syncthreads();
    for (int i = tid; i < R * kDh; i += DEC_THREADS) {           // combine the WPR warps of a row: thread -> (row, dim)
      const int r = i >> 6, dd = i & 63;
      float m = -INFINITY;
#pragma unroll
      for (int j = 0; j < WPR; ++j) m = fmaxf(m, sm.wpart[r + R * j][0]);
      float l = 0.f, o = 0.f;
#pragma unroll
      for (int j = 0; j < WPR; ++j) {
        const float* w = sm.wpart[r + R * j];
        const float c = w[0] == -INFINITY ? 0.f : exp2f((w[0] - m) * kLog2eD);
        l += w[1] * c;
        o += w[4 + dd] * c;
      }
      float* mg = sm.msg[r];
      mg[4 + dd] = o;
      if (dd == 0) { mg[0] = m; mg[1] = l; mg[3] = 0.f; }
      if (own && dd < pk) mg[68 + dd] = sm.vcur[r][dd];
    }
    if (warp < R && lane == 0) sm.msg[warp][2] = d_own;
    fence_proxy_async();
    __syncthreads();
    if (tid < G * R) {                                           // one bulk copy per (group peer, row)
      const int pg = tid / R, r = tid - pg * R;
      const uint32_t peer = (uint32_t)(head + 4 * pg);
      bulk_copy_to_peer(mapa_u32(smem_u32(&sm.part[g][r][0]), peer), &sm.msg[r][0], (uint32_t)(msg_f * 4),
                        mapa_u32(smem_u32(&sm.bar[BAR_P]), peer));
    }
    arrived(BAR_P);
    // every CTA of the group combines the G partial states (+ the step's own key): thread -> (row, head dim)
    for (int i = tid; i < R * kDh; i += DEC_THREADS) {
      const int r = i >> 6, dd = i & 63;
      float m = -INFINITY, s_own = 0.f;
      for (int s = 0; s < G; ++s) { m = fmaxf(m, sm.part[s][r][0]); s_own += sm.part[s][r][2]; }
      if (own) m = fmaxf(m, s_own);
      float l = 0.f, o = 0.f;
      for (int s = 0; s < G; ++s) {
        const float* w = sm.part[s][r];
        const float c = w[0] == -INFINITY ? 0.f : exp2f((w[0] - m) * kLog2eD);
        l += w[1] * c;
        o += w[4 + dd] * c;
      }
      if (own) {
        const float e = exp2f((s_own - m) * kLog2eD);
        l += e;
        o += e * sm.part[dd / pk][r][68 + dd % pk];
      }
      sm.att[r][dd] = l > 0.f ? o / l : 0.f;
    }
    __syncthreads();
  };

  // ---- out-projection of an attention sub-block, K-split over this CTA's head dimensions, + all-reduce send
  auto out_projection = [&](const bf16* wt) {
    gemv_ksplit<R>(wt, (head * kDh + g * pk) >> 3, pk >> 3, &sm.att[0][g * pk], kDh, sm.pout[red_sel]);
    all_reduce_send();
  };

  int first_open = 0;          // (thread 0 of rank 0) first step whose stop verdict is still unknown
  const int pre_n = kD / C, ffn_n = kF / C;                      // column slices of prenet-1 and FFN-1

  for (int t = 0; t < a.Tmax; ++t) {
    // ---- stop rule, evaluated one step late by one thread while the others start the prenet
    if (rank == 0 && tid == 0 && t > 0) {
      const int base = first_open;
      const int n_chk = min(8, t - base);
      int vals[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) vals[i] = i < n_chk ? __ldcg(a.stop_count + base + i) : 0;
      bool stop = false;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (i >= n_chk) break;
        const int v = vals[i];
        if ((v >> 16) != a.B) break;                             // some cluster has not reported this step yet
        if ((v & 0xffff) == a.B) { stop = true; break; }
        first_open = base + i + 1;
      }
      if (stop)
        for (int p = 0; p < C; ++p) *cluster.map_shared_rank(&sm.exit_flag, p) = 1;
    }
    // ---- teacher-forced decode: the fed-back frame comes from a given trajectory (all bytes of the previous step's
    // all-gather into melprev have landed: BAR_M was waited at the end of that step)
    if (a.forced_mel != nullptr && t > 0) {
      for (int i = tid; i < R * kMel; i += DEC_THREADS) {
        const int r = i / kMel, c = i - r * kMel;
        sm.melprev[r][c] = __ldg(a.forced_mel + ((size_t)row_b(r) * a.Tmax + (t - 1)) * kMel + c);
      }
      __syncthreads();
    }
    // ---- decoder prenet (model/model.py:61-67): relu(W1 mel_prev + b1), this CTA's column slice
    gemv_rows<R>(a.w_pre1, kMelPad, a.b_pre1, kMelPad, pre_n, [&](int i) { return rank * pre_n + i; },
                 &sm.melprev[0][0], kMelPad, [&](int i, int, int r, float v) { sm.h1[r][i] = fmaxf(v, 0.f); });
    cluster.sync();
    if (sm.exit_flag) break;
    // relu(W2 . + b2) + alpha * pe[t]   (model/model.py:95, 355-356): K-split over the slice, all-reduce
    gemv_ksplit<R>(a.w_pre2, (rank * pre_n) >> 3, pre_n >> 3, &sm.h1[0][0], kD, sm.pout[red_sel]);
    all_reduce_send();
    {
      const float* buf = all_reduce_wait();
      const float c0 = __ldg(a.b_pre2 + tid), pe = a.alpha * __ldg(a.pe + (size_t)t * kD + tid);
#pragma unroll
      for (int r = 0; r < R; ++r) sm.x[r][tid] = fmaxf(reduced(buf, r, tid) + c0, 0.f) + pe;
      __syncthreads();
    }

#pragma unroll 1
    for (int l = 0; l < 3; ++l) {
      const DecLayerW& w = a.layer[l];
      // ---- self-attention: this head's query in full, this CTA's piece of the step's key / value (cache append)
      {
        bf16* cache = a.self_kv + (((size_t)l * a.B) * a.Tmax + t) * 512 + head * kDh + g * pk;
        gemv_rows<R>(w.w_qkv, kD, w.b_qkv, kD, kDh + 2 * pk,
                        [&](int i) { return i < kDh ? head * kDh + i
                                                    : (i < kDh + pk ? kD + head * kDh + g * pk + (i - kDh)
                                                                    : 2 * kD + head * kDh + g * pk + (i - kDh - pk)); },
                        &sm.x[0][0], kD, [&](int i, int, int r, float v) {
                          if (i < kDh) { sm.q[r][i] = v; return; }
                          const bf16 h = __float2bfloat16(v);    // what later steps read back from the cache
                          const bool is_v = i >= kDh + pk;
                          const int j = is_v ? i - kDh - pk : i - kDh;
                          (is_v ? sm.vcur : sm.kcur)[r][j] = __bfloat162float(h);
                          if (row_valid(r)) cache[(size_t)(row0 + r) * a.Tmax * 512 + (is_v ? kD : 0) + j] = h;
                        });
        __syncthreads();
      }
      attention_phase(a.self_kv + ((size_t)l * a.B + row_b(arow)) * a.Tmax * 512, t, true);
      out_projection(w.w_so);
      residual_layernorm(w.b_so, w.ln1_g, w.ln1_b);
      // ---- cross-attention over the precomputed memory K/V (keys >= phoneme_len masked)
      gemv_rows<R>(w.w_cq, kD, w.b_cq, kD, kDh, [&](int i) { return head * kDh + i; }, &sm.x[0][0], kD,
                   [&](int i, int, int r, float v) { sm.q[r][i] = v; });
      __syncthreads();
      attention_phase(a.cross_kv + ((size_t)l * a.B + row_b(arow)) * a.Tp * 512, min(a.ph_len[row_b(arow)], a.Tp), false);
      out_projection(w.w_co);
      residual_layernorm(w.b_co, w.ln2_g, w.ln2_b);
      // ---- position-wise FFN: hidden slice (column-parallel), second matrix K-split over the slice
      gemv_rows<R>(w.w_f1, kD, w.b_f1, kD, ffn_n, [&](int i) { return rank * ffn_n + i; }, &sm.x[0][0], kD,
                   [&](int i, int, int r, float v) { sm.h1[r][i] = fmaxf(v, 0.f); });
      __syncthreads();
      gemv_ksplit<R>(w.w_f2, (rank * ffn_n) >> 3, ffn_n >> 3, &sm.h1[0][0], kD, sm.pout[red_sel]);
      all_reduce_send();
      residual_layernorm(w.b_f2, w.ln3_g, w.ln3_b);
    }

    // ---- mel + stop heads (model/model.py:376-380), column-parallel; the frame is fed back un-rounded
    {
      const int per = (kMel + 1 + C - 1) / C;
      const int nb = rank * per, ne = min(kMel + 1, nb + per);
      arm(BAR_M, R * kMel * 4);
      int rep = 0, cnt = 0;
      const uint32_t peer_mel = mapa_u32(smem_u32(&sm.melprev[0][0]), (uint32_t)(lane % C));
      const uint32_t peer_bar = mapa_u32(smem_u32(&sm.bar[BAR_M]), (uint32_t)(lane % C));
      gemv_rows_bcast<R>(a.w_head, kD, a.b_head, kD, max(ne - nb, 0), [&](int i) { return nb + i; }, &sm.x[0][0], kD,
                      [&](int, int n, int r, float v) {
                        if (n < kMel) {
                          if (lane < C) st_async_f32(peer_mel + (r * kMelPad + n) * 4, v, peer_bar);   // lane p -> CTA p
                          if (lane == 0 && row_valid(r)) a.mel_out[((size_t)(row0 + r) * a.Tmax + t) * kMel + n] = v;
                        } else if (row_valid(r)) {
                          if (lane == 0) a.stop_out[(size_t)(row0 + r) * a.Tmax + t] = v;
                          ++rep;
                          cnt += v >= a.stop_logit ? 1 : 0;
                        }
                      });
      if (lane == 0 && rep > 0) atomicAdd(a.stop_count + t, (rep << 16) | cnt);
      arrived(BAR_M);            // waited in the same step: nothing is in flight towards a CTA that leaves the loop
    }
  }
  cluster.sync();
}

__global__ void decode_finalize(const int* stop_count, int Tmax, int B, int* n_frames) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  int n = Tmax;
  for (int t = 0; t < Tmax; ++t) {
    const int v = stop_count[t];
    if ((v >> 16) == B && (v & 0xffff) == B) { n = t + 1; break; }
  }
  *n_frames = n;
}

template <int R>
static const char* launch_r(const DecodeArgs& a, int C, cudaStream_t s) {
  const int n_clusters = (a.B + R - 1) / R;
  const size_t smem = dec_smem_bytes<R>(C);
  if (smem > 227 * 1024) return "decode: rows per cluster x cluster size needs more shared memory than an SM has";
  if (cudaFuncSetAttribute(decode_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
    return "cudaFuncSetAttribute(decode smem) failed";
  if (C > 8 &&
      cudaFuncSetAttribute(decode_kernel<R>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess)
    return "cudaFuncSetAttribute(decode non-portable cluster) failed";
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(n_clusters * C);
  cfg.blockDim = dim3(DEC_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = C; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  if (cudaLaunchKernelEx(&cfg, decode_kernel<R>, a) != cudaSuccess) return "decode kernel launch failed";
  decode_finalize<<<1, 32, 0, s>>>(a.stop_count, a.Tmax, a.B, a.n_frames);
  note_launch(2);
  return cudaGetLastError() == cudaSuccess ? nullptr : "decode finalize launch failed";
}

// How many clusters of C CTAs of decode_kernel<R> the device can hold at once (0 on error).  A launch that needs
// more than this runs in waves, i.e. the late clusters decode their utterances after the early ones finished.
template <int R>
static int max_clusters_r(int C) {
  const size_t smem = dec_smem_bytes<R>(C);
  if (smem > 227 * 1024) return 0;
  if (cudaFuncSetAttribute(decode_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 0;
  if (C > 8 && cudaFuncSetAttribute(decode_kernel<R>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess)
    return 0;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(C * 64);
  cfg.blockDim = dim3(DEC_THREADS);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = C; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, decode_kernel<R>, &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int decode_max_clusters(int R, int C) {
  static int cache_dev[kMaxDevices][9][17] = {};     // per device; benign race (same value from every thread)
  if (R < 1 || R > 8 || C < 1 || C > 16) return 0;
  DevInfo di;
  if (device_info(&di) != nullptr) return 0;
  int (*cache)[17] = cache_dev[di.dev];
  if (cache[R][C] == 0) {
    int n = 0;
    switch (R) {
      case 1: n = max_clusters_r<1>(C); break;
      case 2: n = max_clusters_r<2>(C); break;
      case 4: n = max_clusters_r<4>(C); break;
      case 8: n = max_clusters_r<8>(C); break;
      default: n = 0;
    }
    cache[R][C] = n > 0 ? n : -1;
  }
  return cache[R][C] > 0 ? cache[R][C] : 0;
}

// Rows per cluster / cluster size for a batch: everything must be co-resident (one wave); among the feasible plans
// the one with the lowest measured step time wins.  The table is the 1000-frame sweep of this kernel on B200
// (tools/decode_sweep.sh, profiles/r02_decode_plan_sweep.txt), us per step: `lo` with at most one CTA per SM, `hi` with
// two (R <= 2 only; 16-CTA clusters lose most when two CTAs share an SM: their all-reduce traffic doubles).
void decode_plan(int B, int* R_out, int* C_out) {
  static const int Rs[4] = {1, 2, 4, 8}, Cs[3] = {4, 8, 16};
  static const float lo[4][3] = {{62.f, 46.f, 44.f}, {117.f, 76.f, 74.f}, {211.f, 212.f, 179.f}, {398.f, 300.f, 1e9f}};
  static const float hi[4][3] = {{75.f, 58.f, 99.f}, {135.f, 88.f, 137.f}, {1e9f, 1e9f, 1e9f}, {1e9f, 1e9f, 1e9f}};
  DevInfo di;
  const int sms = device_info(&di) == nullptr ? di.sm_count : 148;
  float best = 1e30f;
  int bR = 8, bC = 4;
  for (int ri = 0; ri < 4; ++ri) {
    for (int ci = 0; ci < 3; ++ci) {
      const int need = (B + Rs[ri] - 1) / Rs[ri];
      const int cap = decode_max_clusters(Rs[ri], Cs[ci]);
      if (cap <= 0 || need > cap) continue;
      const float t = need * Cs[ci] <= sms ? lo[ri][ci] : hi[ri][ci];
      if (t < best) { best = t; bR = Rs[ri]; bC = Cs[ci]; }
    }
  }
  *R_out = bR; *C_out = bC;
}

// Utterances one launch can hold co-resident (the largest one-wave plan); larger batches still decode correctly, in
// waves, but the device-side stop rule cannot end the first wave early (its clusters never see the later ones report).
int decode_capacity() {
  int best = 0;
  for (int R = 1; R <= 8; R *= 2)
    for (int C = 4; C <= 16; C *= 2) {
      const int rows = decode_max_clusters(R, C) * R;
      if (rows > best) best = rows;
    }
  return best;
}

const char* decode_launch(const DecodeArgs& a, int R, int C, cudaStream_t s) {
  if (a.B <= 0 || a.Tmax <= 0 || a.Tp <= 0) return "decode: empty problem";
  if (a.B > 65535) return "decode: batch too large for the stop counter";
  if (C != 4 && C != 8 && C != 16) return "decode: cluster size must be 4, 8 or 16";
  switch (R) {
    case 1: return launch_r<1>(a, C, s);
    case 2: return launch_r<2>(a, C, s);
    case 4: return launch_r<4>(a, C, s);
    case 8: return launch_r<8>(a, C, s);
    default: return "decode: rows per cluster must be 1, 2, 4 or 8";
  }
}

}  // namespace tts
