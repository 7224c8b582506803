// Persistent autoregressive greedy mel decode with a per-layer K/V cache (sm_100a).
//
// Replaces the loop of TransformerTTS.inference (model/model.py:354-384).  The reference re-runs
// prenet + PE + all decoder layers over the whole prefix every step and syncs with the host three
// times per step; here ONE kernel launch runs all steps.  A thread-block cluster owns R utterances:
// every CTA keeps the R x 256 fp32 residual stream in shared memory, computes a 1/C slice of each
// projection (bf16 weights streamed from L2, warp-shuffle reductions), and publishes the slice into the
// shared memory of all C CTAs of the cluster (DSMEM) with st.async stores that complete bytes on an mbarrier of
// the destination CTA: a phase boundary is a local wait for the bytes of the phase (no release fence, no cluster
// rendezvous); one barrier.cluster per step remains (stop flag, K/V cache visibility).  Attention over the
// cache is split over heads x key ranges across the CTAs with 16-byte coalesced K/V reads and an online
// softmax; partial (max, sum, o) triples are combined once they have arrived.  The stop rule ("all rows >= threshold at the same step", model/model.py:382-384) is evaluated
// on the device through a per-step counter; no host round trip happens inside the loop.
#include <cooperative_groups.h>

#include "decode.cuh"
#include "ptx.cuh"

namespace cg = cooperative_groups;

namespace tts {

constexpr int DEC_THREADS = 256;
constexpr int DEC_WARPS = DEC_THREADS / 32;
constexpr int LDA = 1024;          // floats per row of the broadcast buffers
constexpr int PSTRIDE = 68;        // floats per attention partial: m, l, pad, pad, o[64]
constexpr int MAX_CLUSTER = 16;
constexpr float kLog2eD = 1.4426950408889634f;

template <int R>
struct DecSmem {
  float x[R][kD];                  // residual stream (replicated in every CTA)
  float att[R][kD];                // combined attention output (local)
  float melprev[R][kMelPad];       // previous frame, zero padded to 128 (broadcast target)
  float bufA[R][LDA];              // broadcast targets, ping-pong
  float bufB[R][LDA];
  float part[MAX_CLUSTER / 4][kH][R][PSTRIDE];   // attention partials [split][head][row]
  float wpart[DEC_WARPS][PSTRIDE]; // per-warp partials before the in-CTA combine
  int exit_flag;
  alignas(8) uint64_t bar[4];      // data-arrival barriers of bufA / bufB / part / melprev (st.async complete_tx from all peers)
};

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

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y), c = unpack_bf16(u.z), d = unpack_bf16(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

// y[r][n] = sum_k in[r][k] * W[n][k] + bias[n] for n in [n_begin, n_end): one warp per output,
// U outputs in flight per warp, lanes stride K in 16-byte weight loads.  emit(n0, r, values[U]) is called
// by all 32 lanes with the reduced values of outputs n0 .. n0+U-1 of row r.
template <int R, int U, class Emit>
__device__ __forceinline__ void gemv_slice(const bf16* __restrict__ W, int ldw, const float* __restrict__ bias, int K,
                                           int n_begin, int n_end, const float* in, int ld_in, Emit&& emit) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int nb = n_begin + warp * U; nb < n_end; nb += DEC_WARPS * U) {
    float acc[U][R];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int r = 0; r < R; ++r) acc[u][r] = 0.f;
    for (int k0 = lane * 8; k0 < K; k0 += 256) {
      uint4 w[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int n = nb + u;
        w[u] = n < n_end ? __ldg(reinterpret_cast<const uint4*>(W + (size_t)n * ldw + k0)) : make_uint4(0, 0, 0, 0);
      }
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
          for (int i = 0; i < 8; ++i) acc[u][r] = fmaf(wf[i], xin[r][i], acc[u][r]);
      }
    }
    // the slices are multiples of U wide (N / C for N = 256, 768, 1024 and U = 1 for the 81 head rows), so a pass is
    // always full: emit gets the U adjacent outputs of a row at once and can publish them as one vector store
    float bn[U];
#pragma unroll
    for (int u = 0; u < U; ++u) bn[u] = __ldg(bias + nb + u);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float vals[U];
#pragma unroll
      for (int u = 0; u < U; ++u) vals[u] = warp_sum(acc[u][r]) + bn[u];
      emit(nb, r, vals);
    }
  }
}

// x[r] = LayerNorm(x[r] + delta[r]) for the R rows, one warp per row (eps 1e-5, biased variance).
template <int R>
__device__ __forceinline__ void add_layernorm(float (*x)[kD], const float (*delta)[LDA], const float* __restrict__ g,
                                              const float* __restrict__ b) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r = warp; r < R; r += DEC_WARPS) {
    float v[8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { v[i] = x[r][lane + 32 * i] + delta[r][lane + 32 * i]; s += v[i]; }
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
//   q         : 64 fp32 query values (already scaled)
//   unit/units: this warp's index / count among the warps sharing the (row, head) key range
//   cur_k/v   : optional extra key held in shared memory (the step's own k, v; fp32, rounded to bf16 here
//               so that it matches what later steps read back from the cache)
__device__ __forceinline__ Partial attend_keys(const bf16* kv, int n_keys, const float* q, int head, int unit,
                                               int units, const float* cur_k, const float* cur_v) {
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
  for (int g4 = unit; g4 < n_groups; g4 += 2 * units) {
    // two independent 4-key groups per iteration keep four 16-byte loads in flight per lane
    const int j0 = g4 * 4 + grp, j1 = (g4 + units) * 4 + grp;
    const bool ok0 = j0 < n_keys, ok1 = j1 < n_keys;
    uint4 k0 = make_uint4(0, 0, 0, 0), v0 = k0, k1 = k0, v1 = k0;
    if (ok0) { k0 = __ldcg(reinterpret_cast<const uint4*>(base + (size_t)j0 * 512));
               v0 = __ldcg(reinterpret_cast<const uint4*>(base + (size_t)j0 * 512 + 256)); }
    if (ok1) { k1 = __ldcg(reinterpret_cast<const uint4*>(base + (size_t)j1 * 512));
               v1 = __ldcg(reinterpret_cast<const uint4*>(base + (size_t)j1 * 512 + 256)); }
    float kf[8], vf[8];
    unpack8(k0, kf);
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s = fmaf(qf[i], kf[i], s);
    s += __shfl_xor_sync(0xffffffffu, s, 1); s += __shfl_xor_sync(0xffffffffu, s, 2); s += __shfl_xor_sync(0xffffffffu, s, 4);
    unpack8(v0, vf);
    if (ok0) partial_update(p, s, vf);
    unpack8(k1, kf);
    s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s = fmaf(qf[i], kf[i], s);
    s += __shfl_xor_sync(0xffffffffu, s, 1); s += __shfl_xor_sync(0xffffffffu, s, 2); s += __shfl_xor_sync(0xffffffffu, s, 4);
    unpack8(v1, vf);
    if (ok1) partial_update(p, s, vf);
  }
  if (cur_k != nullptr && unit == 0) {
    float kf[8], vf[8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      kf[i] = __bfloat162float(__float2bfloat16(cur_k[head * kDh + dl * 8 + i]));
      vf[i] = __bfloat162float(__float2bfloat16(cur_v[head * kDh + dl * 8 + i]));
      s = fmaf(qf[i], kf[i], s);
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1); s += __shfl_xor_sync(0xffffffffu, s, 2); s += __shfl_xor_sync(0xffffffffu, s, 4);
    if (grp == 0) partial_update(p, s, vf);
  }
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
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row0 = (blockIdx.x / C) * R;
  DecSmem<R>* peer = cluster.map_shared_rank(&sm, lane % C);     // lane p publishes into CTA p

  // attention work split: CTA -> (head, key split); warp -> (row, sub-unit)
  const int head = rank & 3, split = rank >> 2, n_split = C >> 2;
  constexpr int WPR = DEC_WARPS / R;                             // warps per row
  const int arow = warp % R, wi = warp / R;
  const int unit = split * WPR + wi, units = n_split * WPR;

  for (int i = tid; i < R * kMelPad; i += DEC_THREADS) (&sm.melprev[0][0])[i] = 0.f;
  if (tid == 0) {
    sm.exit_flag = 0;
    for (int i = 0; i < 4; ++i) mbar_init(&sm.bar[i], 1);
    fence_mbar_init();
  }
  cluster.sync();

  // Phase protocol.  Most phases publish with st.async into every peer and wait on the local barrier of the
  // destination buffer: thread 0 arms it with the bytes the whole cluster will deliver, everybody waits for the
  // phase to complete.  A CTA can only reach phase p after every peer has sent phase p-1 (all its warps emit in
  // every such phase), i.e. after every peer has finished reading what phase p overwrites.  One phase per step keeps
  // the full cluster barrier: prenet 1 (stop flag; its release / acquire also makes the K/V rows that the peers
  // appended to the cache in global memory during the previous step visible before this step's attention).
  enum { BAR_A = 0, BAR_B = 1, BAR_P = 2, BAR_M = 3 };
  const uint32_t sm_base = smem_u32(&sm);
  const uint32_t peer_base = mapa_u32(sm_base, (uint32_t)(lane % C));        // DecSmem of CTA lane % C
  auto raddr = [&](const void* local) { return peer_base + (smem_u32(local) - sm_base); };
  const uint32_t rbarA = raddr(&sm.bar[BAR_A]), rbarB = raddr(&sm.bar[BAR_B]), rbarM = raddr(&sm.bar[BAR_M]);
  uint32_t parA = 0, parB = 0, parP = 0, parM = 0;
  auto arm = [&](int d, uint32_t bytes) { if (tid == 0) mbar_expect_tx(&sm.bar[d], bytes); };
  auto arrived = [&](int d, uint32_t& parity) {
    mbar_wait(&sm.bar[d], parity);
    parity ^= 1u;
    __syncthreads();
  };

  auto slice = [&](int N, int& nb, int& ne) {
    const int per = (N + C - 1) / C;
    nb = rank * per;
    ne = min(N, nb + per);
  };
  auto row_valid = [&](int r) { return row0 + r < a.B; };
  auto row_b = [&](int r) { return min(row0 + r, a.B - 1); };

  // ---- attention over cached keys for (arow, head); publishes the CTA's partial to every peer
  auto attention_phase = [&](const bf16* kv, int n_keys, const float* q, const float* cur_k, const float* cur_v) {
    arm(BAR_P, (uint32_t)(C * R * 66 * 4));                      // n_split x 4 heads x R rows x (m, l, o[64])
    Partial p = attend_keys(kv, n_keys, q, head, unit, units, cur_k, cur_v);
    const int dl = lane & 7;
    if (lane < 8) {
      float* w = sm.wpart[warp];
      if (lane == 0) { w[0] = p.m; w[1] = p.l; }
#pragma unroll
      for (int i = 0; i < 8; ++i) w[4 + dl * 8 + i] = p.o[i];
    }
    __syncthreads();
    if (warp < R) {                                              // wi == 0: combine the WPR warps of row `warp`
      float m = -INFINITY;
#pragma unroll
      for (int i = 0; i < WPR; ++i) m = fmaxf(m, sm.wpart[warp + R * i][0]);
      float l = 0.f, o[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) o[i] = 0.f;
#pragma unroll
      for (int i = 0; i < WPR; ++i) {
        const float* w = sm.wpart[warp + R * i];
        const float c = w[0] == -INFINITY ? 0.f : exp2f((w[0] - m) * kLog2eD);
        l += w[1] * c;
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] += w[4 + dl * 8 + j] * c;
      }
      const uint32_t loc = smem_u32(&sm.part[split][head][warp][0]);
      for (int p0 = 0; p0 < C; p0 += 4) {
        const int pr = p0 + (lane >> 3);
        if (pr < C) {
          const uint32_t dst = mapa_u32(loc, (uint32_t)pr), rb = mapa_u32(smem_u32(&sm.bar[BAR_P]), (uint32_t)pr);
          if (dl == 0) st_async_f32x2(dst, m, l, rb);
          st_async_f32x4(dst + (4 + dl * 8) * 4, o[0], o[1], o[2], o[3], rb);
          st_async_f32x4(dst + (8 + dl * 8) * 4, o[4], o[5], o[6], o[7], rb);
        }
      }
    }
    arrived(BAR_P, parP);
    // every CTA combines the key splits of all heads: thread -> (head, dim)
    {
      const int hh = tid >> 6, d = tid & 63;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        float m = -INFINITY;
        for (int s = 0; s < n_split; ++s) m = fmaxf(m, sm.part[s][hh][r][0]);
        float l = 0.f, o = 0.f;
        for (int s = 0; s < n_split; ++s) {
          const float* w = sm.part[s][hh][r];
          const float c = w[0] == -INFINITY ? 0.f : exp2f((w[0] - m) * kLog2eD);
          l += w[1] * c;
          o += w[4 + d] * c;
        }
        sm.att[r][tid] = l > 0.f ? o / l : 0.f;
      }
    }
    __syncthreads();
  };

  int first_open = 0;          // (thread 0 of rank 0) first step whose stop verdict is still unknown

  for (int t = 0; t < a.Tmax; ++t) {
    int nb, ne;
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
    // broadcast into melprev have landed: BAR_M was waited at the end of that step)
    if (a.forced_mel != nullptr && t > 0) {
      for (int i = tid; i < R * kMel; i += DEC_THREADS) {
        const int r = i / kMel, c = i - r * kMel;
        sm.melprev[r][c] = __ldg(a.forced_mel + ((size_t)row_b(r) * a.Tmax + (t - 1)) * kMel + c);
      }
      __syncthreads();
    }
    // ---- decoder prenet (model/model.py:61-67): relu(W1 mel_prev + b1)
    slice(kD, nb, ne);
    gemv_slice<R, 2>(a.w_pre1, kMelPad, a.b_pre1, kMelPad, nb, ne, &sm.melprev[0][0], kMelPad,
                     [&](int n, int r, const float (&v)[2]) {
                       if (lane < C)
                         *reinterpret_cast<float2*>(&peer->bufA[r][n]) = make_float2(fmaxf(v[0], 0.f), fmaxf(v[1], 0.f));
                     });
    cluster.sync();
    if (sm.exit_flag) break;
    // relu(W2 . + b2) + alpha * pe[t]   (model/model.py:95, 355-356)
    arm(BAR_B, R * kD * 4);
    gemv_slice<R, 2>(a.w_pre2, kD, a.b_pre2, kD, nb, ne, &sm.bufA[0][0], LDA, [&](int n, int r, const float (&v)[2]) {
      const float2 pe2 = __ldg(reinterpret_cast<const float2*>(a.pe + (size_t)t * kD + n));
      if (lane < C)
        st_async_f32x2(raddr(&sm.bufB[r][n]), fmaxf(v[0], 0.f) + a.alpha * pe2.x, fmaxf(v[1], 0.f) + a.alpha * pe2.y, rbarB);
    });
    arrived(BAR_B, parB);
    for (int i = tid; i < R * kD; i += DEC_THREADS) sm.x[i / kD][i % kD] = sm.bufB[i / kD][i % kD];
    __syncthreads();

#pragma unroll 1
    for (int l = 0; l < 3; ++l) {
      const DecLayerW& w = a.layer[l];
      // ---- self-attention: in-proj, cache append
      slice(3 * kD, nb, ne);
      arm(BAR_A, R * 3 * kD * 4);
      gemv_slice<R, 4>(w.w_qkv, kD, w.b_qkv, kD, nb, ne, &sm.x[0][0], kD, [&](int n, int r, const float (&v)[4]) {
        if (lane < C) st_async_vec<4>(raddr(&sm.bufA[r][n]), v, rbarA);
        if (lane == 0 && n >= kD && row_valid(r))
          *reinterpret_cast<uint2*>(a.self_kv + (((size_t)l * a.B + row0 + r) * a.Tmax + t) * 512 + (n - kD)) =
              make_uint2(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]));
      });
      arrived(BAR_A, parA);
      attention_phase(a.self_kv + ((size_t)l * a.B + row_b(arow)) * a.Tmax * 512, t, &sm.bufA[arow][head * kDh],
                      &sm.bufA[arow][kD], &sm.bufA[arow][2 * kD]);
      slice(kD, nb, ne);
      arm(BAR_B, R * kD * 4);
      gemv_slice<R, 2>(w.w_so, kD, w.b_so, kD, nb, ne, &sm.att[0][0], kD,
                       [&](int n, int r, const float (&v)[2]) { if (lane < C) st_async_vec<2>(raddr(&sm.bufB[r][n]), v, rbarB); });
      arrived(BAR_B, parB);
      add_layernorm<R>(sm.x, sm.bufB, w.ln1_g, w.ln1_b);
      // ---- cross-attention over the precomputed memory K/V (keys >= phoneme_len masked)
      arm(BAR_A, R * kD * 4);
      gemv_slice<R, 2>(w.w_cq, kD, w.b_cq, kD, nb, ne, &sm.x[0][0], kD,
                       [&](int n, int r, const float (&v)[2]) { if (lane < C) st_async_vec<2>(raddr(&sm.bufA[r][n]), v, rbarA); });
      arrived(BAR_A, parA);
      attention_phase(a.cross_kv + ((size_t)l * a.B + row_b(arow)) * a.Tp * 512, min(a.ph_len[row_b(arow)], a.Tp),
                      &sm.bufA[arow][head * kDh], nullptr, nullptr);
      arm(BAR_B, R * kD * 4);
      gemv_slice<R, 2>(w.w_co, kD, w.b_co, kD, nb, ne, &sm.att[0][0], kD,
                       [&](int n, int r, const float (&v)[2]) { if (lane < C) st_async_vec<2>(raddr(&sm.bufB[r][n]), v, rbarB); });
      arrived(BAR_B, parB);
      add_layernorm<R>(sm.x, sm.bufB, w.ln2_g, w.ln2_b);
      // ---- position-wise FFN
      slice(kF, nb, ne);
      arm(BAR_A, R * kF * 4);
      gemv_slice<R, 4>(w.w_f1, kD, w.b_f1, kD, nb, ne, &sm.x[0][0], kD, [&](int n, int r, const float (&v)[4]) {
        if (lane < C)
          st_async_f32x4(raddr(&sm.bufA[r][n]), fmaxf(v[0], 0.f), fmaxf(v[1], 0.f), fmaxf(v[2], 0.f), fmaxf(v[3], 0.f), rbarA);
      });
      arrived(BAR_A, parA);
      slice(kD, nb, ne);
      arm(BAR_B, R * kD * 4);
      gemv_slice<R, 2>(w.w_f2, kF, w.b_f2, kF, nb, ne, &sm.bufA[0][0], LDA,
                       [&](int n, int r, const float (&v)[2]) { if (lane < C) st_async_vec<2>(raddr(&sm.bufB[r][n]), v, rbarB); });
      arrived(BAR_B, parB);
      add_layernorm<R>(sm.x, sm.bufB, w.ln3_g, w.ln3_b);
    }

    // ---- mel + stop heads (model/model.py:376-380); the frame is fed back un-rounded
    slice(kMel + 1, nb, ne);
    arm(BAR_M, R * kMel * 4);
    int rep = 0, cnt = 0;
    gemv_slice<R, 1>(a.w_head, kD, a.b_head, kD, nb, ne, &sm.x[0][0], kD, [&](int n, int r, const float (&vv)[1]) {
      const float v = vv[0];
      if (n < kMel) {
        if (lane < C) st_async_f32(raddr(&sm.melprev[r][n]), v, rbarM);
        if (lane == 0 && row_valid(r)) a.mel_out[((size_t)(row0 + r) * a.Tmax + t) * kMel + n] = v;
      } else if (row_valid(r)) {
        if (lane == 0) a.stop_out[(size_t)(row0 + r) * a.Tmax + t] = v;
        ++rep;
        cnt += v >= a.stop_logit ? 1 : 0;
      }
    });
    if (lane == 0 && rep > 0) atomicAdd(a.stop_count + t, (rep << 16) | cnt);
    arrived(BAR_M, parM);        // waited in the same step: nothing is in flight towards a CTA that leaves the loop
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
  const size_t smem = sizeof(DecSmem<R>);
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
  const size_t smem = sizeof(DecSmem<R>);
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
// the cheapest by the measured per-step cost model  t(C, R) = base[C] + slope[C] * (R - 1)  wins
// (B200: 16-CTA clusters have the lowest latency at one row, small clusters scale better with rows; profiles/).
void decode_plan(int B, int* R_out, int* C_out) {
  const int Cs[3] = {16, 8, 4};
  const float base[3] = {58.f, 74.f, 100.f}, slope[3] = {58.f, 14.f, 24.f};
  float best = 1e30f;
  int bR = 8, bC = 4;
  for (int ci = 0; ci < 3; ++ci) {
    for (int R = 1; R <= 8; R *= 2) {
      const int need = (B + R - 1) / R;
      const int cap = decode_max_clusters(R, Cs[ci]);
      if (cap <= 0 || need > cap) continue;
      const float t = base[ci] + slope[ci] * (R - 1);
      if (t < best) { best = t; bR = R; bC = Cs[ci]; }
      break;                                           // larger R with the same C only costs more
    }
  }
  *R_out = bR; *C_out = bC;
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
