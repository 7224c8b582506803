// tcgen05 / TMEM / TMA dense contraction with fused epilogues (sm_100a).
//
// One CTA computes one 128 x BN output tile:  warp 0 = TMA producer, warp 1 = TMEM owner + single-thread
// tcgen05.mma issuer, warps 2..5 = epilogue (TMEM -> registers -> global).  Operands arrive through a
// ring of STAGES shared-memory stages (A 128x64 bf16, W BNx64 bf16, both 128B-swizzled K-major) guarded
// by full/empty mbarriers; the fp32 accumulator lives in BN TMEM columns.  Shared memory is sized so two
// CTAs are resident per SM: one CTA's epilogue overlaps the other's mainloop.
//
// Replaces (reference call sites): nn.Linear / LinearNorm (model/module.py:36-53), Conv1d+BatchNorm1d of
// ConvNormBN (model/module.py:18-19,28-33; BN folded on the host), MultiheadAttention in/out projections
// (torch/nn/functional.py:5798,6665), FFN linear1/linear2 and the post-norm residual+LayerNorm
// (torch/nn/modules/transformer.py:951-982), PositionalEncoding add (model/model.py:95).
#include <stdlib.h>

#include "kernels.h"
#include "ptx.cuh"

namespace tts {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int GEMM_THREADS = 192;

struct GemmDev {
  int B, T, N;
  int taps, tap_off, kc_per_tap;
  int tiles_per_utt, n_tiles, num_kb;
  const float* bias;
  int act;
  const float* pe; float alpha; int pe_period;
  const bf16* residual; int ld_res;
  const float* residual_f32; int ld_res32;
  const float* ln_g; const float* ln_b;
  bf16* out_bf16; int ld_bf16;
  float* out_f32; int ld_f32;
  float* out_last;
  int vec16, vec32;                 // row pitches / bases allow 16-byte stores
};

template <int BN, int ST>
struct GemmSmem {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFF = ST * STAGE_BYTES;
  static constexpr int PAR_OFF = BAR_OFF + 256;       // bias | ln_g | ln_b, BN floats each
  static constexpr int TOTAL = PAR_OFF + 3 * BN * 4 + 1024;  // + manual 1024 B alignment slack
};

// Loads accumulator chunk `c` (32 columns) of this thread's row and applies bias/act/PE/residual.
template <int BN>
__device__ __forceinline__ void epi_chunk(const GemmDev& p, const float* s_bias, uint32_t taddr, int c, int n0,
                                          bool valid, size_t grow, int t_pe, float (&v)[32]) {
  uint32_t r[32];
  tmem_ld32(taddr + c * 32, r);
  tmem_ld_wait();
  const int col0 = n0 + c * 32;
  const float4* sb = reinterpret_cast<const float4*>(s_bias + c * 32);   // zero beyond N (W rows there are zero too)
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
    for (int j = 0; j < 32; ++j) v[j] = tanhf(v[j]);
  }
  if (!valid) return;
  if (p.pe != nullptr) {
    const float4* pr = reinterpret_cast<const float4*>(p.pe + (size_t)t_pe * kD + col0);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (col0 + j * 4 < p.N) {
        float4 q = __ldg(pr + j);
        v[j * 4 + 0] += p.alpha * q.x; v[j * 4 + 1] += p.alpha * q.y;
        v[j * 4 + 2] += p.alpha * q.z; v[j * 4 + 3] += p.alpha * q.w;
      }
    }
  }
  if (p.residual != nullptr) {
    const uint4* rr = reinterpret_cast<const uint4*>(p.residual + grow * p.ld_res + col0);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (col0 + j * 8 < p.N) {
        uint4 u = __ldg(rr + j);
        float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y), cc = unpack_bf16(u.z), d = unpack_bf16(u.w);
        v[j * 8 + 0] += a.x; v[j * 8 + 1] += a.y; v[j * 8 + 2] += b.x; v[j * 8 + 3] += b.y;
        v[j * 8 + 4] += cc.x; v[j * 8 + 5] += cc.y; v[j * 8 + 6] += d.x; v[j * 8 + 7] += d.y;
      }
    }
  }
  if (p.residual_f32 != nullptr) {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (col0 + j < p.N) v[j] += __ldg(p.residual_f32 + grow * p.ld_res32 + col0 + j);
  }
}

__device__ __forceinline__ void epi_store(const GemmDev& p, int n0, int c, size_t grow, const float (&v)[32]) {
  const int col0 = n0 + c * 32;
  const int n_main = p.out_last != nullptr ? p.N - 1 : p.N;
  if (p.out_last != nullptr) {
    const int j = p.N - 1 - col0;
    if (j >= 0 && j < 32) {
#pragma unroll
      for (int jj = 0; jj < 32; ++jj) if (jj == j) p.out_last[grow] = v[jj];
    }
  }
  if (p.out_bf16 != nullptr) {
    bf16* o = p.out_bf16 + grow * p.ld_bf16 + col0;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const int cg = col0 + g * 8;
      if (p.vec16 && cg + 8 <= n_main) {
        uint4 u;
        u.x = pack_bf16(v[g * 8 + 0], v[g * 8 + 1]); u.y = pack_bf16(v[g * 8 + 2], v[g * 8 + 3]);
        u.z = pack_bf16(v[g * 8 + 4], v[g * 8 + 5]); u.w = pack_bf16(v[g * 8 + 6], v[g * 8 + 7]);
        *reinterpret_cast<uint4*>(o + g * 8) = u;
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (cg + j < n_main) o[g * 8 + j] = __float2bfloat16(v[g * 8 + j]);
      }
    }
  }
  if (p.out_f32 != nullptr) {
    float* o = p.out_f32 + grow * p.ld_f32 + col0;
#pragma unroll
    for (int g = 0; g < 8; ++g) {
      const int cg = col0 + g * 4;
      if (p.vec32 && cg + 4 <= n_main) {
        *reinterpret_cast<float4*>(o + g * 4) = make_float4(v[g * 4], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (cg + j < n_main) o[g * 4 + j] = v[g * 4 + j];
      }
    }
  }
}

template <int BN, int ST>
__global__ void __launch_bounds__(GEMM_THREADS)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const GemmDev p) {
  using S = GemmSmem<BN, ST>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1024(smem_raw);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + S::BAR_OFF);
  uint64_t* empty = full + ST;
  uint64_t* accum_full = empty + ST;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_full + 1);
  float* s_bias = reinterpret_cast<float*>(smem + S::PAR_OFF);
  float* s_lng = s_bias + BN;
  float* s_lnb = s_lng + BN;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tile = blockIdx.x;
  const int n_tile = tile % p.n_tiles;
  const int m_tile = tile / p.n_tiles;
  const int b = m_tile / p.tiles_per_utt;
  const int t0 = (m_tile % p.tiles_per_utt) * BM;
  const int n0 = n_tile * BN;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
    for (int s = 0; s < ST; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
    mbar_init(accum_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int s = 0; uint32_t ph = 0;
      for (int kb = 0; kb < p.num_kb; ++kb) {
        const int tap = kb / p.kc_per_tap;
        const int kc = kb - tap * p.kc_per_tap;
        mbar_wait(empty + s, ph ^ 1);
        mbar_expect_tx(full + s, S::STAGE_BYTES);
        uint8_t* sa = smem + s * S::STAGE_BYTES;
        tma_load_3d(sa, &tmA, full + s, kc * BK, t0 + tap + p.tap_off, b);
        tma_load_2d(sa + S::A_BYTES, &tmW, full + s, kb * BK, n0);
        if (++s == ST) { s = 0; ph ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (one thread)
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
      int s = 0; uint32_t ph = 0;
      for (int kb = 0; kb < p.num_kb; ++kb) {
        mbar_wait(full + s, ph);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + s * S::STAGE_BYTES);
        const uint64_t da = umma_desc_sw128(sa);
        const uint64_t db = umma_desc_sw128(sa + S::A_BYTES);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k)
          tc_mma_bf16(tmem_base, da + 2 * k, db + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
        tc_commit(empty + s);
        if (kb == p.num_kb - 1) tc_commit(accum_full);
        if (++s == ST) { s = 0; ph ^= 1; }
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ epilogue: thread == output row
    const int q = warp & 3;                       // TMEM lane quadrant this warp may touch
    const int row = q * 32 + lane;
    const int t = t0 + row;
    const bool valid = t < p.T;
    const size_t grow = (size_t)b * p.T + (valid ? t : 0);
    const int t_pe = p.pe_period > 0 ? (int)(grow % (size_t)p.pe_period) : 0;
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    // stage the per-column epilogue vectors once per CTA while the mainloop runs
    for (int i = threadIdx.x - 64; i < BN; i += 128) {
      const int col = n0 + i;
      const bool in = col < p.N;
      s_bias[i] = in ? __ldg(p.bias + col) : 0.f;
      if (p.ln_g != nullptr) { s_lng[i] = in ? __ldg(p.ln_g + col) : 0.f; s_lnb[i] = in ? __ldg(p.ln_b + col) : 0.f; }
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    mbar_wait(accum_full, 0);
    tc_fence_after();
    float v[32];
    if (p.ln_g != nullptr) {
      float sum = 0.f, sq = 0.f;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        epi_chunk<BN>(p, s_bias, taddr, c, n0, valid, grow, t_pe, v);
#pragma unroll
        for (int j = 0; j < 32; ++j) { sum += v[j]; sq += v[j] * v[j]; }
      }
      const float mean = sum * (1.f / BN);
      const float var = fmaxf(sq * (1.f / BN) - mean * mean, 0.f);
      const float rstd = rsqrtf(var + 1e-5f);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        epi_chunk<BN>(p, s_bias, taddr, c, n0, valid, grow, t_pe, v);
#pragma unroll
        for (int j = 0; j < 32; ++j)
          v[j] = (v[j] - mean) * rstd * s_lng[c * 32 + j] + s_lnb[c * 32 + j];
        if (valid) epi_store(p, n0, c, grow, v);
      }
    } else {
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        if (n0 + c * 32 >= p.N) break;            // warp-uniform
        epi_chunk<BN>(p, s_bias, taddr, c, n0, valid, grow, t_pe, v);
        if (valid) epi_store(p, n0, c, grow, v);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, BN);
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  }
  return fn;
}

const char* make_tmap_3d(CUtensorMap* m, const void* base, uint64_t d0, uint64_t d1, uint64_t d2,
                         uint64_t stride1_bytes, uint64_t stride2_bytes, uint32_t box0, uint32_t box1, int swizzle_bytes,
                         int elem_bytes) {
  EncodeTiledFn enc = get_encode();
  if (enc == nullptr) return "cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)";
  cuuint64_t dims[3] = {d0, d1, d2};
  cuuint64_t strides[2] = {stride1_bytes, stride2_bytes};
  cuuint32_t box[3] = {box0, box1, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(m, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3,
                   const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? nullptr : "cuTensorMapEncodeTiled failed (alignment / stride constraints)";
}

static const char* make_tmap_2d(CUtensorMap* m, const void* base, uint64_t d0, uint64_t d1, uint64_t stride1_bytes,
                                uint32_t box0, uint32_t box1) {
  EncodeTiledFn enc = get_encode();
  if (enc == nullptr) return "cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)";
  cuuint64_t dims[2] = {d0, d1};
  cuuint64_t strides[1] = {stride1_bytes};
  cuuint32_t box[2] = {box0, box1};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? nullptr : "cuTensorMapEncodeTiled failed (alignment / stride constraints)";
}

template <int BN, int ST>
static const char* launch_t(const CUtensorMap& tmA, const CUtensorMap& tmW, const GemmDev& p, int grid,
                            cudaStream_t s) {
  using S = GemmSmem<BN, ST>;
  static bool attr_done = false;
  if (!attr_done) {
    if (cudaFuncSetAttribute(gemm_kernel<BN, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL) !=
        cudaSuccess)
      return "cudaFuncSetAttribute(gemm) failed";
    attr_done = true;
  }
  gemm_kernel<BN, ST><<<grid, GEMM_THREADS, S::TOTAL, s>>>(tmA, tmW, p);
  note_launch();
  return cudaGetLastError() == cudaSuccess ? nullptr : "gemm kernel launch failed";
}

const char* gemm_launch(const GemmArgs& a, cudaStream_t s) {
  if (a.B <= 0 || a.T <= 0 || a.N <= 0) return "gemm: empty problem";
  if ((a.C * 2) % 16 != 0) return "gemm: activation row pitch must be a multiple of 16 bytes";
  const bool ln = a.ln_g != nullptr;
  if (ln && a.N != 256) return "gemm: LayerNorm epilogue needs N == 256";
  if (a.pe != nullptr && (a.N != kD || a.pe_period <= 0)) return "gemm: PE epilogue needs N == 256 and a period";
  {
    static const bool force_v1 = getenv("TTS_GEMM_V1") != nullptr;
    if (!force_v1 && gemm2_supports(a)) return gemm2_launch(a, s);
  }
  const int ktot = a.taps * a.kc_per_tap * BK;
  GemmDev p;
  p.B = a.B; p.T = a.T; p.N = a.N;
  p.taps = a.taps; p.tap_off = a.tap_off; p.kc_per_tap = a.kc_per_tap;
  p.tiles_per_utt = (a.T + BM - 1) / BM;
  p.num_kb = a.taps * a.kc_per_tap;
  p.bias = a.bias; p.act = a.act;
  p.pe = a.pe; p.alpha = a.alpha; p.pe_period = a.pe != nullptr ? a.pe_period : 0;
  p.residual = a.residual; p.ld_res = a.ld_res;
  p.residual_f32 = a.residual_f32; p.ld_res32 = a.ld_res32;
  p.ln_g = a.ln_g; p.ln_b = a.ln_b;
  p.out_bf16 = a.out_bf16; p.ld_bf16 = a.ld_bf16;
  p.out_f32 = a.out_f32; p.ld_f32 = a.ld_f32;
  p.out_last = a.out_last;
  p.vec16 = a.out_bf16 != nullptr && a.ld_bf16 % 8 == 0 && (reinterpret_cast<uintptr_t>(a.out_bf16) & 15) == 0;
  p.vec32 = a.out_f32 != nullptr && a.ld_f32 % 4 == 0 && (reinterpret_cast<uintptr_t>(a.out_f32) & 15) == 0;
  if (a.residual != nullptr && (a.ld_res % 8 != 0 || (reinterpret_cast<uintptr_t>(a.residual) & 15) != 0))
    return "gemm: bf16 residual must be 16-byte aligned with a row pitch that is a multiple of 8";
  if ((reinterpret_cast<uintptr_t>(a.A) & 15) != 0 || (reinterpret_cast<uintptr_t>(a.W) & 15) != 0)
    return "gemm: operands must be 16-byte aligned";
  const bool wide = ln || (a.N % 256 == 0);
  const int BN = wide ? 256 : 128;
  p.n_tiles = (a.N + BN - 1) / BN;
  CUtensorMap tmA, tmW;
  const char* e = make_tmap_3d(&tmA, a.A, (uint64_t)a.C, (uint64_t)a.T, (uint64_t)a.B, (uint64_t)a.C * 2,
                               (uint64_t)a.T * a.C * 2, BK, BM);
  if (e) return e;
  e = make_tmap_2d(&tmW, a.W, (uint64_t)ktot, (uint64_t)a.N, (uint64_t)ktot * 2, BK, BN);
  if (e) return e;
  const int grid = a.B * p.tiles_per_utt * p.n_tiles;
  return wide ? launch_t<256, 2>(tmA, tmW, p, grid, s) : launch_t<128, 3>(tmA, tmW, p, grid, s);
}

}  // namespace tts
