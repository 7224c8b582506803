// Host-side plumbing shared by the kernel translation units: the tensor-map encoder (driver entry point fetched
// through the runtime, no libcuda link dependency), the per-device cache of SM counts / function-attribute opt-ins,
// and the two dispatchers c_api.cu calls (gemm_launch, attn_launch).  There is ONE kernel family per operator: a
// shape the sm_100a kernels do not take is an error, never a fallback.
#include <mutex>

#include "kernels.h"

namespace tts {

// ------------------------------------------------------------------ per-device state
// cudaFuncSetAttribute and the SM count are per DEVICE; a process may drive several GPUs, so both are
// cached per device ordinal under a mutex.
static std::mutex g_dev_mu;
static int g_sm_count[kMaxDevices] = {};

const char* device_info(DevInfo* out) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return "cudaGetDevice failed";
  if (dev < 0 || dev >= kMaxDevices) return "device ordinal out of range (kMaxDevices)";
  std::lock_guard<std::mutex> lock(g_dev_mu);
  if (g_sm_count[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      return "cannot query the SM count";
    g_sm_count[dev] = n;
  }
  out->dev = dev;
  out->sm_count = g_sm_count[dev];
  return nullptr;
}

bool PerDeviceOnce::need(int dev) {
  std::lock_guard<std::mutex> lock(g_dev_mu);
  return !done_[dev];
}
void PerDeviceOnce::mark(int dev) {
  std::lock_guard<std::mutex> lock(g_dev_mu);
  done_[dev] = true;
}

// ------------------------------------------------------------------ tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  std::lock_guard<std::mutex> lock(g_dev_mu);
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

// ------------------------------------------------------------------ dispatchers
const char* gemm_launch(const GemmArgs& a, cudaStream_t s) {
  if (a.B <= 0 || a.T <= 0 || a.N <= 0) return "gemm: empty problem";
  if ((a.C * 2) % 16 != 0) return "gemm: activation row pitch must be a multiple of 16 bytes";
  if (a.ln_g != nullptr && a.N != 256) return "gemm: LayerNorm epilogue needs N == 256";
  if (a.pe != nullptr && (a.N != kD || a.pe_period <= 0)) return "gemm: PE epilogue needs N == 256 and a period";
  if ((reinterpret_cast<uintptr_t>(a.A) & 15) != 0 || (reinterpret_cast<uintptr_t>(a.W) & 15) != 0)
    return "gemm: operands must be 16-byte aligned";
  if (!gemm2_supports(a))
    return "gemm: unsupported shape (bf16 output: N a multiple of 256, 16-byte aligned rows; fp32 output: N <= 128, "
           "no LayerNorm / PE / bf16 residual)";
  return gemm2_launch(a, s);
}

const char* attn_launch(const AttnArgs& a, cudaStream_t s) {
  if (a.B <= 0 || a.Tq <= 0 || a.Tk <= 0) return "attention: empty problem";
  if ((a.ldq | a.ldk | a.ldv | a.ldo) % 8 != 0) return "attention: leading dimensions must be multiples of 8";
  // probabilities wanted and short keys: single-pass softmax with the block store (without the store the
  // streaming kernel is as fast: 18.8 vs 22.6 us on the encoder shape)
  if (a.probs != nullptr && attn3_supports(a)) return attn3_launch(a, s);
  if (attn2_supports(a)) return attn2_launch(a, s);
  return "attention: q / k / v / out must be 16-byte aligned";
}

}  // namespace tts
