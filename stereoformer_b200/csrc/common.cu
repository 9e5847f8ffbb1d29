#include "common.cuh"

#include <stdarg.h>

#include <mutex>

#include "../../include/stereoformer_b200.h"

namespace sf {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static encode_tiled_fn get_encode() {
  static encode_tiled_fn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<encode_tiled_fn>(p);
  });
  return fn;
}

// Encoded tensor maps are a pure function of (base pointer, element type, dims, box, swizzle): a caller that runs the
// same operator on the same buffers step after step (the reference's eager LowCNN.forward does; so does any serving
// loop with static buffers) gets them from a small per-thread table instead of a driver call per operand per launch.
struct TmapKey {
  const void* base;
  uint64_t n0, n1, n2;
  uint32_t b0, b1, b2;
  int dt, swz;
  bool operator==(const TmapKey& o) const {
    return base == o.base && n0 == o.n0 && n1 == o.n1 && n2 == o.n2 && b0 == o.b0 && b1 == o.b1 && b2 == o.b2 &&
           dt == o.dt && swz == o.swz;
  }
};
struct TmapSlot { TmapKey key; CUtensorMap map; bool used; };
constexpr int TMAP_SLOTS = 128;

static int make_tmap_3d(CUtensorMap* out, CUtensorMapDataType dt, int esz, const void* base, uint64_t n0,
                        uint64_t n1, uint64_t n2, uint32_t b0, uint32_t b1, uint32_t b2,
                        CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_NONE) {
  static thread_local TmapSlot cache[TMAP_SLOTS] = {};
  const TmapKey key{base, n0, n1, n2, b0, b1, b2, static_cast<int>(dt), static_cast<int>(swz)};
  uint64_t h = reinterpret_cast<uintptr_t>(base) >> 8;
  h = (h ^ (n0 * 0x9E3779B97F4A7C15ULL) ^ (n1 << 20) ^ (n2 << 40) ^ (static_cast<uint64_t>(b0) << 8) ^ b1 ^
       (static_cast<uint64_t>(b2) << 16) ^ (static_cast<uint64_t>(swz) << 3)) * 0xD6E8FEB86659FD93ULL;
  TmapSlot& slot = cache[(h >> 40) % TMAP_SLOTS];
  if (slot.used && slot.key == key) {
    *out = slot.map;
    return SF_OK;
  }
  encode_tiled_fn enc = get_encode();
  if (!enc) {
    set_error("cuTensorMapEncodeTiled driver entry point not available");
    return SF_ERR_CUDA;
  }
  SF_CHECK_ARG((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base pointer must be 16-byte aligned");
  SF_CHECK_ARG((n0 * esz) % 16 == 0 && (b0 * esz) % 16 == 0 && b0 <= 256 && b1 <= 256 && b2 <= 256,
               "TMA shape constraint violated (n0=%llu b0=%u)", (unsigned long long)n0, b0);
  cuuint64_t dims[3] = {n0, n1, n2};
  cuuint64_t strides[2] = {n0 * esz, n0 * n1 * esz};
  cuuint32_t box[3] = {b0, b1, b2};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, dt, 3, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (dims %llu,%llu,%llu box %u,%u,%u)", (int)r,
              (unsigned long long)n0, (unsigned long long)n1, (unsigned long long)n2, b0, b1, b2);
    return SF_ERR_CUDA;
  }
  slot.key = key;
  slot.map = *out;
  slot.used = true;
  return SF_OK;
}

int make_tmap_3d_f32(CUtensorMap* out, const void* base, uint64_t n0, uint64_t n1, uint64_t n2, uint32_t b0,
                     uint32_t b1, uint32_t b2) {
  return make_tmap_3d(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, base, n0, n1, n2, b0, b1, b2);
}
int make_tmap_3d_bf16(CUtensorMap* out, const void* base, uint64_t n0, uint64_t n1, uint64_t n2, uint32_t b0,
                      uint32_t b1, uint32_t b2) {
  return make_tmap_3d(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, n0, n1, n2, b0, b1, b2);
}

// fp32 with an explicit shared-memory swizzle (0 / 32 / 64 / 128 bytes; the box's inner extent must then be exactly
// that many bytes) -- used by the TMA-store epilogues whose staging tiles are written conflict-free by row.
int make_tmap_3d_f32_swz(CUtensorMap* out, const void* base, uint64_t n0, uint64_t n1, uint64_t n2, uint32_t b0,
                         uint32_t b1, uint32_t b2, int swizzle_bytes) {
  CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_NONE;
  if (swizzle_bytes == 32) swz = CU_TENSOR_MAP_SWIZZLE_32B;
  else if (swizzle_bytes == 64) swz = CU_TENSOR_MAP_SWIZZLE_64B;
  else if (swizzle_bytes == 128) swz = CU_TENSOR_MAP_SWIZZLE_128B;
  return make_tmap_3d(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, base, n0, n1, n2, b0, b1, b2, swz);
}

// bf16, 128-byte swizzle: the box's inner extent must be exactly 64 elements (128 B) -- the UMMA SW128 atom.
int make_tmap_3d_bf16_sw128(CUtensorMap* out, const void* base, uint64_t n0, uint64_t n1, uint64_t n2, uint32_t b1,
                            uint32_t b2) {
  return make_tmap_3d(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, n0, n1, n2, 64, b1, b2,
                      CU_TENSOR_MAP_SWIZZLE_128B);
}

}  // namespace sf

extern "C" const char* sf_last_error(void) { return sf::g_err; }
extern "C" int sf_abi_version(void) { return SF_ABI_VERSION; }
extern "C" int sf_device_sm_count(void) { return sf::sm_count(); }
