// Host utilities: error string, TMA tensor-map construction, device info.
#include <stdarg.h>
#include <stdlib.h>
#include <map>
#include <mutex>
#include <set>
#include <string.h>
#include <utility>

#include "../include/tfd_b200.h"
#include "tfd_common.cuh"

static thread_local char g_err[1024] = "";

void tfd_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* tfd_last_error(void) { return g_err; }

extern "C" int tfd_version(void) { return 100; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess || !p) {
    return nullptr;
  }
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

static CUtensorMapSwizzle swz(int code) {
  switch (code) {
    case 129: return CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B;  // 128 B span, 32 B atoms: MN-major tf32 operands
    case 128: return CU_TENSOR_MAP_SWIZZLE_128B;
    case 64: return CU_TENSOR_MAP_SWIZZLE_64B;
    case 32: return CU_TENSOR_MAP_SWIZZLE_32B;
    default: return CU_TENSOR_MAP_SWIZZLE_NONE;
  }
}
static int swz_span(int code) { return code == 129 ? 128 : code; }

int tfd_make_tmap_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                     uint32_t box_cols, int swizzle, int elem_bytes) {
  EncodeTiledFn fn = get_encode_tiled();
  TFD_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled entry point not available");
  TFD_REQUIRE(elem_bytes == 2 || elem_bytes == 4, "TMA element size %d unsupported", elem_bytes);
  TFD_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base %p not 16-byte aligned", base);
  TFD_REQUIRE((ld * elem_bytes) % 16 == 0, "TMA row pitch %llu B not a multiple of 16",
              (unsigned long long)(ld * elem_bytes));
  TFD_REQUIRE(box_cols * elem_bytes <= (uint32_t)(swizzle ? swz_span(swizzle) : 256) && box_rows <= 256 && box_cols <= 256,
              "TMA box %ux%u (%d-byte elements) does not fit swizzle %d", box_rows, box_cols, elem_bytes, swizzle);
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstr[1] = {ld * elem_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                  const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz(swizzle),
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TFD_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(2d %llux%llu ld %llu box %ux%u esz %d swz %d) failed: %d",
              (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, box_rows, box_cols, elem_bytes,
              swizzle, (int)r);
  return 0;
}

int tfd_make_tmap_nhwc(CUtensorMap* out, const void* base, uint64_t N, uint64_t H, uint64_t W, uint64_t C, uint32_t bn,
                       uint32_t bh, uint32_t bw, uint32_t bc, int swizzle, int elem_bytes) {
  EncodeTiledFn fn = get_encode_tiled();
  TFD_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled entry point not available");
  TFD_REQUIRE(elem_bytes == 2 || elem_bytes == 4, "TMA element size %d unsupported", elem_bytes);
  TFD_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base %p not 16-byte aligned", base);
  TFD_REQUIRE((C * elem_bytes) % 16 == 0 && bc <= C && (int)(bc * elem_bytes) <= swz_span(swizzle),
              "NHWC channel box %u of %llu (%d-byte elements) unsupported with swizzle %d", bc, (unsigned long long)C,
              elem_bytes, swizzle);
  const uint64_t e = (uint64_t)elem_bytes;
  cuuint64_t gdim[4] = {C, W, H, N};
  cuuint64_t gstr[3] = {C * e, W * C * e, H * W * C * e};
  cuuint32_t box[4] = {bc, bw, bh, bn};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(out, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4,
                  const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz(swizzle),
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  TFD_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(nhwc %llu,%llu,%llu,%llu box %u,%u,%u,%u esz %d) failed: %d",
              (unsigned long long)N, (unsigned long long)H, (unsigned long long)W, (unsigned long long)C, bn, bh, bw, bc,
              elem_bytes, (int)r);
  return 0;
}

extern "C" int tfd_device_info(int device, int* sm_count, int* cc_major, int* cc_minor, size_t* smem_optin) {
  cudaDeviceProp p;
  TFD_CUDA_CHECK(cudaGetDeviceProperties(&p, device));
  if (sm_count) *sm_count = p.multiProcessorCount;
  if (cc_major) *cc_major = p.major;
  if (cc_minor) *cc_minor = p.minor;
  if (smem_optin) *smem_optin = p.sharedMemPerBlockOptin;
  return 0;
}

extern "C" int tfd_set_device(int device) {
  TFD_CUDA_CHECK(cudaSetDevice(device));
  return 0;
}

// PDL switch.  Measured on B200 (bench.py, 1 GPU): 121.9 us/step with PDL vs 119.9 us without -- the step kernels
// hold ~200 KB of shared memory per SM, so a successor's CTAs cannot become resident before the predecessor's
// leave and there is no prologue to overlap.  Default off; TFD_PDL=1 or tfd_set_pdl(1) turns it on.
static int g_pdl = -1;
int tfd_pdl_enabled() {
  if (g_pdl < 0) {
    const char* e = getenv("TFD_PDL");
    g_pdl = (e && e[0] == '1') ? 1 : 0;
  }
  return g_pdl;
}
extern "C" int tfd_set_pdl(int on) {
  g_pdl = on ? 1 : 0;
  return 0;
}

// cudaFuncSetAttribute applies to the CURRENT device only, so "already configured" is remembered per
// (device, kernel): a second GPU driven from the same process gets its own opt-in to > 48 KB of dynamic shared memory.
int tfd_ensure_dyn_smem(const void* kernel, int bytes) {
  static std::mutex mu;
  static std::map<std::pair<int, const void*>, int> configured;
  int dev = 0;
  TFD_CUDA_CHECK(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(mu);
  int& have = configured[std::make_pair(dev, kernel)];
  if (have >= bytes) return 0;
  TFD_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  have = bytes;
  return 0;
}

int tfd_num_sms() {
  static std::mutex mu;
  static std::map<int, int> sms;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  std::lock_guard<std::mutex> lock(mu);
  int& n = sms[dev];
  if (n <= 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  }
  return n;
}

// Hint, once per (device, kernel): run this kernel with the largest shared-memory carve-out, the configuration the
// tcgen05 kernels (200+ KB of shared memory) put an SM into -- so that its blocks can become resident NEXT to such a
// CTA instead of waiting for the SM to drain and reconfigure.
void tfd_prefer_max_carveout(const void* kernel) {
  static std::mutex mu;
  static std::set<std::pair<int, const void*>> done;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return;
  std::lock_guard<std::mutex> lock(mu);
  if (!done.insert(std::make_pair(dev, kernel)).second) return;
  if (cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared) !=
      cudaSuccess)
    cudaGetLastError();  // a hint only: never fail a launch over it
}
