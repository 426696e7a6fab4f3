// common.cuh — shared device helpers for the lcb200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cuda_fp8.h>
#include <stdint.h>
#include <atomic>
#include "../../include/lcb200.h"

namespace lcb {

void set_last_error(const char* what, cudaError_t e);

#define LCB_CUDA_CHECK_LAUNCH(what)                     \
  do {                                                  \
    cudaError_t _e = cudaGetLastError();                \
    if (_e != cudaSuccess) {                            \
      ::lcb::set_last_error(what, _e);                  \
      return LCB_E_CUDA;                                \
    }                                                   \
  } while (0)

constexpr int kNumSMs = 148;  // B200

// cudaFuncSetAttribute is a per-device (per-context) setting: one "configured" bit per device, so that a process that
// drives several GPUs opts every one of them in. Setting an attribute twice from two racing threads is harmless.
struct DeviceOnce {
  std::atomic<unsigned long long> mask{0};
  int dev = 0;
  bool needed() {
    int d = 0;
    cudaGetDevice(&d);
    if (d < 0 || d >= 64) return true;
    return ((mask.load(std::memory_order_acquire) >> d) & 1ull) == 0;
  }
  void done() {
    int d = 0;
    cudaGetDevice(&d);
    if (d >= 0 && d < 64) mask.fetch_or(1ull << d, std::memory_order_release);
  }
};

// ---------------------------------------------------------------------------------------------
// Storage-type traits: 16-byte vectors, unpack to fp32, pack with round-to-nearest-even.
// ---------------------------------------------------------------------------------------------
template <typename T>
struct TI;

template <>
struct TI<float> {
  static constexpr int VEC = 4;
  static constexpr int ID = LCB_F32;
  __device__ __forceinline__ static void unpack(const uint4& v, float* f) {
    f[0] = __uint_as_float(v.x);
    f[1] = __uint_as_float(v.y);
    f[2] = __uint_as_float(v.z);
    f[3] = __uint_as_float(v.w);
  }
  __device__ __forceinline__ static uint4 pack(const float* f) {
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]),
                      __float_as_uint(f[3]));
  }
  __device__ __forceinline__ static float rd(float x) { return x; }
  __device__ __forceinline__ static float load1(const float* p) { return *p; }
  __device__ __forceinline__ static void store1(float* p, float v) { *p = v; }
};

template <>
struct TI<__half> {
  static constexpr int VEC = 8;
  static constexpr int ID = LCB_F16;
  __device__ __forceinline__ static void unpack(const uint4& v, float* f) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __half2 h = *reinterpret_cast<const __half2*>(&w[i]);
      float2 t = __half22float2(h);
      f[2 * i] = t.x;
      f[2 * i + 1] = t.y;
    }
  }
  __device__ __forceinline__ static uint4 pack(const float* f) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __half2 h = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
  __device__ __forceinline__ static float rd(float x) { return __half2float(__float2half_rn(x)); }
  __device__ __forceinline__ static float load1(const __half* p) { return __half2float(*p); }
  __device__ __forceinline__ static void store1(__half* p, float v) { *p = __float2half_rn(v); }
};

template <>
struct TI<__nv_bfloat16> {
  static constexpr int VEC = 8;
  static constexpr int ID = LCB_BF16;
  __device__ __forceinline__ static void unpack(const uint4& v, float* f) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ __forceinline__ static uint4 pack(const float* f) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
  __device__ __forceinline__ static float rd(float x) {
    uint32_t d;  // F2FP.BF16.F32.PACK_AB (full-rate) instead of the conversion-pipe F2F
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(0.0f), "f"(x));
    return __uint_as_float(d << 16);
  }
  __device__ __forceinline__ static float load1(const __nv_bfloat16* p) {
    return __bfloat162float(*p);
  }
  __device__ __forceinline__ static void store1(__nv_bfloat16* p, float v) {
    *p = __float2bfloat16_rn(v);
  }
};

// streaming 128-bit global accesses (read-once / write-once data: keep it out of L1)
__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

// NaN-propagating min / max (torch.amax / amin propagate NaN)
__device__ __forceinline__ float max_nan(float a, float b) {
  float r;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ float min_nan(float a, float b) {
  float r;
  asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}

// round a fp32 value to the runtime math dtype and back
__device__ __forceinline__ float rd_rt(float x, int dt) {
  if (dt == LCB_BF16) return __bfloat162float(__float2bfloat16_rn(x));
  if (dt == LCB_F16) return __half2float(__float2half_rn(x));
  return x;
}

__device__ __forceinline__ float load_rt(const void* p, int64_t i, int dt) {
  if (dt == LCB_BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
  if (dt == LCB_F16) return __half2float(reinterpret_cast<const __half*>(p)[i]);
  return reinterpret_cast<const float*>(p)[i];
}
__device__ __forceinline__ void store_rt(void* p, int64_t i, float v, int dt) {
  if (dt == LCB_BF16)
    reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(v);
  else if (dt == LCB_F16)
    reinterpret_cast<__half*>(p)[i] = __float2half_rn(v);
  else
    reinterpret_cast<float*>(p)[i] = v;
}

// fp32 -> fp8 (RNE, saturate-to-finite) -> fp32: the rounding of Tensor.to(float8_*) for in-range values
__device__ __forceinline__ uint8_t f32_to_fp8_byte(float x, int kind) {
  return (uint8_t)__nv_cvt_float_to_fp8(x, __NV_SATFINITE, kind == LCB_KIND_FP8_E4M3 ? __NV_E4M3 : __NV_E5M2);
}
__device__ __forceinline__ float fp8_byte_to_f32(uint8_t b, int kind) {
  __half_raw h = __nv_cvt_fp8_to_halfraw((__nv_fp8_storage_t)b, kind == LCB_KIND_FP8_E4M3 ? __NV_E4M3 : __NV_E5M2);
  return __half2float(*reinterpret_cast<__half*>(&h));
}

}  // namespace lcb
