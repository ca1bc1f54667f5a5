// CUDA vector types, 16-bit pair conversions and the warp reductions of csrc/common.cuh for kernels that run under cuda_emu.h
#pragma once
#include "cuda_emu.h"

struct uint4 { uint32_t x, y, z, w; };
inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return uint4{x, y, z, w}; }
struct float2 { float x, y; };

template <typename T> inline float emu_up16(uint16_t bits);
template <> inline float emu_up16<__nv_bfloat16>(uint16_t bits) { return __bfloat162float(__nv_bfloat16{bits}); }
template <> inline float emu_up16<__half>(uint16_t bits) { return __half2float(__half{bits}); }
template <typename T> inline uint16_t emu_down16(float f);
template <> inline uint16_t emu_down16<__nv_bfloat16>(float f) { return __float2bfloat16_rn(f).v; }
template <> inline uint16_t emu_down16<__half>(float f) { return __float2half_rn(f).v; }
template <typename T> inline float2 emu_unpack2(uint32_t u) {
  return float2{emu_up16<T>(static_cast<uint16_t>(u)), emu_up16<T>(static_cast<uint16_t>(u >> 16))};
}
template <typename T> inline uint32_t emu_pack2(float lo, float hi) {
  return static_cast<uint32_t>(emu_down16<T>(lo)) | (static_cast<uint32_t>(emu_down16<T>(hi)) << 16);
}

inline float warp_sum(float v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
