// CPU emulation of the warp-level matrix primitives csrc/window_attn_tc.cu is written against, on top of cuda_emu.h:
//   mma16816<T>   = mma.sync.aligned.m16n8k16.row.col.f32.{bf16,f16}.{bf16,f16}.f32
//   ldsm_x4       = ldmatrix.sync.aligned.m8n8.x4.shared.b16
//   ldsm_x4_t     = ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16
//   cp_async16 / cp_async4 / cp_async_commit / cp_async_wait<N>   (copies happen immediately here)
// The fragment layouts are the ones of the PTX ISA tables: with g = lane / 4 and t = lane % 4,
//   A (16 x 16, row):  a[i] at row g + 8 * ((i / 2) % 2), column 2 t + (i % 2) + 8 * (i / 4)            i = 0..7, two per register
//   B (16 x 8,  col):  b[i] at row (k) 2 t + (i % 2) + 8 * (i / 2), column (n) g                        i = 0..3
//   C (16 x 8):        c[i] at row g + 8 * (i / 2), column 2 t + (i % 2)                                i = 0..3
//   ldmatrix: lane l supplies the address of row l % 8 of matrix l / 8; register i of lane (g, t) holds elements
//             M_i[g][2 t], M_i[g][2 t + 1]   (plain)   or   M_i[2 t][g], M_i[2 t + 1][g]   (.trans)
// Every lane of a warp must call these together (they are .sync.aligned on hardware as well).
#pragma once
#include "cuda_emu.h"

namespace emu_mma {
inline emu::Warp& warp() { return emu::cur()->warps[threadIdx.x >> 5]; }
inline void wbar() { pthread_barrier_wait(&warp().bar); }
struct Scratch { float a[16][16]; float b[16][8]; };
inline Scratch& scratch() { static Scratch s[32]; return s[threadIdx.x >> 5]; }
template <typename T> inline float up(uint16_t bits);
template <> inline float up<__nv_bfloat16>(uint16_t bits) { return __bfloat162float(__nv_bfloat16{bits}); }
template <> inline float up<__half>(uint16_t bits) { return __half2float(__half{bits}); }
}  // namespace emu_mma

template <typename T>
inline void mma16816(float c[4], const uint32_t a[4], uint32_t b0, uint32_t b1) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  emu_mma::Scratch& s = emu_mma::scratch();
  for (int i = 0; i < 8; ++i) {
    const uint16_t bits = static_cast<uint16_t>(a[i >> 1] >> ((i & 1) * 16));
    s.a[g + 8 * ((i >> 1) & 1)][2 * t + (i & 1) + 8 * (i >> 2)] = emu_mma::up<T>(bits);
  }
  const uint32_t b[2] = {b0, b1};
  for (int i = 0; i < 4; ++i) {
    const uint16_t bits = static_cast<uint16_t>(b[i >> 1] >> ((i & 1) * 16));
    s.b[2 * t + (i & 1) + 8 * (i >> 1)][g] = emu_mma::up<T>(bits);
  }
  emu_mma::wbar();
  for (int i = 0; i < 4; ++i) {
    const int row = g + 8 * (i >> 1), col = 2 * t + (i & 1);
    float acc = c[i];
    for (int k = 0; k < 16; ++k) acc += s.a[row][k] * s.b[k][col];
    c[i] = acc;
  }
  emu_mma::wbar();
}

inline void ldsm_x4_impl(uint32_t r[4], const void* row_ptr, bool trans) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  emu::Warp& w = emu_mma::warp();
  w.slots[lane] = reinterpret_cast<uint64_t>(row_ptr);
  emu_mma::wbar();
  for (int i = 0; i < 4; ++i) {
    auto row = [&](int rr) { return reinterpret_cast<const uint16_t*>(w.slots[i * 8 + rr]); };
    const uint16_t lo = trans ? row(2 * t)[g] : row(g)[2 * t];
    const uint16_t hi = trans ? row(2 * t + 1)[g] : row(g)[2 * t + 1];
    r[i] = static_cast<uint32_t>(lo) | (static_cast<uint32_t>(hi) << 16);
  }
  emu_mma::wbar();
}
inline void ldsm_x4(uint32_t r[4], const void* row_ptr) { ldsm_x4_impl(r, row_ptr, false); }
inline void ldsm_x4_t(uint32_t r[4], const void* row_ptr) { ldsm_x4_impl(r, row_ptr, true); }

inline void cp_async16(void* smem_dst, const void* gmem_src) { std::memcpy(smem_dst, gmem_src, 16); }
inline void cp_async4(void* smem_dst, const void* gmem_src) { std::memcpy(smem_dst, gmem_src, 4); }
inline void cp_async_commit() {}
template <int N> inline void cp_async_wait() {}

template <typename T> inline T __ldg(const T* p) { return *p; }

// 16-bit pairs <-> fp32
template <typename T> inline uint32_t pack2(float lo, float hi);
template <> inline uint32_t pack2<__nv_bfloat16>(float lo, float hi) {
  return static_cast<uint32_t>(__float2bfloat16_rn(lo).v) | (static_cast<uint32_t>(__float2bfloat16_rn(hi).v) << 16);
}
template <> inline uint32_t pack2<__half>(float lo, float hi) {
  return static_cast<uint32_t>(__float2half_rn(lo).v) | (static_cast<uint32_t>(__float2half_rn(hi).v) << 16);
}
template <typename T> inline void unpack2(uint32_t u, float& lo, float& hi) {
  lo = emu_mma::up<T>(static_cast<uint16_t>(u));
  hi = emu_mma::up<T>(static_cast<uint16_t>(u >> 16));
}
