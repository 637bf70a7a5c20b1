// Shared device helpers for the delay-spectrum Gibbs sampler (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dg {

constexpr int TILE = 128;     // block-column width and row-tile height of the factorisation
constexpr int KC = 32;        // K chunk staged per pipeline stage (one CTA barrier per chunk: 256 DMMAs per warp between barriers)
constexpr int THREADS = 256;  // 8 warps per CTA, one CTA per SM
constexpr int NWARP = THREADS / 32;
// window ids beyond the cosine-sum table (window.cuh, include/dgibbs.h): "triangular", "tukey-r" with r in 1e-6 units
constexpr int DG_WINDOW_TRIANGULAR = 8;
constexpr int DG_WINDOW_TUKEY_BASE = 1 << 20;
__host__ __device__ __forceinline__ bool window_id_valid(int id) {
  return (id >= -1 && id <= DG_WINDOW_TRIANGULAR) || (id >= DG_WINDOW_TUKEY_BASE && id <= DG_WINDOW_TUKEY_BASE + 1000000);
}
constexpr int LDA_NT = KC + 4;    // smem row pitch (doubles) of a [rows][KC] tile: conflict-free fragments
constexpr int LDB_NN = TILE + 4;  // smem row pitch of a [KC][TILE] tile
constexpr int LDW = TILE + 4;     // smem row pitch of a full 128x128 tile

// ---------------------------------------------------------------- fp64 tensor pipe
// D(8x8) += A(8x4) * B(4x8).  Lane l: g = l>>2, q = l&3.
//   a = A[g][q]   b = B[q][g]   c0 = C[g][2q]  c1 = C[g][2q+1]
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// ---------------------------------------------------------------- async copies
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// ---------------------------------------------------------------- visibilities
// Element `idx` of the caller's visibility array: complex128, or complex64 widened exactly (draco's containers store
// complex64, core/containers.py:503; shipping them as they are halves the PCIe bytes of the host-pointer path).
__device__ __forceinline__ double2 ld_vis(const double2* data, int c64, long long idx) {
  if (c64) {
    const float2 v = reinterpret_cast<const float2*>(data)[idx];
    return make_double2((double)v.x, (double)v.y);
  }
  return data[idx];
}

// ---------------------------------------------------------------- Philox4x32-10
struct Philox {
  uint32_t k0, k1;
  __device__ __forceinline__ Philox(uint64_t key) : k0((uint32_t)key), k1((uint32_t)(key >> 32)) {}
  __device__ __forceinline__ uint4 operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const {
    uint32_t a = k0, b = k1;
#pragma unroll
    for (int r = 0; r < 10; r++) {
      uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      uint32_t n0 = hi1 ^ c1 ^ a, n1 = lo1, n2 = hi0 ^ c3 ^ b, n3 = lo0;
      c0 = n0; c1 = n1; c2 = n2; c3 = n3;
      a += 0x9E3779B9u; b += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
};

// uniform in (0, 1], 53 bits
__device__ __forceinline__ double u01(uint32_t lo, uint32_t hi) {
  uint64_t v = ((uint64_t)hi << 32) | lo;
  return ((double)(v >> 11) + 1.0) * (1.0 / 9007199254740992.0);
}

// Two independent standard normals from one Philox block (Box-Muller, fp64).
__device__ __forceinline__ void normal_pair(const uint4& r, double& z0, double& z1) {
  double u1 = u01(r.x, r.y), u2 = u01(r.z, r.w);
  double rad = sqrt(-2.0 * log(u1));
  double s, c;
  sincospi(2.0 * u2, &s, &c);
  z0 = rad * c;
  z1 = rad * s;
}

// Same, with the transcendental part in fp32 (32-bit uniforms, fast intrinsics): the variates have ~1e-7
// granularity and reach 6.7 sigma -- used for the N(0,1) perturbations w1, w2 of the native-RNG mode, where
// parity is statistical (3 sigma / KS); ~15x fewer instructions than the fp64 version.
__device__ __forceinline__ void normal_pair_fast(const uint4& r, double& z0, double& z1, double& z2, double& z3) {
  const float k = 2.3283064365386963e-10f;  // 2^-32
  const float ua = ((float)r.x + 0.5f) * k, ub = ((float)r.z + 0.5f) * k;
  const float ra = sqrtf(-2.0f * __logf(fminf(ua, 1.0f))), rb = sqrtf(-2.0f * __logf(fminf(ub, 1.0f)));
  float s, c;
  __sincosf(6.2831853071795865f * ((float)r.y * k), &s, &c);
  z0 = (double)(ra * c);
  z1 = (double)(ra * s);
  __sincosf(6.2831853071795865f * ((float)r.w * k), &s, &c);
  z2 = (double)(rb * c);
  z3 = (double)(rb * s);
}

// RNG stream ids inside one (baseline, iteration)
enum { STREAM_W1 = 0, STREAM_W2 = 1, STREAM_CHI2 = 2, STREAM_W1_ODD = 9, STREAM_W2_ODD = 10 };

// The four N(0,1) perturbations that share one counter (c0, c1, iteration): two for the even sample of a pair (z0, z1)
// and two for the odd one (z2, z3).  Default: fp64 Box-Muller (53-bit uniforms, fp64 log / sincospi -- the resolution of
// the reference's numpy normals), one Philox block per sample: `stream` for the even, `stream_odd` for the odd sample.
// -DDGIBBS_FAST_NORMALS=1 restores the fp32 variant (all four from the `stream` block; 1 % faster at cfg2).
#ifndef DGIBBS_FAST_NORMALS
#define DGIBBS_FAST_NORMALS 0
#endif
__device__ __forceinline__ void normal_quad(const Philox& ph, uint32_t c0, uint32_t c1, uint32_t it, uint32_t stream,
                                            uint32_t stream_odd, bool even, bool odd, double& z0, double& z1, double& z2,
                                            double& z3) {
#if DGIBBS_FAST_NORMALS
  normal_pair_fast(ph(c0, c1, it, stream), z0, z1, z2, z3);
#else
  z0 = z1 = z2 = z3 = 0.0;
  if (even) normal_pair(ph(c0, c1, it, stream), z0, z1);
  if (odd) normal_pair(ph(c0, c1, it, stream_odd), z2, z3);
#endif
}

__device__ __forceinline__ uint64_t mix_key(uint64_t seed, uint64_t gid) {
  // splitmix64 of (seed, global baseline id): the Philox key of a baseline
  uint64_t z = seed + 0x9E3779B97F4A7C15ull * (gid + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// Standard normal number `idx` of stream `stream` of iteration `iter` (pair-wise generated:
// element idx uses block idx>>1, lane idx&1).
__device__ __forceinline__ double philox_normal(const Philox& ph, uint32_t iter, uint32_t stream, uint64_t idx) {
  uint64_t blk = idx >> 1;
  uint4 r = ph((uint32_t)blk, (uint32_t)(blk >> 32), iter, stream);
  double z0, z1;
  normal_pair(r, z0, z1);
  return (idx & 1) ? z1 : z0;
}

// Entry (r, c) of FTNiF = F^T N^-1 F, which is circulant (real mode) or 2x2-block circulant
// (complex time domain); ct = first column c[k] (cplx: cr[0..N) then ci[0..N)).
__device__ __forceinline__ double circ_entry(const double* ct, int N, int cplx, int r, int c) {
  if (!cplx) {
    int k = r - c;
    if (k < 0) k += N;
    return ct[k];
  }
  int k = (r >> 1) - (c >> 1);
  if (k < 0) k += N;
  int pr = r & 1, pc = c & 1;
  return (pr == pc) ? ct[k] : (pr ? ct[N + k] : -ct[N + k]);
}

}  // namespace dg
