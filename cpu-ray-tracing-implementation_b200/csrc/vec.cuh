// vec.cuh — binary64 3-vectors, rays and the Philox4x32-10 draw addressing of the B200 backend.
//
// Arithmetic follows the reference's vec3 (src/vec3.h:61-84) operation for operation: the library is
// compiled with -fmad=false so that nvcc emits separate DMUL/DADD exactly where x86-64 SSE2 does, which
// is what makes hit records bit-identical to the reference (DESIGN.md "Numerics").  Where a fused
// multiply-add is wanted (conservative box tests) it is written explicitly with __fma_rn.
#pragma once
#include <cstdint>
#ifdef B200_HOST_EMU
#include "host_emu.h" // tests/emu only: lets the CPU test-suite execute these device functions without a GPU
#else
#include <cuda_runtime.h>
#endif

#define HD __host__ __device__ __forceinline__
#define DEV __device__ __forceinline__
// Experiment (-DB200_COLD_NOINLINE=1): cold paths (volumes, noise textures) called instead of inlined, to shrink the hot loops'
// instruction footprint (ncu: stall_no_instruction was 24 % of the samples of the RTNW extend kernel).  Measured on B200
// (profiles/r02_g_builds.jsonl, r02_h_builds.jsonl): SLOWER on every scene, 2-9 %, whether the scene is passed by reference
// (the kernel parameter gets a local-memory copy per thread) or through DevScene::self: an out-of-line call forces the ray
// and the entry it takes by reference into local memory.  Off.
#ifndef B200_COLD_NOINLINE
#define B200_COLD_NOINLINE 0
#endif
#if B200_COLD_NOINLINE
#define DEV_COLD __device__ __noinline__
#else
#define DEV_COLD __device__ __forceinline__
#endif

namespace b200 {

struct V {
  double x, y, z;
};
HD V v3(double a, double b, double c) { return V{a, b, c}; }
HD V operator+(V a, V b) { return V{a.x + b.x, a.y + b.y, a.z + b.z}; }
HD V operator-(V a, V b) { return V{a.x - b.x, a.y - b.y, a.z - b.z}; }
HD V operator*(V a, V b) { return V{a.x * b.x, a.y * b.y, a.z * b.z}; }
HD V operator*(V a, double c) { return V{a.x * c, a.y * c, a.z * c}; }
HD V operator*(double c, V a) { return V{a.x * c, a.y * c, a.z * c}; }
// vec3 / double is three IEEE divisions by the same number (src/vec3.h:75).  A binary64 division costs ~35 instructions on the
// GPU and unit_vector() alone made a third of the shading kernels' instruction stream, so on the device the three quotients
// share ONE correctly rounded reciprocal:  y = RN(1/c);  q0 = RN(a y);  r = a - c q0 (exact, one fma);  q = RN(q0 + r y).
// With y correctly rounded and no overflow / underflow / subnormal anywhere in the sequence, q is the correctly rounded a / c
// (Markstein's division theorem; checked bit for bit against x86 division on 2.5e9 random, special-mantissa and near-midpoint
// cases).  Everything outside that range (zero or non-finite numerators, exponents beyond +-500) takes the plain divisions.
// Measured on B200 (profiles/r01_k_layout.txt): all 140 GPU parity tests stay bit-exact and the lambertian shade kernel
// executes ~20 % fewer instructions, but the frame time does not move (5.61 vs 5.62 Grays/s): the shade kernels wait on
// dependent gathers, not on the issue slots.  Off by default (no gain, one more thing to trust); -DB200_SHARED_RCP_DIV=1.
#ifndef B200_SHARED_RCP_DIV
#define B200_SHARED_RCP_DIV 0
#endif
HD V operator/(V a, double c) {
#if defined(__CUDA_ARCH__) && B200_SHARED_RCP_DIV
  const int ec = (__double2hiint(c) >> 20) & 0x7ff, ex = (__double2hiint(a.x) >> 20) & 0x7ff, ey = (__double2hiint(a.y) >> 20) & 0x7ff,
            ez = (__double2hiint(a.z) >> 20) & 0x7ff;
  const int lo = min(min(ec, ex), min(ey, ez)), hi = max(max(ec, ex), max(ey, ez));
  if (lo > 523 && hi < 1523) { // every operand normal, |.| in [2^-499, 2^500)
    const double y = __drcp_rn(c);
    const double qx = __dmul_rn(a.x, y), qy = __dmul_rn(a.y, y), qz = __dmul_rn(a.z, y);
    return V{__fma_rn(__fma_rn(-c, qx, a.x), y, qx), __fma_rn(__fma_rn(-c, qy, a.y), y, qy), __fma_rn(__fma_rn(-c, qz, a.z), y, qz)};
  }
#endif
  return V{a.x / c, a.y / c, a.z / c};
}
HD V operator-(V a) { return V{-a.x, -a.y, -a.z}; }
HD double dot(V a, V b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
HD V cross(V a, V b) { return V{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
HD double length_squared(V a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
HD double length(V a) { return sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }
HD V unit_vector(V a) { return a / length(a); }
HD V vfloor(V a) { return V{floor(a.x), floor(a.y), floor(a.z)}; }
HD double getc(const V &a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
HD void setc(V &a, int i, double v) {
  if (i == 0)
    a.x = v;
  else if (i == 1)
    a.y = v;
  else
    a.z = v;
}

struct Ray { // src/ray.h:14-16
  V o, d;
  double tm;
};
HD V at(const Ray &r, double t) { return r.o + t * r.d; } // src/ray.h:11

#define B200_PI 3.1415926535897932385 /* src/utility.h:15 */
#define B200_INF (__longlong_as_double(0x7ff0000000000000LL))

// ------------------------------------------------------------------------------------------------
// Random draws.  The reference consumes one global std::rand stream (src/utility.h:20); here every draw
// is a Philox4x32-10 output addressed by (seed; pixel, sample, bounce, purpose, aux), so a sample's value
// does not depend on scheduling, wave size or the number of GPUs.  Each 32-bit word maps to [0,1) with
// 2^-32 resolution (the reference's is 2^-31, never 1.0 either).
enum Purpose { P_CAMERA = 1, P_LENS = 2, P_MATERIAL = 3, P_SAMPLE = 4, P_VOLUME = 5 };

struct Draws {
  double u0, u1, u2, u3;
};

HD Draws philox4x32_10(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) {
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; r++) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c1 = (uint32_t)p1;
    c3 = (uint32_t)p0;
    c0 = n0;
    c2 = n2;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  const double s = 1.0 / 4294967296.0;
  return Draws{c0 * s, c1 * s, c2 * s, c3 * s};
}

struct Stream { // the path whose draws are being consumed
  uint64_t seed;
  uint32_t pixel, sample;
  HD Draws get(int bounce, int purpose, uint32_t aux) const {
    return philox4x32_10(seed, pixel, sample, ((uint32_t)bounce << 8) | (uint32_t)purpose, aux);
  }
};

} // namespace b200
