// host_emu.h — TEST INFRASTRUCTURE.  Minimal stand-ins for the CUDA built-ins used by the device FUNCTIONS of
// cpu-ray-tracing-implementation_b200/csrc (not the kernels), so that tests/emu/emu.cc can execute the very same
// traversal / intersection / shading source on the CPU of a box without a GPU.  Never part of libb200rt.so.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#define __host__
#define __device__
#define __forceinline__ inline
#define __noinline__
#define __align__(n) alignas(n)
struct alignas(16) double2 {
  double x, y;
};
struct alignas(8) int2 {
  int x, y;
};
struct alignas(16) int4 {
  int x, y, z, w;
};
template <class T> inline T __ldg(const T *p) { return *p; }
inline double __longlong_as_double(long long v) {
  double d;
  std::memcpy(&d, &v, 8);
  return d;
}
inline long long __double_as_longlong(double d) {
  long long v;
  std::memcpy(&v, &d, 8);
  return v;
}
using std::isnan;
struct alignas(16) float4 {
  float x, y, z, w;
};
inline float __fmaf_rn(float a, float b, float c) { return std::fmaf(a, b, c); }
inline float __int_as_float(int v) {
  float f;
  std::memcpy(&f, &v, 4);
  return f;
}
inline float __double2float_rd(double x) {
  float f = (float)x;
  if ((double)f > x) f = std::nextafterf(f, -INFINITY);
  return f;
}
inline float __double2float_ru(double x) {
  float f = (float)x;
  if ((double)f < x) f = std::nextafterf(f, INFINITY);
  return f;
}
// fabsf / fmaxf / fminf: the C library's global-namespace functions
inline double2 make_double2(double x, double y) { return double2{x, y}; }
inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
inline float __uint_as_float(unsigned v) {
  float f;
  std::memcpy(&f, &v, 4);
  return f;
}
