// Shared device-side definitions of the B200 stencil-sweep kernels.
//
// A *body functor* F (generated from the apply region by frontend/codegen.cpp, one per body
// structure) exposes
//     typedef T;  RANK, NIN, NCONST, STAGED, RZ, RY, RX, FLOPS
//     template <bool FMA, class Acc> static T eval(const Acc &u, const T *c);
// where `u.template tap<K, DZ, DY, DX>()` is the value of input K at the current point plus the
// canonical (z, y, x) offset and c[] are the body's constants in order of appearance.  eval()
// is the body's op list verbatim — the reference clones the same ops into its scf.for nest
// (lib/Passes/DataflowLowering.cpp:413-423) — so with FMA = false the result is bit-identical
// to the CPU lowering.
#pragma once
#ifdef __CUDACC_RTC__
// NVRTC (runtime/jit.cpp compiles bodies outside the prebuilt catalogue): no host headers
typedef unsigned char uint8_t;
typedef unsigned int uint32_t;
typedef int int32_t;
typedef unsigned long long uint64_t;
typedef long long int64_t;
typedef unsigned long long uintptr_t;
namespace std {
template <class T, T V> struct integral_constant {
  static constexpr T value = V;
  typedef T value_type;
  constexpr operator T() const noexcept { return V; }
};
}  // namespace std
#else
#include <cuda_runtime.h>
#include <stdint.h>
#include <type_traits>
#endif

namespace neptune {

// IEEE round-to-nearest ops that ptxas may never contract (FMA = false), or plain operators
// that it may (FMA = true, tolerance tier).
template <class T, bool FMA> struct Arith;
template <> struct Arith<double, false> {
  static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
  static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
  static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
  static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
  static __device__ __forceinline__ double neg(double a) { return -a; }
};
template <> struct Arith<double, true> {
  static __device__ __forceinline__ double add(double a, double b) { return a + b; }
  static __device__ __forceinline__ double sub(double a, double b) { return a - b; }
  static __device__ __forceinline__ double mul(double a, double b) { return a * b; }
  static __device__ __forceinline__ double div(double a, double b) { return a / b; }
  static __device__ __forceinline__ double neg(double a) { return -a; }
};
template <> struct Arith<float, false> {
  static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
  static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
  static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
  static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
  static __device__ __forceinline__ float neg(float a) { return -a; }
};
template <> struct Arith<float, true> {
  static __device__ __forceinline__ float add(float a, float b) { return a + b; }
  static __device__ __forceinline__ float sub(float a, float b) { return a - b; }
  static __device__ __forceinline__ float mul(float a, float b) { return a * b; }
  static __device__ __forceinline__ float div(float a, float b) { return a / b; }
  static __device__ __forceinline__ float neg(float a) { return -a; }
};

// Kernel parameters, canonical (z, y, x) coordinates of the RESULT box's storage:
// rank 3 -> (d0, d1, d2); rank 2 -> (1, d0, d1); rank 1 -> (1, 1, d0).
// in[k] is pre-shifted by (out_lb - in_lb_k) so that the tap of input k at result point (z,y,x)
// plus offset (dz,dy,dx) is in[k][(z+dz)*in_sz[k] + (y+dy)*in_sy[k] + (x+dx)]
// (index algebra of DataflowLowering.cpp:381-411: iv + off - inLb).
template <class F> struct KParams {
  typedef typename F::T T;
  T *out;
  const T *copy_src;     // input 0, NOT shifted: copy-through is storage-index identity (memref.copy, :282-287)
  const T *in[F::NIN];
  long long in_sz[F::NIN], in_sy[F::NIN];
  long long out_sz, out_sy;
  int nz, ny, nx;        // result extents
  int blo[3], bhi[3];    // apply bounds, result storage coordinates
  int wlo[3], whi[3];    // box this launch writes (whole result box, or a slab of planes)
  // persistent-tile kernels only
  int zchunk, nchunks, ncx, ncy, cxo, cyo, nitems;
  int zgap;              // planes skipped between chunk 0 and chunk 1 (two boundary ranges in one launch), else 0
  T c[F::NCONST > 0 ? F::NCONST : 1];
};

template <int N, class Fn> __device__ __forceinline__ void static_for(Fn &&fn) {
  if constexpr (N > 0) {
    static_for<N - 1>(fn);
    fn(std::integral_constant<int, N - 1>{});
  }
}

}  // namespace neptune
