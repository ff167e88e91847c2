// Universal sweep kernel: one thread per result point, every tap a read-only global load served
// by L1/L2.  Handles any rank (1..3), any tap radius, any number of inputs with their own lower
// bounds — it is the correctness anchor on the device and the path for bodies/geometries the
// TMA-staged kernel (sweep_tma.cuh) does not take.
//
// Semantics = ApplyToSCFForLowering (lib/Passes/DataflowLowering.cpp:258-448): points of the
// result box outside `bounds` take input 0's value (copy-through, :282-287), points inside
// evaluate the body and store the yielded scalar (:427-444).
#pragma once
#include "common.cuh"

namespace neptune {

template <class F> struct DirectAcc {
  typedef typename F::T T;
  const KParams<F> &p;
  int z, y, x;
  template <int K, int DZ, int DY, int DX> __device__ __forceinline__ T tap() const {
    return __ldg(p.in[K] + (long long)(z + DZ) * p.in_sz[K] + (long long)(y + DY) * p.in_sy[K] + (x + DX));
  }
};

constexpr int DIRECT_BX = 128, DIRECT_BY = 2;

template <class F, bool FMA>
__global__ void __launch_bounds__(DIRECT_BX *DIRECT_BY) sweep_direct(const __grid_constant__ KParams<F> p) {
  typedef typename F::T T;
  const int x = p.wlo[2] + blockIdx.x * DIRECT_BX + threadIdx.x;
  const int y = p.wlo[1] + blockIdx.y * DIRECT_BY + threadIdx.y;
  const int z = p.wlo[0] + blockIdx.z;
  if (x >= p.whi[2] || y >= p.whi[1]) return;
  const bool inb = z >= p.blo[0] && z < p.bhi[0] && y >= p.blo[1] && y < p.bhi[1] && x >= p.blo[2] && x < p.bhi[2];
  T v;
  if (inb) {
    DirectAcc<F> acc{p, z, y, x};
    v = F::template eval<FMA>(acc, p.c);
  } else {
    v = __ldg(p.copy_src + (long long)z * p.in_sz[0] + (long long)y * p.in_sy[0] + x);
  }
  p.out[(long long)z * p.out_sz + (long long)y * p.out_sy + x] = v;
}

}  // namespace neptune
