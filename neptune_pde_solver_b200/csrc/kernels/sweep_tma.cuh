// TMA-staged persistent sweep kernel for sm_100a (the fast path of `neptune_ir.apply`).
//
// Data movement
//   * the one input that is read at neighbour offsets (the "staged" input) is brought into shared
//     memory by the Tensor Memory Accelerator: one cp.async.bulk.tensor.3d per plane-tile, a box of
//     (TX + 2*HX) x (TY + 2*RY) x 1 elements including its x/y halo, completion signalled on an
//     mbarrier (no thread ever issues a global load for it, out-of-box halo elements are
//     zero-filled by the TMA unit and never feed an in-bounds point);
//   * a ring of NS such stages, filled by a dedicated producer warp (one elected thread) and
//     handed back by the consumer warps through full[]/empty[] mbarriers — no CTA-wide barrier in
//     the steady state — lets the TMA unit run NS - 2*RZ - 1 planes ahead of the math;
//   * inputs read only at offset 0 (coefficient fields, right-hand sides) stream straight from
//     global memory with 128-bit read-only loads; results leave with 128-bit stores.
// Work decomposition
//   * the result box is cut into columns of TY x TX points; a column is cut into chunks of
//     `zchunk` planes; work item = (chunk, column), chunk-major, so that all columns of a chunk
//     are swept in lockstep by neighbouring CTAs and their shared halos are L2 hits;
//   * the grid is persistent: CTA b takes items b, b + gridDim.x, ...; its plane loads form ONE
//     continuous stream, so the ring never drains between items;
//   * every thread owns VY x VX points of the tile (VX * sizeof(T) = 16 bytes) and evaluates the
//     body functor on a register window read from the ring with 128-bit LDS.
// Semantics = ApplyToSCFForLowering (lib/Passes/DataflowLowering.cpp:258-448): out = copy of
// input 0, then body over `bounds`; here both in one pass (points outside `bounds` store the
// staged centre value, which IS input 0 — the host only selects this kernel when input 0 is the
// staged input).
#pragma once
#ifdef __CUDACC_RTC__
struct alignas(64) CUtensorMap_st { unsigned long long opaque[16]; };   // <cuda.h>'s definition; NVRTC has no host headers
typedef CUtensorMap_st CUtensorMap;
#else
#include <cuda.h>
#endif

#include "common.cuh"

namespace neptune {

template <int TX_, int TY_, int VY_, int NS_, int MINB_, bool STREAM_ = false, bool GENERIC_ = false> struct TileCfg {
  static constexpr int TX = TX_, TY = TY_, VY = VY_, NS = NS_, MINB = MINB_;
  static constexpr bool STREAM = STREAM_;    // results leave with st.global.cs (evict-first in L2)
  static constexpr bool GENERIC = GENERIC_;  // ring read through generic pointers (LD.E) instead of LDS: see the kernel
};

template <class F, class C> struct TmaGeom {
  typedef typename F::T T;
  static constexpr int S = sizeof(T);
  static constexpr int VX = 16 / S;
  static constexpr int RZ = F::RZ, RY = F::RY, RX = F::RX;
  static constexpr int HX = RX > 0 ? ((RX + VX - 1) / VX) * VX : 0;  // x halo, kept 16-byte aligned
  static constexpr int SXP = C::TX + 2 * HX;                        // stage row pitch (elements)
  static constexpr int SY = C::TY + 2 * RY;                         // stage rows
  static constexpr int STAGE_TX_BYTES = SXP * SY * S;               // bytes one TMA box delivers
  static constexpr int STAGE_BYTES = (STAGE_TX_BYTES + 127) / 128 * 128;
  static constexpr int NTX = C::TX / VX, NTY = C::TY / C::VY;
  static constexpr int CONSUMERS = NTX * NTY;        // math threads
  static constexpr int THREADS = CONSUMERS + 32;     // + one producer warp
  static constexpr int NS = C::NS;
  static constexpr int SMEM_BYTES = NS * STAGE_BYTES + 2 * NS * 8 + 128;  // + full/empty barriers + alignment slack
  static_assert(NS >= 2 * RZ + 2, "ring must hold the z window plus one plane in flight");
  static_assert(C::TX % VX == 0 && C::TY % C::VY == 0, "tile must be a multiple of the thread tile");
  static_assert(SXP <= 256 && SY <= 256, "TMA box dimensions are limited to 256 elements");
  static_assert(CONSUMERS % 32 == 0 && THREADS <= 1024, "bad thread count");
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(done)
      : "r"(bar), "r"(parity)
      : "memory");
  return done != 0;
}
// A wait that cannot hang the GPU: a lost TMA completion must fail loudly.  The bound is wall-clock time
// (%globaltimer, 20 s), not a poll count, so a slow but legitimate completion — compute-sanitizer, an ncu replay,
// a debugger, an oversubscribed GPU — is waited for; the trap takes the context down (sticky error) and is the
// last resort only.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  unsigned long long t0 = 0;
  for (unsigned polls = 1;; ++polls) {
    if (mbar_try_wait(bar, parity)) return;
    if ((polls & 0x3ffu) == 0) {
      unsigned long long now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now;
      else if (now - t0 > 20000000000ull) __trap();
      __nanosleep(64);
    }
  }
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, int x, int y, int z, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(dst), "l"((uint64_t)map), "r"(x), "r"(y), "r"(z), "r"(bar)
      : "memory");
}

__device__ __forceinline__ void tma_prefetch_3d(const CUtensorMap *map, int x, int y, int z) {
  asm volatile("cp.async.bulk.prefetch.tensor.3d.L2.global.tile [%0, {%1, %2, %3}];" ::"l"((uint64_t)map), "r"(x), "r"(y), "r"(z) : "memory");
}

template <class T> struct Vec16;
template <> struct Vec16<double> { typedef double2 type; };
template <> struct Vec16<float> { typedef float4 type; };

template <class T, int VX> __device__ __forceinline__ void lds_vec(T (&dst)[VX], const T *src) {
  typedef typename Vec16<T>::type V;
  union { V v; T e[VX]; } u;
  u.v = *reinterpret_cast<const V *>(src);
#pragma unroll
  for (int i = 0; i < VX; ++i) dst[i] = u.e[i];
}
template <class T, int VX> __device__ __forceinline__ void ldg_vec(T (&dst)[VX], const T *src) {
  typedef typename Vec16<T>::type V;
  union { V v; T e[VX]; } u;
  u.v = __ldg(reinterpret_cast<const V *>(src));
#pragma unroll
  for (int i = 0; i < VX; ++i) dst[i] = u.e[i];
}
template <class T, int VX, bool STREAM> __device__ __forceinline__ void stg_vec(T *dst, const T (&src)[VX]) {
  typedef typename Vec16<T>::type V;
  union { V v; T e[VX]; } u;
#pragma unroll
  for (int i = 0; i < VX; ++i) u.e[i] = src[i];
  if constexpr (STREAM) __stcs(reinterpret_cast<V *>(dst), u.v);
  else *reinterpret_cast<V *>(dst) = u.v;
}

// register window accessor for thread-tile point (IY, IX)
template <class F, class G, int VY, int IY, int IX> struct WinAcc {
  typedef typename F::T T;
  static constexpr int NPW = F::NIN > 1 ? F::NIN - 1 : 1;
  const T (&w)[2 * G::RZ + 1][VY + 2 * G::RY][G::VX + 2 * G::RX];
  const T (&pw)[NPW][VY][G::VX];
  template <int K, int DZ, int DY, int DX> __device__ __forceinline__ T tap() const {
    if constexpr (K == F::STAGED) {
      return w[DZ + G::RZ][IY + DY + G::RY][IX + DX + G::RX];
    } else {
      static_assert(DZ == 0 && DY == 0 && DX == 0, "non-staged inputs are read at offset 0 only");
      return pw[K < F::STAGED ? K : K - 1][IY][IX];
    }
  }
};

// the same with the pointwise inputs of one thread-tile row only (many-input bodies load them row by row)
template <class F, class G, int VY, int IY, int IX> struct WinAccRow {
  typedef typename F::T T;
  static constexpr int NPW = F::NIN > 1 ? F::NIN - 1 : 1;
  const T (&w)[2 * G::RZ + 1][VY + 2 * G::RY][G::VX + 2 * G::RX];
  const T (&pw)[NPW][G::VX];
  template <int K, int DZ, int DY, int DX> __device__ __forceinline__ T tap() const {
    if constexpr (K == F::STAGED) {
      return w[DZ + G::RZ][IY + DY + G::RY][IX + DX + G::RX];
    } else {
      static_assert(DZ == 0 && DY == 0 && DX == 0, "non-staged inputs are read at offset 0 only");
      return pw[K < F::STAGED ? K : K - 1][IX];
    }
  }
};

template <class F, class C, bool FMA>
__global__ void __launch_bounds__((TmaGeom<F, C>::THREADS), C::MINB)
sweep_tma(const __grid_constant__ KParams<F> p, const __grid_constant__ CUtensorMap tmap) {
  typedef TmaGeom<F, C> G;
  typedef typename F::T T;
  constexpr int RZ = G::RZ, RY = G::RY, RX = G::RX, HX = G::HX, VX = G::VX, VY = C::VY, NS = G::NS;
  constexpr int NPW = F::NIN > 1 ? F::NIN - 1 : 1;
  constexpr int NCW = G::CONSUMERS / 32;  // consumer warps
  constexpr bool LATE_PW = F::NIN > 4;

  extern __shared__ __align__(128) unsigned char smem_raw[];
  // (pointer arithmetic, not a round trip through an integer: the latter loses the shared address space and every
  // window read became a generic LD.E instead of LDS — r02 SASS census)
  unsigned char *smem = C::GENERIC ? reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127))
                                   : smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  const uint32_t full0 = smem_u32(smem + NS * G::STAGE_BYTES);  // full[s]: TMA bytes of stage s have landed
  const uint32_t empty0 = full0 + 8 * NS;                       // empty[s]: every consumer warp is done with stage s
  const uint32_t stage0 = smem_u32(smem);

  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int s = 0; s < NS; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, NCW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // work item -> geometry
  auto geom = [&](int item, int &zb, int &ze, int &y0, int &x0) {
    const int ncols = p.ncx * p.ncy;
    const int chunk = item / ncols, col = item - chunk * ncols;
    const int cx = col / p.ncy, cy = col - cx * p.ncy;
    zb = p.wlo[0] + chunk * p.zchunk + (chunk ? p.zgap : 0);
    ze = min(p.whi[0], zb + p.zchunk);
    y0 = (p.cyo + cy) * C::TY;
    x0 = (p.cxo + cx) * C::TX;
  };

  if (tid >= G::CONSUMERS) {
    // ---- producer warp: one elected thread streams plane-tiles into the ring, gated only by the
    // empty barriers, so loads run up to NS planes ahead of the math and never wait for a CTA-wide
    // barrier ----------------------------------------------------------------------------------------
    if (tid == G::CONSUMERS) {
      unsigned n = 0;
      for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
        int zb, ze, y0, x0;
        geom(item, zb, ze, y0, x0);
        const int nslots = ze - zb + 2 * RZ;
        for (int j = 0; j < nslots; ++j, ++n) {
          const unsigned s = n % NS;
          mbar_wait(empty0 + 8 * s, ((n / NS) & 1) ^ 1);  // fresh barrier: the wait on the previous phase passes
          mbar_expect_tx(full0 + 8 * s, G::STAGE_TX_BYTES);
          tma_load_3d(stage0 + s * G::STAGE_BYTES, &tmap, x0 - HX, y0 - RY, zb - RZ + j, full0 + 8 * s);
        }
      }
    }
    return;
  }

  // ---- consumer warps ---------------------------------------------------------------------------
  const int tx = tid % G::NTX, ty = tid / G::NTX;
  const int lx = tx * VX, ly = ty * VY;  // tile-local origin of this thread's points
  const bool lane0 = (tid & 31) == 0;
  unsigned n = 0;  // slots consumed so far
  for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
    int zb, ze, y0, x0;
    geom(item, zb, ze, y0, x0);
    const int gx = x0 + lx, gy = y0 + ly;
    const bool live = gx < p.whi[2] && gy < p.whi[1] && gx + VX > p.wlo[2] && gy + VY > p.wlo[1];
    const int nslots = ze - zb + 2 * RZ;
    for (int j = 0; j < nslots; ++j, ++n) {
      const bool compute = j >= 2 * RZ;
      const int z = zb + j - 2 * RZ;

      // pointwise inputs: 128-bit streaming loads.  Few inputs: issued before the wait so they overlap it.
      // Many inputs (one coefficient field per tap): issued right before the body instead, in the same
      // basic block, so that ptxas can interleave loads and arithmetic within the register budget of two
      // CTAs per SM (r01 ncu of the 27-field body: all 27 vectors preloaded = 168 registers, one CTA of 9
      // warps per SM, DRAM at 4.4 TB/s waiting on `long_scoreboard`).
      T pw[NPW][VY][VX];
      auto load_pw = [&]() {
        if (compute && live) {
          static_for<F::NIN>([&](auto kc) {
            constexpr int K = decltype(kc)::value;
            if constexpr (K != F::STAGED) {
              constexpr int PK = K < F::STAGED ? K : K - 1;
#pragma unroll
              for (int iy = 0; iy < VY; ++iy) {
                if (gy + iy < p.ny) {
                  const T *src = p.in[K] + (long long)z * p.in_sz[K] + (long long)(gy + iy) * p.in_sy[K] + gx;
                  if (gx + VX <= p.nx) {
                    ldg_vec<T, VX>(pw[PK][iy], src);
                  } else {
#pragma unroll
                    for (int ix = 0; ix < VX; ++ix) pw[PK][iy][ix] = (gx + ix < p.nx) ? __ldg(src + ix) : T(0);
                  }
                }
              }
            }
          });
        }
      };
      if constexpr (F::NIN > 1 && !LATE_PW) load_pw();

      mbar_wait(full0 + 8 * (n % NS), (n / NS) & 1);
      if (!compute) continue;

      // register window: planes z-RZ..z+RZ are slots n-2RZ..n
      T w[2 * RZ + 1][VY + 2 * RY][VX + 2 * RX];
      if (live) {
#pragma unroll
        for (int dz = 0; dz <= 2 * RZ; ++dz) {
          const unsigned s = (n - 2 * RZ + dz) % NS;
          const T *sp = reinterpret_cast<const T *>(smem + s * G::STAGE_BYTES);
#pragma unroll
          for (int r = 0; r < VY + 2 * RY; ++r) {
            const T *row = sp + (ly + r) * G::SXP + HX + lx;  // (ly + r - RY) + RY rows down
            T ctr[VX];
            lds_vec<T, VX>(ctr, row);
#pragma unroll
            for (int i = 0; i < VX; ++i) w[dz][r][RX + i] = ctr[i];
#pragma unroll
            for (int i = 0; i < RX; ++i) {
              w[dz][r][i] = row[i - RX];
              w[dz][r][RX + VX + i] = row[VX + i];
            }
          }
        }
      }
      // this warp has its window in registers: hand the oldest stage (plane z-RZ) back to the
      // producer, and at the end of an item the 2RZ younger ones too (the next item restarts its z window)
      __syncwarp();
      if (lane0) {
        mbar_arrive(empty0 + 8 * ((n - 2 * RZ) % NS));
        if (j == nslots - 1) {
#pragma unroll
          for (int q = 1; q <= 2 * RZ; ++q) mbar_arrive(empty0 + 8 * ((n - 2 * RZ + q) % NS));
        }
      }

      if constexpr (LATE_PW) {
        if (live) {
          const bool zin = z >= p.blo[0] && z < p.bhi[0];
          const bool xin = gx >= p.blo[2] && gx + VX <= p.bhi[2] && gx >= p.wlo[2] && gx + VX <= p.whi[2];
          static_for<VY>([&](auto iyc) {
            constexpr int IY = decltype(iyc)::value;
            const int y = gy + IY;
            if (y < p.whi[1] && y >= p.wlo[1]) {
              T *dst = p.out + (long long)z * p.out_sz + (long long)y * p.out_sy + gx;
              if (zin && xin && y >= p.blo[1] && y < p.bhi[1]) {
                // whole vector inside the apply bounds (which lie inside every input's box): unconditional
                // vector loads and the body in ONE basic block
                T pr[NPW][VX];
                static_for<F::NIN>([&](auto kc) {
                  constexpr int K = decltype(kc)::value;
                  if constexpr (K != F::STAGED)
                    ldg_vec<T, VX>(pr[K < F::STAGED ? K : K - 1],
                                   p.in[K] + (long long)z * p.in_sz[K] + (long long)y * p.in_sy[K] + gx);
                });
                T res[VX];
                static_for<VX>([&](auto ixc) {
                  constexpr int IX = decltype(ixc)::value;
                  WinAccRow<F, G, VY, IY, IX> acc{w, pr};
                  res[IX] = F::template eval<FMA>(acc, p.c);
                });
                stg_vec<T, VX, C::STREAM>(dst, res);
              } else {
                // edge of the bounds or of the written box: point by point
                static_for<VX>([&](auto ixc) {
                  constexpr int IX = decltype(ixc)::value;
                  const int x = gx + IX;
                  if (x >= p.wlo[2] && x < p.whi[2]) {
                    T v = w[RZ][IY + RY][IX + RX];
                    if (zin && y >= p.blo[1] && y < p.bhi[1] && x >= p.blo[2] && x < p.bhi[2]) {
                      T pr[NPW][VX];
                      static_for<F::NIN>([&](auto kc) {
                        constexpr int K = decltype(kc)::value;
                        if constexpr (K != F::STAGED)
                          pr[K < F::STAGED ? K : K - 1][IX] =
                              __ldg(p.in[K] + (long long)z * p.in_sz[K] + (long long)y * p.in_sy[K] + x);
                      });
                      WinAccRow<F, G, VY, IY, IX> acc{w, pr};
                      v = F::template eval<FMA>(acc, p.c);
                    }
                    dst[IX] = v;
                  }
                });
              }
            }
          });
        }
        continue;
      }
      if (live) {
        const bool zin = z >= p.blo[0] && z < p.bhi[0];
        static_for<VY>([&](auto iyc) {
          constexpr int IY = decltype(iyc)::value;
          const int y = gy + IY;
          if (y < p.whi[1] && y >= p.wlo[1]) {
            const bool yin = zin && y >= p.blo[1] && y < p.bhi[1];
            T res[VX];
            static_for<VX>([&](auto ixc) {
              constexpr int IX = decltype(ixc)::value;
              const int x = gx + IX;
              T v = w[RZ][IY + RY][IX + RX];  // copy-through value: input 0 at this point
              if (yin && x >= p.blo[2] && x < p.bhi[2]) {
                WinAcc<F, G, VY, IY, IX> acc{w, pw};
                v = F::template eval<FMA>(acc, p.c);
              }
              res[IX] = v;
            });
            T *dst = p.out + (long long)z * p.out_sz + (long long)y * p.out_sy + gx;
            if (gx >= p.wlo[2] && gx + VX <= p.whi[2]) {
              stg_vec<T, VX, C::STREAM>(dst, res);
            } else {
#pragma unroll
              for (int ix = 0; ix < VX; ++ix)
                if (gx + ix >= p.wlo[2] && gx + ix < p.whi[2]) dst[ix] = res[ix];
            }
          }
        });
      }
    }
  }
}

}  // namespace neptune
