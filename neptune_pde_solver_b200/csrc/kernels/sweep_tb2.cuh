// Temporally blocked sweep for sm_100a: TWO time steps of a radius-1, single-input apply per launch.
//
//   u2 = apply(apply(u0))   with   u1 = apply(u0) never leaving the SM.
//
// The host time loop of the reference calls the compiled entry once per step and swaps the
// buffers; every call streams the whole field through memory.  Here a CTA streams
// level-0 plane tiles (u0) in with TMA exactly like sweep_tma.cuh, evaluates the body once to get a
// level-1 plane (u1) over the tile plus a one-point ring, keeps the last planes of u1 in shared
// memory, evaluates the body a second time on them and stores the level-2 plane (u2): one read and
// one write of the field per TWO steps.  Arithmetic is untouched — both evaluations are the body's
// own op list (functor F), and points outside `bounds` carry input 0 at both levels (copy-through,
// lib/Passes/DataflowLowering.cpp:282-287) — so the result is bit-identical to two single sweeps.
//
// Geometry (VX * sizeof(T) = 16 bytes, thread block = VY x VX points):
//   * the consumer threads form a grid of (TX/VX + 2) x (TY/VY + 2) blocks covering the output tile
//     plus one block all round; every thread computes its level-1 block (phase A); only the inner
//     TX/VX x TY/VY threads compute and store a level-2 block (phase B).  The outer half of the
//     ring blocks is redundant work whose inputs may be stale shared memory: never selected.
//   * level-0 stage = TMA box (TX + 4 VX) x (TY + 4) at (x0 - 2 VX, y0 - 2); level-1 planes live in a
//     ring of 4 so that one consumer barrier per plane orders both the RAW and the WAR hazard.
//   * z: an item of zc output planes consumes zc + 4 level-0 planes and produces zc + 2 level-1 planes.
#pragma once
#include "sweep_tma.cuh"

namespace neptune {

template <int TX_, int TY_, int VY_, int NS_, int MINB_> struct Tb2Cfg {
  static constexpr int TX = TX_, TY = TY_, VY = VY_, NS = NS_, MINB = MINB_;
};

template <class F, class C> struct Tb2Geom {
  typedef typename F::T T;
  static constexpr int S = sizeof(T);
  static constexpr int VX = 16 / S, VY = C::VY;
  static constexpr int NBX = C::TX / VX + 2, NBY = C::TY / VY + 2;  // thread-block grid
  static constexpr int ACTIVE = NBX * NBY;
  static constexpr int CONSUMERS = (ACTIVE + 31) / 32 * 32;
  static constexpr int THREADS = CONSUMERS + 32;
  static constexpr int NS = C::NS;
  // level 0
  static constexpr int P0 = C::TX + 4 * VX;        // stage row pitch = TMA box width
  static constexpr int BOX_Y = C::TY + 4;
  static constexpr int STAGE_TX_BYTES = P0 * BOX_Y * S;
  static constexpr int STAGE_BYTES = (STAGE_TX_BYTES + 127) / 128 * 128;
  static constexpr int GUARD_BYTES = (P0 * S * (VY > 2 ? VY - 1 : 1) + 127) / 128 * 128;  // rows read above/below a box
  // level 1
  static constexpr int P1 = NBX * VX, H1 = NBY * VY;
  static constexpr int L1_BYTES = (P1 * H1 * S + 127) / 128 * 128;
  static constexpr int NL1 = 4;
  static constexpr int SMEM_BYTES = GUARD_BYTES + NS * STAGE_BYTES + GUARD_BYTES + NL1 * L1_BYTES + 2 * NS * 8 + 128;
  static_assert(F::RZ == 1 && F::RY <= 1 && F::RX <= 1, "temporal blocking is built for radius-1 3-D bodies");
  static_assert(F::NIN == 1 && F::STAGED == 0, "temporal blocking needs a single input");
  static_assert(C::TX % VX == 0 && C::TY % VY == 0, "tile must be a multiple of the thread block");
  static_assert(P0 <= 256 && BOX_Y <= 256, "TMA box dimensions are limited to 256 elements");
  static_assert(THREADS <= 1024, "too many threads");
};

// accessor over a 3 x (VY+2) x (VX+2) register window, radius 1 in every direction
template <class F, int VY, int VX, int IY, int IX> struct Win1Acc {
  typedef typename F::T T;
  const T (&w)[3][VY + 2][VX + 2];
  template <int K, int DZ, int DY, int DX> __device__ __forceinline__ T tap() const {
    static_assert(K == 0, "single input");
    return w[DZ + 1][IY + DY + 1][IX + DX + 1];
  }
};

template <class T, int VX, int VY, int PITCH, class PL>
__device__ __forceinline__ void load_window(T (&w)[3][VY + 2][VX + 2], const PL &pl) {
#pragma unroll
  for (int dz = 0; dz < 3; ++dz) {
#pragma unroll
    for (int r = 0; r < VY + 2; ++r) {
      const T *row = pl[dz] + (r - 1) * PITCH;  // pl[] points at the thread block's first point
      T ctr[VX];
      lds_vec<T, VX>(ctr, row);
#pragma unroll
      for (int i = 0; i < VX; ++i) w[dz][r][1 + i] = ctr[i];
      w[dz][r][0] = row[-1];
      w[dz][r][VX + 1] = row[VX];
    }
  }
}

template <class F, class C, bool FMA>
__global__ void __launch_bounds__((Tb2Geom<F, C>::THREADS), C::MINB)
sweep_tb2(const __grid_constant__ KParams<F> p, const __grid_constant__ CUtensorMap tmap) {
  typedef Tb2Geom<F, C> G;
  typedef typename F::T T;
  constexpr int VX = G::VX, VY = G::VY, NS = G::NS, NL1 = G::NL1;
  constexpr int NCW = G::CONSUMERS / 32;

  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char *smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  unsigned char *l0 = smem + G::GUARD_BYTES;                              // NS level-0 stages
  unsigned char *l1 = l0 + NS * G::STAGE_BYTES + G::GUARD_BYTES;           // NL1 level-1 planes
  const uint32_t full0 = smem_u32(l1 + NL1 * G::L1_BYTES);
  const uint32_t empty0 = full0 + 8 * NS;
  const uint32_t stage0 = smem_u32(l0);

  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int s = 0; s < NS; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, NCW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

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
    // ---- producer: level-0 plane tiles zb-2 .. ze+1 of every item -------------------------------
    if (tid == G::CONSUMERS) {
      unsigned n = 0;
      for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
        int zb, ze, y0, x0;
        geom(item, zb, ze, y0, x0);
        const int nslots = ze - zb + 4;
        for (int j = 0; j < nslots; ++j, ++n) {
          const unsigned s = n % NS;
          mbar_wait(empty0 + 8 * s, ((n / NS) & 1) ^ 1);
          mbar_expect_tx(full0 + 8 * s, G::STAGE_TX_BYTES);
          tma_load_3d(stage0 + s * G::STAGE_BYTES, &tmap, x0 - 2 * VX, y0 - 2, zb - 2 + j, full0 + 8 * s);
        }
      }
    }
    return;
  }

  // ---- consumers ----------------------------------------------------------------------------------
  const bool active = tid < G::ACTIVE;
  const int bx = active ? tid % G::NBX : 0, by = active ? tid / G::NBX : 0;
  const bool inner = active && bx >= 1 && bx <= G::NBX - 2 && by >= 1 && by <= G::NBY - 2;
  const bool lane0 = (tid & 31) == 0;
  const T *const ring0 = reinterpret_cast<const T *>(l0);
  T *const ring1 = reinterpret_cast<T *>(l1);
  constexpr int STAGE_ELEMS = G::STAGE_BYTES / (int)sizeof(T);
  constexpr int L1_ELEMS = G::L1_BYTES / (int)sizeof(T);
  const int soff0 = (by * VY - VY + 2) * G::P0 + VX + bx * VX;   // thread block's first point inside a level-0 stage
  const int soff1 = (by * VY) * G::P1 + bx * VX;                 // ... inside a level-1 plane
  unsigned s = 0, ph = 0;   // level-0 ring slot / phase of the next plane to arrive
  unsigned c1 = 0;          // level-1 planes produced so far (slot = c1 % NL1)
  for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
    int zb, ze, y0, x0;
    geom(item, zb, ze, y0, x0);
    const int gx = x0 - VX + bx * VX, gy = y0 - VY + by * VY;  // global coords of the thread block
    unsigned bmask = 0, wmask = 0;  // bit iy*VX+ix: inside apply bounds (y,x) / inner thread and inside the written box
#pragma unroll
    for (int iy = 0; iy < VY; ++iy)
#pragma unroll
      for (int ix = 0; ix < VX; ++ix) {
        const int y = gy + iy, x = gx + ix;
        if (y >= p.blo[1] && y < p.bhi[1] && x >= p.blo[2] && x < p.bhi[2]) bmask |= 1u << (iy * VX + ix);
        if (inner && y >= p.wlo[1] && y < p.whi[1] && x >= p.wlo[2] && x < p.whi[2]) wmask |= 1u << (iy * VX + ix);
      }
    const int nslots = ze - zb + 4;
    T *dst = p.out + (long long)zb * p.out_sz + (long long)gy * p.out_sy + gx;
    const T *pl[3] = {ring0 + soff0, ring0 + soff0, ring0 + soff0};
    for (int j = 0; j < nslots; ++j) {
      const int k = zb - 2 + j;  // level-0 plane arriving now
      mbar_wait(full0 + 8 * s, ph);
      pl[0] = pl[1]; pl[1] = pl[2]; pl[2] = ring0 + s * STAGE_ELEMS + soff0;
      const unsigned s_old = (s + NS - 2) % NS;  // slot of plane k-2
      if (++s == NS) { s = 0; ph ^= 1; }
      if (j < 2) continue;

      // ---- phase A: level-1 plane k-1 over the whole thread grid ------------------------------
      T w[3][VY + 2][VX + 2];
      if (active) load_window<T, VX, VY, G::P0>(w, pl);
      __syncwarp();
      if (lane0) {
        mbar_arrive(empty0 + 8 * s_old);
        if (j == nslots - 1) {
          mbar_arrive(empty0 + 8 * ((s_old + 1) % NS));
          mbar_arrive(empty0 + 8 * ((s_old + 2) % NS));
        }
      }
      if (active) {
        const unsigned m = (k - 1 >= p.blo[0] && k - 1 < p.bhi[0]) ? bmask : 0u;
        T *l1p = ring1 + (c1 % NL1) * L1_ELEMS + soff1;
        static_for<VY>([&](auto iyc) {
          constexpr int IY = decltype(iyc)::value;
          T res[VX];
          static_for<VX>([&](auto ixc) {
            constexpr int IX = decltype(ixc)::value;
            Win1Acc<F, VY, VX, IY, IX> acc{w};
            const T v = F::template eval<FMA>(acc, p.c);
            res[IX] = ((m >> (IY * VX + IX)) & 1u) ? v : w[1][IY + 1][IX + 1];
          });
          stg_vec<T, VX, false>(l1p + IY * G::P1, res);  // shared-memory store of the level-1 block
        });
      }
      ++c1;
      // one consumer-wide barrier per plane: orders this plane's level-1 stores before the reads
      // below, and the reads of the previous iteration before the stores of the next (ring of 4)
      asm volatile("bar.sync 1, %0;" ::"n"(G::CONSUMERS) : "memory");
      if (j < 4) continue;

      // ---- phase B: level-2 plane k-2 from level-1 planes k-3, k-2, k-1 -------------------------
      if (wmask) {
        const T *q[3] = {ring1 + ((c1 + NL1 - 3) % NL1) * L1_ELEMS + soff1,
                         ring1 + ((c1 + NL1 - 2) % NL1) * L1_ELEMS + soff1,
                         ring1 + ((c1 + NL1 - 1) % NL1) * L1_ELEMS + soff1};
        load_window<T, VX, VY, G::P1>(w, q);
        const unsigned m = (k - 2 >= p.blo[0] && k - 2 < p.bhi[0]) ? bmask : 0u;
        static_for<VY>([&](auto iyc) {
          constexpr int IY = decltype(iyc)::value;
          T res[VX];
          static_for<VX>([&](auto ixc) {
            constexpr int IX = decltype(ixc)::value;
            Win1Acc<F, VY, VX, IY, IX> acc{w};
            const T v = F::template eval<FMA>(acc, p.c);
            res[IX] = ((m >> (IY * VX + IX)) & 1u) ? v : w[1][IY + 1][IX + 1];
          });
          constexpr unsigned ROW = ((1u << VX) - 1u) << (IY * VX);
          T *d = dst + IY * p.out_sy;
          if ((wmask & ROW) == ROW) {
            stg_vec<T, VX, false>(d, res);
          } else {
#pragma unroll
            for (int ix = 0; ix < VX; ++ix)
              if ((wmask >> (IY * VX + ix)) & 1u) d[ix] = res[ix];
          }
        });
      }
      dst += p.out_sz;
    }
  }
}

}  // namespace neptune
