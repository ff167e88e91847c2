// Temporally blocked z-march, grouped form (sm_100a): sweep_tb2m's data movement with the plane loop
// restructured the way sweep_rows2 is (profiles/r01_c3_tb2m_stall_breakdown.md):
//   * the march advances THREE planes per loop iteration, fully unrolled: the three-deep z windows of
//     both levels rotate through statically named register blocks (no per-plane register moves:
//     ~50 of tb2m's 308 warp instructions per plane);
//   * phase A of the three planes, ONE consumer barrier, phase B of the three planes, one barrier: two
//     barriers per three planes instead of three (the level-1 ring holds five planes);
//   * level-1 centre blocks are re-read from shared memory in phase B instead of being held in
//     registers across phase A and the barrier;
//   * groups in the steady state of an item (no prologue / tail plane) run a copy of the group code
//     without the per-plane j tests.
// Geometry, TMA producer, ring protocol, masks and arithmetic are those of sweep_tb2m.cuh: the result is
// bit-identical to two single sweeps.  Single-input bodies only.
#pragma once
#include "sweep_tb2m.cuh"

namespace neptune {

template <class F, class C> struct Tb2gGeom : Tb2mGeom<F, C> {
  typedef Tb2mGeom<F, C> B;
  static constexpr int NL1 = 5;
  static constexpr int SMEM_BYTES = B::GUARD_BYTES + B::NS * B::SLOT_BYTES + B::GUARD_BYTES + B::L1_GUARD + NL1 * B::L1_BYTES +
                                    B::L1_GUARD + 2 * B::NS * 8 + 128;
  static_assert(F::NIN == 1, "grouped form: single-input bodies");
  static_assert(SMEM_BYTES <= 227 * 1024, "rings do not fit in shared memory");
};

template <class F, class C, bool FMA>
__global__ void __launch_bounds__((Tb2gGeom<F, C>::THREADS), C::MINB)
sweep_tb2g(const __grid_constant__ KParams<F> p, const __grid_constant__ CUtensorMap tmap) {
  typedef Tb2gGeom<F, C> G;
  typedef typename F::T T;
  constexpr int VX = G::VX, VY = G::VY, NS = G::NS;
  constexpr int NCW = G::CONSUMERS / 32;

  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char *smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  unsigned char *l0 = smem + G::GUARD_BYTES;                                          // NS level-0 stages
  unsigned char *l1 = l0 + NS * G::SLOT_BYTES + G::GUARD_BYTES + G::L1_GUARD;          // 5 level-1 planes
  const uint32_t full0 = smem_u32(l1 + G::NL1 * G::L1_BYTES + G::L1_GUARD);
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
    zb = p.wlo[0] + chunk * p.zchunk;
    ze = min(p.whi[0], zb + p.zchunk);
    y0 = (p.cyo + cy) * C::TY;
    x0 = (p.cxo + cx) * C::TX;
  };

  if (tid >= G::CONSUMERS) {
    // ---- producer: level-0 plane tiles zb-2 .. ze+1 of every item (as sweep_tb2m) ------------------
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
          tma_load_3d(stage0 + s * G::SLOT_BYTES, &tmap, x0 - 2 * VX, y0 - 2, zb - 2 + j, full0 + 8 * s);
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
  const bool swap = ((tid >> 3) & 1) != 0;
  const T *const ring0 = reinterpret_cast<const T *>(l0);
  T *const ring1 = reinterpret_cast<T *>(l1);
  constexpr int SLOT_ELEMS = G::SLOT_BYTES / (int)sizeof(T);
  constexpr int L1_ELEMS = G::L1_BYTES / (int)sizeof(T);
  const int soff0 = (by * VY - VY + 2) * G::P0 + VX + bx * VX;
  const int soff1 = (by * VY) * G::P1 + bx * VX;
  unsigned s = 0, ph = 0;   // level-0 ring slot / phase of the next plane to arrive
  unsigned c1 = 0;          // level-1 ring slot (0..4) of the next plane phase A produces
  // z windows: a[R], a[R+1], a[R+2] = level-0 centre blocks of planes k-2, k-1, k at step R of a group;
  // b[R], b[R+1], b[R+2] = level-1 centre blocks of planes k-3, k-2, k-1
  T a[5][VY][VX], b[5][VY][VX];
#pragma unroll
  for (int q = 0; q < 5; ++q)
#pragma unroll
    for (int iy = 0; iy < VY; ++iy)
#pragma unroll
      for (int ix = 0; ix < VX; ++ix) a[q][iy][ix] = b[q][iy][ix] = T(0);
  T nopw[1][VY][VX];
#pragma unroll
  for (int iy = 0; iy < VY; ++iy)
#pragma unroll
    for (int ix = 0; ix < VX; ++ix) nopw[0][iy][ix] = T(0);

  for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
    int zb, ze, y0, x0;
    geom(item, zb, ze, y0, x0);
    const int gx = x0 - VX + bx * VX, gy = y0 - VY + by * VY;
    unsigned bmask = 0, wmask = 0;
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
    const T *prv = ring0 + soff0;
    unsigned s_prv = 0;

    // one group of up to three arriving planes j0, j0+1, j0+2; STEADY: all three exist and j0 >= 4
    for (int j0 = 0; j0 < nslots; j0 += 3) {
      const bool steady = j0 >= 4 && j0 + 3 <= nslots;
      // two copies of the group code, each instantiated and entered from exactly one place (a closure
      // called from two sites is not inlined, and then the register windows live in local memory)
      static_for<2>([&](auto steadyc) {
      constexpr bool STEADY = decltype(steadyc)::value == 1;
      if (steady != STEADY) return;
      const unsigned cA = c1;   // level-1 planes produced before this group
      // ---- phase A ---------------------------------------------------------------------------------
      static_for<3>([&](auto rc) {
        constexpr int R = decltype(rc)::value;
        const int j = j0 + R;
        if (STEADY || j < nslots) {
          const int k = zb - 2 + j;
          mbar_wait(full0 + 8 * s, ph);
          const T *cur = ring0 + s * SLOT_ELEMS + soff0;
          T wa[VY + 2][VX + 2];
          const bool do_a = STEADY || j >= 2;
          // (threads beyond the thread grid run as copies of thread 0: same reads, same values stored to
          // the same level-1 addresses, no global writes — cheaper than a divergent region around the loads)
#pragma unroll
          for (int r = 0; r < VY; ++r) lds_vec<T, VX>(a[R + 2][r], cur + r * G::P0);
          if (do_a) load_halo<T, VX, VY, G::P0>(wa, prv, swap);
          __syncwarp();
          if (lane0) {
            if (STEADY || j >= 1) mbar_arrive(empty0 + 8 * s_prv);
            if (j == nslots - 1) mbar_arrive(empty0 + 8 * s);
          }
          s_prv = s;
          prv = cur;
          if (++s == NS) { s = 0; ph ^= 1; }
          if (do_a) {
            {
#pragma unroll
              for (int iy = 0; iy < VY; ++iy)
#pragma unroll
                for (int ix = 0; ix < VX; ++ix) wa[1 + iy][1 + ix] = a[R + 1][iy][ix];
              const unsigned m = (k - 1 >= p.blo[0] && k - 1 < p.bhi[0]) ? bmask : 0u;
              T *l1p = ring1 + c1 * L1_ELEMS + soff1;
              static_for<VY>([&](auto iyc) {
                constexpr int IY = decltype(iyc)::value;
                T ra[VX];
                static_for<VX>([&](auto ixc) {
                  constexpr int IX = decltype(ixc)::value;
                  MarchAcc<F, VY, VX, IY, IX> acc{a[R], wa, a[R + 2], nopw};
                  const T v = F::template eval<FMA>(acc, p.c);
                  ra[IX] = ((m >> (IY * VX + IX)) & 1u) ? v : a[R + 1][IY][IX];
                });
                stg_vec<T, VX, false>(l1p + IY * G::P1, ra);
              });
            }
            if (++c1 == 5u) c1 = 0;
          }
        }
      });
      // this group's level-1 planes are visible to the neighbours
      asm volatile("bar.sync 1, %0;" ::"n"(G::CONSUMERS) : "memory");
      // ---- phase B ---------------------------------------------------------------------------------
      // the level-1 window does not live in registers across phase A: its two oldest planes (produced by the
      // previous group) are re-read from the five-plane ring (own blocks only, so no halo is needed of them)
      auto slot = [](unsigned base, unsigned d) { const unsigned v = base + d; return v >= 5u ? v - 5u : v; };
      if (wmask != 0) {
#pragma unroll
        for (int r = 0; r < VY; ++r) {
          lds_vec<T, VX>(b[0][r], ring1 + slot(cA, 3u) * L1_ELEMS + soff1 + r * G::P1);
          lds_vec<T, VX>(b[1][r], ring1 + slot(cA, 4u) * L1_ELEMS + soff1 + r * G::P1);
        }
      }
      unsigned cB = cA;   // ring slot of the level-1 plane this step pushes = the one phase A produced at the same step
      static_for<3>([&](auto rc) {
        constexpr int R = decltype(rc)::value;
        const int j = j0 + R;
        if ((STEADY || j < nslots) && (STEADY || j >= 2)) {
          const int k = zb - 2 + j;
          // keep ptxas from hoisting the shared-memory loads of all three steps to the top of the phase
          // (it does, for latency, and then spills: 364 bytes at 96 registers)
          if (R > 0) __syncwarp();
          const T *pnew = ring1 + cB * L1_ELEMS + soff1;                 // level-1 plane k-1 (this thread's block)
          const T *pmid = ring1 + slot(cB, 4u) * L1_ELEMS + soff1;       // level-1 plane k-2
          cB = slot(cB, 1u);
          if (wmask != 0) {
#pragma unroll
            for (int r = 0; r < VY; ++r) lds_vec<T, VX>(b[R + 2][r], pnew + r * G::P1);
            if (STEADY || j >= 4) {
              T wb[VY + 2][VX + 2];
              load_halo<T, VX, VY, G::P1>(wb, pmid, swap);
#pragma unroll
              for (int iy = 0; iy < VY; ++iy)
#pragma unroll
                for (int ix = 0; ix < VX; ++ix) wb[1 + iy][1 + ix] = b[R + 1][iy][ix];
              const unsigned m = (k - 2 >= p.blo[0] && k - 2 < p.bhi[0]) ? bmask : 0u;
              static_for<VY>([&](auto iyc) {
                constexpr int IY = decltype(iyc)::value;
                T res[VX];
                static_for<VX>([&](auto ixc) {
                  constexpr int IX = decltype(ixc)::value;
                  MarchAcc<F, VY, VX, IY, IX> acc{b[R], wb, b[R + 2], nopw};
                  const T v = F::template eval<FMA>(acc, p.c);
                  res[IX] = ((m >> (IY * VX + IX)) & 1u) ? v : b[R + 1][IY][IX];
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
          }
          if (STEADY || j >= 4) dst += p.out_sz;
        }
      });
      // the level-1 plane read as `pmid` at step 0 is the next group's first write: everybody must be done
      asm volatile("bar.sync 1, %0;" ::"n"(G::CONSUMERS) : "memory");
      // rotate the windows by three planes (the only register moves of the group)
#pragma unroll
      for (int q = 0; q < 2; ++q)
#pragma unroll
        for (int iy = 0; iy < VY; ++iy)
#pragma unroll
          for (int ix = 0; ix < VX; ++ix) a[q][iy][ix] = a[q + 3][iy][ix];
      });
    }
  }
}

}  // namespace neptune
