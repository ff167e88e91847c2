// Temporally blocked z-march, software-pipelined grouped form (sm_100a).
//
// sweep_tb2g (three planes per iteration, no register rotation, fewer barriers) lost to sweep_tb2m because
// it runs phase A and phase B as separate instruction streams: half the independent fp64 chains per warp
// (profiles/r01_c3_tb2m_stall_breakdown.md).  Here phase B of the PREVIOUS group rides in the same steps as
// phase A of the current one: step R of a group
//     A : level-0 plane k arrives  -> level-1 plane k-1 (stored to the level-1 ring)
//     B': level-2 plane k-LAG        <- level-1 planes k-6, k-5, k-4 (all produced before the last barrier)
// so both levels' chains interleave in one stream as in tb2m, with ONE consumer barrier per three planes and
// no per-plane register moves.  The level-1 ring holds seven planes (the three being written, the four
// being read); an item runs three extra steps (B' only) to drain.  Geometry, TMA producer, masks and
// arithmetic are those of sweep_tb2m.cuh: bit-identical to two single sweeps.  Single-input bodies only.
#pragma once
#include "sweep_tb2m.cuh"

namespace neptune {

// GS = planes per barrier group (2 or 3).  Level-1 ring = 2*GS + 1 planes (GS being written, GS + 1 being read);
// phase B' lags phase A by GS + 1 planes.  GS = 3 needs no register rotation with a 3-step loop body; GS = 2
// unrolls six steps (three groups) for the same effect and leaves shared memory for two more level-0 TMA slots
// — but ptxas spills 2.4 KB on the six-step body at 96 registers (r01), so only GS = 3 is registered.
template <int TX_, int TY_, int VY_, int NS_, int MINB_, int GS_ = 3> struct Tb2pCfg : Tb2Cfg<TX_, TY_, VY_, NS_, MINB_> {
  static constexpr int GS = GS_;
};

template <class F, class C> struct Tb2pGeom : Tb2mGeom<F, C> {
  typedef Tb2mGeom<F, C> B;
  static constexpr int GS = C::GS;
  static constexpr int NL1 = 2 * GS + 1;
  static_assert(GS == 2 || GS == 3, "two or three planes per barrier group");
  static constexpr int SMEM_BYTES = B::GUARD_BYTES + B::NS * B::SLOT_BYTES + B::GUARD_BYTES + B::L1_GUARD + NL1 * B::L1_BYTES +
                                    B::L1_GUARD + 2 * B::NS * 8 + 128;
  static_assert(F::NIN == 1, "pipelined grouped form: single-input bodies");
  static_assert(B::NS >= 3, "ring too short");
  static_assert(SMEM_BYTES <= 227 * 1024, "rings do not fit in shared memory");
};

template <class F, class C, bool FMA>
__global__ void __launch_bounds__((Tb2pGeom<F, C>::THREADS), C::MINB)
sweep_tb2p(const __grid_constant__ KParams<F> p, const __grid_constant__ CUtensorMap tmap) {
  typedef Tb2pGeom<F, C> G;
  typedef typename F::T T;
  constexpr int VX = G::VX, VY = G::VY, NS = G::NS;
  constexpr int NCW = G::CONSUMERS / 32;
  constexpr int GS = G::GS, NL1 = G::NL1;
  constexpr int U = GS == 3 ? 3 : 6;     // steps per loop iteration: a multiple of GS and of the 3-deep z windows
  constexpr int LAG = GS + 2;            // phase B' of a step computes level-2 plane k - LAG

  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char *smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  unsigned char *l0 = smem + G::GUARD_BYTES;                                          // NS level-0 stages
  unsigned char *l1 = l0 + NS * G::SLOT_BYTES + G::GUARD_BYTES + G::L1_GUARD;          // 2*GS + 1 level-1 planes
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
  unsigned c1 = 0;          // level-1 ring slot (0..6) phase A writes at the current step (advances every step with j >= 2)
  // z windows: a[R], a[R+1], a[R+2] = level-0 centre blocks of planes k-2, k-1, k at step R of a group;
  // b[R], b[R+1], b[R+2] = level-1 centre blocks of planes k-LAG-1, k-LAG, k-LAG+1
  T a[U + 2][VY][VX], b[U + 2][VY][VX];
#pragma unroll
  for (int q = 0; q < U + 2; ++q)
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

    // one group of three steps j0, j0+1, j0+2; an item has nslots arrivals and three drain steps.
    // STEADY: every step has an arrival and j0 >= 7 (both phases live, no tests on j)
    for (int j0 = 0; j0 < nslots + GS; j0 += U) {
      const bool steady = j0 >= LAG + 2 && j0 + U <= nslots;
      static_for<2>([&](auto steadyc) {
      constexpr bool STEADY = decltype(steadyc)::value == 1;
      if (steady != STEADY) return;
      auto slot = [](unsigned base, unsigned d) { const unsigned v = base + d; return v >= (unsigned)NL1 ? v - (unsigned)NL1 : v; };
      static_for<U>([&](auto rc) {
        constexpr int R = decltype(rc)::value;
        const int j = j0 + R;
        const int k = zb - 2 + j;
        const bool arrive = STEADY || j < nslots;        // a level-0 plane arrives at this step
        const bool do_a = STEADY || (arrive && j >= 2);  // ... and completes level-1 plane k-1
        const bool push_b = STEADY || (j >= LAG && j < nslots + GS);       // level-1 plane k-LAG+1 exists
        const bool do_b = STEADY || (j >= LAG + 2 && j < nslots + GS);     // level-2 plane k-LAG is wanted
        T wa[VY + 2][VX + 2], wb[VY + 2][VX + 2];
        // ---- level-0 reads ----------------------------------------------------------------------------
        if (arrive) {
          mbar_wait(full0 + 8 * s, ph);
          const T *cur = ring0 + s * SLOT_ELEMS + soff0;
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
        }
        // ---- phase A: level-1 plane k-1 --------------------------------------------------------------
        if (do_a) {
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
        // ---- phase B': level-2 plane k-LAG -------------------------------------------------------------
        // (its shared-memory reads come after phase A's arithmetic in program order, as in sweep_tb2m: reading
        // both windows up front keeps 12 more fp64 values live through phase A and spills at 96 registers)
        if (push_b && wmask != 0) {
          const T *pnew = ring1 + slot(c1, GS + 1) * L1_ELEMS + soff1;   // level-1 plane k-LAG+1 (this thread's block): slot c1 - GS
#pragma unroll
          for (int r = 0; r < VY; ++r) lds_vec<T, VX>(b[R + 2][r], pnew + r * G::P1);
          if (do_b) load_halo<T, VX, VY, G::P1>(wb, ring1 + slot(c1, GS) * L1_ELEMS + soff1, swap);   // plane k-LAG: slot c1 - GS - 1
        }
        if (do_b) {
          if (wmask != 0) {
#pragma unroll
            for (int iy = 0; iy < VY; ++iy)
#pragma unroll
              for (int ix = 0; ix < VX; ++ix) wb[1 + iy][1 + ix] = b[R + 1][iy][ix];
            const unsigned m = (k - LAG >= p.blo[0] && k - LAG < p.bhi[0]) ? bmask : 0u;
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
          dst += p.out_sz;
        }
        if (STEADY || (j >= 2 && j < nslots + GS)) c1 = slot(c1, 1u);
        // one barrier per group of GS steps: the group's level-1 planes become visible, and the planes phase B'
        // read (the next group's writes, a full ring on) are free
        if constexpr ((R + 1) % GS == 0) asm volatile("bar.sync 1, %0;" ::"n"(G::CONSUMERS) : "memory");
      });
      // rotate the windows by U planes (the only register moves of an iteration)
#pragma unroll
      for (int q = 0; q < 2; ++q)
#pragma unroll
        for (int iy = 0; iy < VY; ++iy)
#pragma unroll
          for (int ix = 0; ix < VX; ++ix) { a[q][iy][ix] = a[q + U][iy][ix]; b[q][iy][ix] = b[q + U][iy][ix]; }
      });
    }
  }
}

}  // namespace neptune
