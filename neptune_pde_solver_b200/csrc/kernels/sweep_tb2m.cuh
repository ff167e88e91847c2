// Temporally blocked sweep, register z-marching form (sm_100a): TWO time steps of a radius-1,
// single-input apply per launch, for bodies whose z taps sit on the centre column ("z-star":
// 7-point stars, and any body whose dz != 0 taps have dy = dx = 0).
//
// Why a second form.  r01 ncu of sweep_tb2 (profiles/r01_c3_tb2_ncu_summary.md): 0.55 shared-memory
// wavefronts per point-update, L1/shared pipe 82 % busy, DRAM at 3.2 TB/s — the kernel is bound by
// shared-memory bandwidth, because every thread re-reads its whole 3 x (VY+2) x (VX+2) window from
// shared memory at both levels, every plane.  Here a thread keeps what it already has in registers:
//   level 0: the centre blocks of planes k-2 and k-1 stay in registers while the thread marches in
//            z; per plane it reads only the centre block of the new plane k and the in-plane halo
//            (rows -1 / VY, columns -1 / VX) of plane k-1;
//   level 1: the centre blocks of planes k-3, k-2, k-1 are the thread's OWN phase-A results and never
//            come back from shared memory; only the in-plane halo of plane k-2 does.
// Left/right halo reads (8 bytes per lane at a 16-byte lane stride) would be 2-way bank conflicts;
// lanes alternate in groups of eight between "left first" and "right first" so each half-warp covers
// all 32 banks once.  Shared-memory wavefronts per thread-plane drop from 104 to 48.
//
// Since phase B no longer reads the level-1 plane written in the same iteration, the level-1 ring
// needs two planes only and phases A and B of one iteration are independent instruction streams
// (one consumer barrier per plane, at the end).  Arithmetic is the body's own op list at both
// levels (functor F), points outside `bounds` carry input 0 (copy-through,
// lib/Passes/DataflowLowering.cpp:282-287): the result is bit-identical to two single sweeps.
// Geometry, work decomposition and the TMA producer are those of sweep_tb2.cuh.
#pragma once
#include "sweep_tb2.cuh"

namespace neptune {

template <class F, class C> struct Tb2mGeom {
  typedef typename F::T T;
  static constexpr int S = sizeof(T);
  static constexpr int VX = 16 / S, VY = C::VY;
  static constexpr int NBX = C::TX / VX + 2, NBY = C::TY / VY + 2;  // thread-block grid
  static constexpr int ACTIVE = NBX * NBY;
  static constexpr int CONSUMERS = (ACTIVE + 31) / 32 * 32;
  static constexpr int THREADS = CONSUMERS + 32;
  static constexpr int NS = C::NS;
  static constexpr int P0 = C::TX + 4 * VX;        // level-0 stage row pitch = TMA box width
  static constexpr int BOX_Y = C::TY + 4;
  static constexpr int STAGE_TX_BYTES = P0 * BOX_Y * S;
  static constexpr int STAGE_BYTES = (STAGE_TX_BYTES + 127) / 128 * 128;
  static constexpr int NSTG = F::NIN;                      // staged tiles per ring slot: input 0, then the pointwise inputs
  static constexpr int SLOT_BYTES = NSTG * STAGE_BYTES;
  static constexpr int GUARD_BYTES = (P0 * S * (VY > 2 ? VY - 1 : 1) + 127) / 128 * 128;
  static constexpr int P1 = NBX * VX, H1 = NBY * VY;  // level-1 plane
  static constexpr int L1_BYTES = (P1 * H1 * S + 127) / 128 * 128;
  static constexpr int L1_GUARD = (P1 * S + 127) / 128 * 128;  // row -1 of the top ring blocks
  static constexpr int NL1 = 2;
  static constexpr int SMEM_BYTES = GUARD_BYTES + NS * SLOT_BYTES + GUARD_BYTES + L1_GUARD + NL1 * L1_BYTES + L1_GUARD +
                                    2 * NS * 8 + 128;
  static_assert(F::RZ == 1 && F::RY <= 1 && F::RX <= 1 && F::ZSTAR, "z-marching needs a radius-1 body with centre-column z taps");
  static_assert(F::STAGED == 0, "input 0 must be the only input read at neighbour offsets (the others are pointwise)");
  static_assert(F::NIN <= 2, "one pointwise input at most (its tiles share the ring slots)");
  static_assert(SMEM_BYTES <= 227 * 1024, "ring does not fit in shared memory");
  static_assert(C::TX % VX == 0 && C::TY % VY == 0, "tile must be a multiple of the thread block");
  static_assert(P0 <= 256 && BOX_Y <= 256, "TMA box dimensions are limited to 256 elements");
  static_assert(THREADS <= 1024, "too many threads");
};

// accessor: centre-plane window c[][], centre blocks m[][] (plane below) and p[][] (plane above) of
// input 0; pw[][][] = the pointwise inputs (coefficients, right-hand sides) at the thread's own points
template <class F, int VY, int VX, int IY, int IX> struct MarchAcc {
  typedef typename F::T T;
  static constexpr int NPW = F::NIN > 1 ? F::NIN - 1 : 1;
  const T (&m)[VY][VX];
  const T (&c)[VY + 2][VX + 2];
  const T (&p)[VY][VX];
  const T (&pw)[NPW][VY][VX];
  template <int K, int DZ, int DY, int DX> __device__ __forceinline__ T tap() const {
    if constexpr (K != 0) {
      static_assert(DZ == 0 && DY == 0 && DX == 0, "inputs other than input 0 are read at offset 0 only");
      return pw[K - 1][IY][IX];
    } else if constexpr (DZ == 0) {
      return c[IY + DY + 1][IX + DX + 1];
    } else {
      static_assert(DY == 0 && DX == 0, "z taps sit on the centre column");
      return DZ < 0 ? m[IY][IX] : p[IY][IX];
    }
  }
};

// in-plane halo of a VY x VX block whose first point is at `blk` (row pitch PITCH) into window c[][];
// the centre of c[][] is filled by the caller from registers.  `swap` alternates the order of the
// left/right reads between lane groups (bank-conflict-free, see header).
template <class T, int VX, int VY, int PITCH>
__device__ __forceinline__ void load_halo(T (&c)[VY + 2][VX + 2], const T *blk, bool swap) {
  {
    const T *top = blk - PITCH, *bot = blk + VY * PITCH;
    T t[VX], b[VX];
    lds_vec<T, VX>(t, top);
    lds_vec<T, VX>(b, bot);
#pragma unroll
    for (int i = 0; i < VX; ++i) { c[0][1 + i] = t[i]; c[VY + 1][1 + i] = b[i]; }
    // corners (only bodies with in-plane diagonal taps keep these loads alive)
    c[0][0] = top[-1]; c[0][VX + 1] = top[VX];
    c[VY + 1][0] = bot[-1]; c[VY + 1][VX + 1] = bot[VX];
  }
  const int o1 = swap ? VX : -1, o2 = swap ? -1 : VX;
#pragma unroll
  for (int r = 0; r < VY; ++r) {
    const T *row = blk + r * PITCH;
    const T v1 = row[o1], v2 = row[o2];
    c[1 + r][0] = swap ? v2 : v1;
    c[1 + r][VX + 1] = swap ? v1 : v2;
  }
}

template <class F, class C, bool FMA>
__global__ void __launch_bounds__((Tb2mGeom<F, C>::THREADS), C::MINB)
sweep_tb2m(const __grid_constant__ KParams<F> p, const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_pw) {
  typedef Tb2mGeom<F, C> G;
  typedef typename F::T T;
  constexpr int VX = G::VX, VY = G::VY, NS = G::NS;
  constexpr int NCW = G::CONSUMERS / 32;
  constexpr int NPW = F::NIN > 1 ? F::NIN - 1 : 1;

  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char *smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  unsigned char *l0 = smem + G::GUARD_BYTES;                                          // NS level-0 stages
  unsigned char *l1 = l0 + NS * G::SLOT_BYTES + G::GUARD_BYTES + G::L1_GUARD;          // 2 level-1 planes
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
          mbar_expect_tx(full0 + 8 * s, G::NSTG * G::STAGE_TX_BYTES);
          tma_load_3d(stage0 + s * G::SLOT_BYTES, &tmap, x0 - 2 * VX, y0 - 2, zb - 2 + j, full0 + 8 * s);
          if constexpr (F::NIN > 1)   // the pointwise input's tile of the same plane, same box (out-of-box elements arrive as zeros)
            tma_load_3d(stage0 + s * G::SLOT_BYTES + G::STAGE_BYTES, &tmap_pw, x0 - 2 * VX, y0 - 2, zb - 2 + j, full0 + 8 * s);
        }
      }
    }
    return;
  }


  // ---- consumers ----------------------------------------------------------------------------------
  // (r01 notes, C3 512^3 f64: this plain form measured 427 Gupd/s.  Three variations were tried and
  // dropped: a 3x unrolled plane loop with statically indexed planes instead of the register moves
  // below (ptxas spilled, 245 Gupd/s); an `all in bounds` fast path without the copy-through selects
  // (the extra branches keep ptxas from interleaving the DP chains of phases A and B, 293 Gupd/s);
  // explicit 32-bit shared addresses (390 Gupd/s); unconditional evaluation with the copy-through selects moved
  // into a branch taken only at the bounds' edge (402 Gupd/s: same story, the branch splits the stream).  The dependent DADD chain of the body is what the
  // warps wait on, so instruction-level parallelism across the two phases matters more than count.)
  const bool active = tid < G::ACTIVE;
  const int bx = active ? tid % G::NBX : 0, by = active ? tid / G::NBX : 0;
  const bool inner = active && bx >= 1 && bx <= G::NBX - 2 && by >= 1 && by <= G::NBY - 2;
  const bool lane0 = (tid & 31) == 0;
  const bool swap = ((tid >> 3) & 1) != 0;
  const T *const ring0 = reinterpret_cast<const T *>(l0);
  T *const ring1 = reinterpret_cast<T *>(l1);
  constexpr int STAGE_ELEMS = G::STAGE_BYTES / (int)sizeof(T);   // offset of the pointwise tile inside a slot
  constexpr int SLOT_ELEMS = G::SLOT_BYTES / (int)sizeof(T);
  constexpr int L1_ELEMS = G::L1_BYTES / (int)sizeof(T);
  const int soff0 = (by * VY - VY + 2) * G::P0 + VX + bx * VX;   // thread block's first point inside a level-0 stage
  const int soff1 = (by * VY) * G::P1 + bx * VX;                 // ... inside a level-1 plane
  unsigned s = 0, ph = 0;   // level-0 ring slot / phase of the next plane to arrive
  unsigned c1 = 0;          // level-1 planes produced so far (slot = c1 & 1)
  // z-marching registers: level-0 centre blocks of planes k-2 (am) and k-1 (ac); level-1 centre
  // blocks of planes k-3 (bm) and k-2 (bc)
  T am[VY][VX], ac[VY][VX], bm[VY][VX], bc[VY][VX];
#pragma unroll
  for (int iy = 0; iy < VY; ++iy)
#pragma unroll
    for (int ix = 0; ix < VX; ++ix) am[iy][ix] = ac[iy][ix] = bm[iy][ix] = bc[iy][ix] = T(0);
  // pointwise input at the thread's points: plane k-1 (qc, read this iteration from the tile the TMA
  // producer stages beside input 0's, used by phase A) and plane k-2 (qm, last iteration's qc, used
  // by phase B).  It is the same field at both levels: a time loop replaces input 0 only.  (r01: reading
  // it with __ldg instead left the warps waiting on DRAM latency once per plane: 16 warps per SM,
  // IPC 1.8, DRAM at 3.8 TB/s, and its halo re-reads missed L2: +12 % DRAM reads.)
  T qc[NPW][VY][VX], qm[NPW][VY][VX];
#pragma unroll
  for (int q = 0; q < NPW; ++q)
#pragma unroll
    for (int iy = 0; iy < VY; ++iy)
#pragma unroll
      for (int ix = 0; ix < VX; ++ix) qc[q][iy][ix] = qm[q][iy][ix] = T(0);

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
    const T *prv = ring0 + soff0;  // this thread's block inside the stage of plane k-1
    unsigned s_prv = 0;
    for (int j = 0; j < nslots; ++j) {
      const int k = zb - 2 + j;  // level-0 plane arriving now
      mbar_wait(full0 + 8 * s, ph);
      const T *cur = ring0 + s * SLOT_ELEMS + soff0;

      // ---- level-0 reads of this iteration ----------------------------------------------------------
      T ap[VY][VX];                 // level-0 centre block of plane k
      T wa[VY + 2][VX + 2];         // level-0 window of plane k-1 (halo from shared memory, centre = ac)
      if (active) {
#pragma unroll
        for (int r = 0; r < VY; ++r) lds_vec<T, VX>(ap[r], cur + r * G::P0);
        if (j >= 2) {
          load_halo<T, VX, VY, G::P0>(wa, prv, swap);
          if constexpr (F::NIN > 1) {   // pointwise input at the thread's points of plane k-1 (phase A; phase B next iteration)
#pragma unroll
            for (int r = 0; r < VY; ++r) lds_vec<T, VX>(qc[0][r], prv + STAGE_ELEMS + r * G::P0);
          }
        }
      }
      // the level-0 stage of plane k-1 is now in registers as far as this warp needs it
      __syncwarp();
      if (lane0) {
        if (j >= 1) mbar_arrive(empty0 + 8 * s_prv);
        if (j == nslots - 1) mbar_arrive(empty0 + 8 * s);
      }
      s_prv = s;
      prv = cur;
      if (++s == NS) { s = 0; ph ^= 1; }

      if (j >= 2) {
        // ---- phase A: level-1 plane k-1 over the whole thread grid ------------------------------
        T ra[VY][VX];
        if (active) {
#pragma unroll
          for (int iy = 0; iy < VY; ++iy)
#pragma unroll
            for (int ix = 0; ix < VX; ++ix) wa[1 + iy][1 + ix] = ac[iy][ix];
          const unsigned m = (k - 1 >= p.blo[0] && k - 1 < p.bhi[0]) ? bmask : 0u;
          T *l1p = ring1 + (c1 & 1u) * L1_ELEMS + soff1;
          static_for<VY>([&](auto iyc) {
            constexpr int IY = decltype(iyc)::value;
            static_for<VX>([&](auto ixc) {
              constexpr int IX = decltype(ixc)::value;
              MarchAcc<F, VY, VX, IY, IX> acc{am, wa, ap, qc};
              const T v = F::template eval<FMA>(acc, p.c);
              ra[IY][IX] = ((m >> (IY * VX + IX)) & 1u) ? v : ac[IY][IX];
            });
            stg_vec<T, VX, false>(l1p + IY * G::P1, ra[IY]);  // shared-memory store of the level-1 block
          });
        }
        // ---- phase B: level-2 plane k-2 from level-1 planes k-3 (bm), k-2 (bc + halo), k-1 (ra) --
        if (j >= 4 && wmask != 0) {
          T wb[VY + 2][VX + 2];  // level-1 window of plane k-2 (halo from shared memory, centre = bc)
          load_halo<T, VX, VY, G::P1>(wb, ring1 + ((c1 + 1) & 1u) * L1_ELEMS + soff1, swap);
#pragma unroll
          for (int iy = 0; iy < VY; ++iy)
#pragma unroll
            for (int ix = 0; ix < VX; ++ix) wb[1 + iy][1 + ix] = bc[iy][ix];
          const unsigned m = (k - 2 >= p.blo[0] && k - 2 < p.bhi[0]) ? bmask : 0u;
          static_for<VY>([&](auto iyc) {
            constexpr int IY = decltype(iyc)::value;
            T res[VX];
            static_for<VX>([&](auto ixc) {
              constexpr int IX = decltype(ixc)::value;
              MarchAcc<F, VY, VX, IY, IX> acc{bm, wb, ra, qm};
              const T v = F::template eval<FMA>(acc, p.c);
              res[IX] = ((m >> (IY * VX + IX)) & 1u) ? v : bc[IY][IX];
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
        if (j >= 4) dst += p.out_sz;
#pragma unroll
        for (int iy = 0; iy < VY; ++iy)
#pragma unroll
          for (int ix = 0; ix < VX; ++ix) { bm[iy][ix] = bc[iy][ix]; bc[iy][ix] = ra[iy][ix]; }
        if constexpr (F::NIN > 1) {
#pragma unroll
          for (int q = 0; q < NPW; ++q)
#pragma unroll
            for (int iy = 0; iy < VY; ++iy)
#pragma unroll
              for (int ix = 0; ix < VX; ++ix) qm[q][iy][ix] = qc[q][iy][ix];
        }
        ++c1;
      }
#pragma unroll
      for (int iy = 0; iy < VY; ++iy)
#pragma unroll
        for (int ix = 0; ix < VX; ++ix) { am[iy][ix] = ac[iy][ix]; ac[iy][ix] = ap[iy][ix]; }
      // one consumer-wide barrier per plane: this plane's level-1 stores become visible to the
      // halo reads of the next iteration, and this iteration's halo reads are over before the
      // iteration after next overwrites their plane (ring of 2)
      if (j >= 2) asm volatile("bar.sync 1, %0;" ::"n"(G::CONSUMERS) : "memory");
    }
  }
}

}  // namespace neptune
