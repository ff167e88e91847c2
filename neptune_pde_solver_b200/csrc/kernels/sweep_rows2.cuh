// Temporally blocked row marching for rank-2 applies (sm_100a): TWO time steps of a radius-1,
// single-input body per launch — the 2-D counterpart of sweep_tb2m.cuh, built on sweep_rows.cuh.
//
// A CTA owns a strip of TX result columns and marches down the rows.  Per stage of RB level-0 rows
// (bulk-copied row segments of TX + 4*VX elements, ring of NS stages, producer warp as in sweep_rows):
//   phase A  every thread pushes the RB new level-0 rows through its 3 x (VX+2) register window and
//            evaluates the level-1 row behind it for its VX columns; the thread grid is TX/VX + 2
//            wide (one ring thread per side), so level 1 exists on the strip plus one column each side;
//            level-1 rows go to a double-buffered shared-memory row block;
//   barrier  one consumer-wide barrier per stage (the block written two stages ago is the one being
//            overwritten, and everybody has passed the barrier in between);
//   phase B  the inner threads push those RB level-1 rows (own columns and the neighbours' edge
//            columns, from shared memory) through a second register window and evaluate the level-2 row
//            behind it, written to global memory with 128-bit stores.
// Both levels run the body functor F; points outside `bounds` carry input 0 at both levels (the
// reference's copy-through, lib/Passes/DataflowLowering.cpp:282-287), so the result is bit-identical to
// two single sweeps.  DRAM traffic per point: one read and one write per TWO updates.
// Stages strictly inside the apply bounds take a path without row, bounds or edge tests.
#pragma once
#include "sweep_rows.cuh"

namespace neptune {

template <int CONSUMERS_, int RB_, int NS_, int MINB_> struct Rows2Cfg {
  static constexpr int CONSUMERS = CONSUMERS_, RB = RB_, NS = NS_, MINB = MINB_;
};

template <class F, class C> struct Rows2Geom {
  typedef typename F::T T;
  static constexpr int S = sizeof(T);
  static constexpr int VX = 16 / S;
  static constexpr int CONSUMERS = C::CONSUMERS;
  static constexpr int NBX = CONSUMERS;                 // thread grid: ring thread, TX/VX inner threads, ring thread
  static constexpr int TX = (NBX - 2) * VX;             // result columns per strip (not a power of two: every thread is used)
  static constexpr int SXP = TX + 4 * VX;               // staged level-0 row: two-point halo kept 16-byte aligned
  static constexpr int ROW_BYTES = SXP * S;
  static constexpr int RB = C::RB, NS = C::NS;
  static constexpr int STAGE_BYTES = (RB * ROW_BYTES + 127) / 128 * 128;
  static constexpr int P1 = NBX * VX;                   // level-1 row (elements)
  static constexpr int L1_BYTES = (RB * P1 * S + 127) / 128 * 128;   // one level-1 row block
  static constexpr int THREADS = CONSUMERS + 32;
  static constexpr int SMEM_BYTES = NS * STAGE_BYTES + 2 * L1_BYTES + 2 * NS * 8 + 128;
  static_assert(F::RANK == 2 && F::RZ == 0 && F::RY == 1 && F::RX <= 1, "rank-2 radius-1 bodies");
  static_assert(F::NIN == 1 && F::STAGED == 0, "temporal blocking of rank-2 bodies takes a single input");
  static_assert(CONSUMERS % 32 == 0 && THREADS <= 1024, "bad thread count");
  static_assert(SMEM_BYTES <= 227 * 1024, "ring does not fit in shared memory");
};

template <class F, class C, bool FMA>
__global__ void __launch_bounds__((Rows2Geom<F, C>::THREADS), C::MINB)
sweep_rows2(const __grid_constant__ KParams<F> p) {
  typedef Rows2Geom<F, C> G;
  typedef typename F::T T;
  constexpr int VX = G::VX, NS = G::NS, RB = G::RB, S = G::S, TX = G::TX;
  constexpr int NCW = G::CONSUMERS / 32;
  constexpr unsigned XFULL = (1u << VX) - 1u;

  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char *smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  unsigned char *l1 = smem + NS * G::STAGE_BYTES;             // two level-1 row blocks
  const uint32_t full0 = smem_u32(l1 + 2 * G::L1_BYTES);
  const uint32_t empty0 = full0 + 8 * NS;
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

  // work item -> result rows [yb, ye) of strip x0 .. x0 + TX
  auto geom = [&](int item, int &yb, int &ye, int &x0) {
    const int chunk = item / p.ncx, cx = item - chunk * p.ncx;
    yb = p.wlo[1] + chunk * p.zchunk + (chunk ? p.zgap : 0);
    ye = min(p.whi[1], yb + p.zchunk);
    x0 = (p.cxo + cx) * TX;
  };
  // Stage sequence of an item: a head stage with level-0 rows yb-2, yb-1 (they fill the window), then
  // stages of RB level-0 rows q = yb .. ye+1; row q completes level-1 row q-1 and level-2 row q-2.

  if (tid >= G::CONSUMERS) {
    // ---- producer -----------------------------------------------------------------------------------
    if (tid == G::CONSUMERS) {
      const T *src0 = p.in[0];
      const long long sy = p.in_sy[0];
      unsigned n = 0;
      for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
        int yb, ye, x0;
        geom(item, yb, ye, x0);
        const int xs = max(0, x0 - 2 * VX), xe = min(p.nx, x0 + TX + 2 * VX);
        const uint32_t row_bytes = (uint32_t)max(0, xe - xs) * S;
        const uint32_t doff = (uint32_t)(xs - (x0 - 2 * VX)) * S;
        for (int ya = yb - 2, cnt = 2; ya <= ye + 1; ya += cnt, cnt = min(RB, ye + 2 - ya), ++n) {
          const unsigned s = n % NS;
          mbar_wait(empty0 + 8 * s, ((n / NS) & 1) ^ 1);
          const int r_lo = max(0, -ya), r_hi = min(cnt, p.ny - ya);   // rows that exist in the box
          mbar_expect_tx(full0 + 8 * s, (uint32_t)max(0, r_hi - r_lo) * row_bytes);
          if (row_bytes)
            for (int r = r_lo; r < r_hi; ++r)
              bulk_load_1d(stage0 + s * G::STAGE_BYTES + r * G::ROW_BYTES + doff, src0 + (long long)(ya + r) * sy + xs, row_bytes,
                           full0 + 8 * s);
        }
      }
    }
    return;
  }

  // ---- consumers ------------------------------------------------------------------------------------
  const int bx = tid;
  const bool inner = bx >= 1 && bx <= G::NBX - 2;
  const bool lane0 = (tid & 31) == 0;
  const bool swap = ((tid >> 3) & 1) != 0;   // alternate left/right halo reads between lane groups (bank conflicts)
  unsigned n = 0, lb = 0;
  T w0[3][VX + 2], w1[3][VX + 2];   // level-0 / level-1 rows y-1, y, y+1 around the thread's VX columns
#pragma unroll
  for (int a = 0; a < 3; ++a)
#pragma unroll
    for (int b = 0; b < VX + 2; ++b) w0[a][b] = w1[a][b] = T(0);

  auto push = [&](T (&w)[3][VX + 2], const T *row) {
#pragma unroll
    for (int b = 0; b < VX + 2; ++b) { w[0][b] = w[1][b]; w[1][b] = w[2][b]; }
    T ctr[VX];
    lds_vec<T, VX>(ctr, row);
#pragma unroll
    for (int i = 0; i < VX; ++i) w[2][1 + i] = ctr[i];
    if constexpr (F::RX > 0) {
      const T v1 = row[swap ? VX : -1], v2 = row[swap ? -1 : VX];
      w[2][0] = swap ? v2 : v1;
      w[2][VX + 1] = swap ? v1 : v2;
    }
  };
  T dummy_pw[1][VX];
#pragma unroll
  for (int i = 0; i < VX; ++i) dummy_pw[0][i] = T(0);

  for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
    int yb, ye, x0;
    geom(item, yb, ye, x0);
    const int gx = x0 - VX + bx * VX;
    unsigned xmask = 0, wmask = 0;   // bit ix: column inside the apply bounds / inner thread and column inside the written box
#pragma unroll
    for (int ix = 0; ix < VX; ++ix) {
      if (gx + ix >= p.blo[2] && gx + ix < p.bhi[2]) xmask |= 1u << ix;
      if (inner && gx + ix >= p.wlo[2] && gx + ix < p.whi[2]) wmask |= 1u << ix;
    }
    const bool xfast = xmask == XFULL;
    const int so0 = VX + bx * VX;   // the thread's columns inside a staged level-0 row
    const int so1 = bx * VX;        // ... inside a level-1 row

    // head stage: level-0 rows yb-2, yb-1
    {
      const unsigned s = n % NS;
      mbar_wait(full0 + 8 * s, (n / NS) & 1);
      const T *sp = reinterpret_cast<const T *>(smem + s * G::STAGE_BYTES) + so0;
      push(w0, sp);
      push(w0, sp + G::SXP);
      __syncwarp();
      if (lane0) mbar_arrive(empty0 + 8 * s);
      ++n;
    }

    for (int q0 = yb; q0 <= ye + 1; q0 += RB, ++n, lb ^= 1u) {
      const unsigned s = n % NS;
      const int cnt = min(RB, ye + 2 - q0);
      mbar_wait(full0 + 8 * s, (n / NS) & 1);
      const T *sp = reinterpret_cast<const T *>(smem + s * G::STAGE_BYTES) + so0;
      T *l1p = reinterpret_cast<T *>(l1 + lb * G::L1_BYTES) + so1;

      // ---- phase A: level-1 rows q0-1 .. q0+cnt-2 ---------------------------------------------------
      if (xfast && cnt == RB && q0 - 1 >= p.blo[1] && q0 - 1 + RB <= p.bhi[1]) {
#pragma unroll
        for (int r = 0; r < RB; ++r) {
          push(w0, sp + r * G::SXP);
          T res[VX];
          static_for<VX>([&](auto ixc) {
            constexpr int IX = decltype(ixc)::value;
            RowAcc<F, VX, IX> acc{w0, dummy_pw};
            res[IX] = F::template eval<FMA>(acc, p.c);
          });
          stg_vec<T, VX, false>(l1p + r * G::P1, res);   // shared-memory store
        }
      } else {
        for (int r = 0; r < cnt; ++r) {
          push(w0, sp + r * G::SXP);
          const int y1 = q0 + r - 1;
          const bool yin = y1 >= p.blo[1] && y1 < p.bhi[1];
          T res[VX];
          static_for<VX>([&](auto ixc) {
            constexpr int IX = decltype(ixc)::value;
            T v = w0[1][IX + 1];   // copy-through: input 0 at this point
            if (yin && ((xmask >> IX) & 1u)) {
              RowAcc<F, VX, IX> acc{w0, dummy_pw};
              v = F::template eval<FMA>(acc, p.c);
            }
            res[IX] = v;
          });
          stg_vec<T, VX, false>(l1p + r * G::P1, res);
        }
      }
      // the level-0 stage is in registers as far as this warp needs it
      __syncwarp();
      if (lane0) mbar_arrive(empty0 + 8 * s);
      // level-1 rows of this stage visible to the neighbours (and the block written two stages ago free again)
      asm volatile("bar.sync 1, %0;" ::"n"(G::CONSUMERS) : "memory");

      // ---- phase B: level-2 rows q0-2 .. q0+cnt-3 ---------------------------------------------------
      if (inner) {
        const T *l1r = l1p;
        T *dst = p.out + (long long)(q0 - 2) * p.out_sy + gx;
        if (xfast && wmask == XFULL && cnt == RB && q0 - 2 >= max(yb, p.blo[1]) && q0 - 2 + RB <= min(ye, p.bhi[1])) {
#pragma unroll
          for (int r = 0; r < RB; ++r) {
            push(w1, l1r + r * G::P1);
            T res[VX];
            static_for<VX>([&](auto ixc) {
              constexpr int IX = decltype(ixc)::value;
              RowAcc<F, VX, IX> acc{w1, dummy_pw};
              res[IX] = F::template eval<FMA>(acc, p.c);
            });
            stg_vec<T, VX, false>(dst + (long long)r * p.out_sy, res);
          }
        } else {
          for (int r = 0; r < cnt; ++r) {
            push(w1, l1r + r * G::P1);
            const int y2 = q0 + r - 2;
            if (y2 >= yb && y2 < ye && wmask) {
              const bool yin = y2 >= p.blo[1] && y2 < p.bhi[1];
              T res[VX];
              static_for<VX>([&](auto ixc) {
                constexpr int IX = decltype(ixc)::value;
                T v = w1[1][IX + 1];   // copy-through: level 1 there is input 0 as well
                if (yin && ((xmask >> IX) & 1u)) {
                  RowAcc<F, VX, IX> acc{w1, dummy_pw};
                  v = F::template eval<FMA>(acc, p.c);
                }
                res[IX] = v;
              });
              T *d = dst + (long long)r * p.out_sy;
              if (wmask == XFULL) {
                stg_vec<T, VX, false>(d, res);
              } else {
#pragma unroll
                for (int ix = 0; ix < VX; ++ix)
                  if ((wmask >> ix) & 1u) d[ix] = res[ix];
              }
            }
          }
        }
      }
    }
  }
}

}  // namespace neptune
