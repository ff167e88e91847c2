// Row-marching sweep for rank-2 applies (sm_100a): the 2-D analogue of the z-march of sweep_tma.cuh.
//
// Why a second single-step form.  sweep_tma sees a rank-2 box as ONE plane (canonical z extent 1), so a
// 2-D sweep degenerates into independent TX x TY tiles: every tile re-reads a y halo (34 rows for 32),
// nothing is marched, and r01 measured 0.70-0.75 of the copy peak on C1/C2 against 0.86 for the 3-D
// z-march.  Here a CTA owns a STRIP of TX columns and marches down the rows:
//   * rows are staged into shared memory by the bulk-copy engine, one `cp.async.bulk` (1-D, not
//     tensor: a row segment of TX + 2*HX elements is contiguous, and a 1-D bulk copy has no 256-element
//     box limit, so strips can be 512+ columns wide and the x halo costs < 1 % extra traffic);
//     RB rows form a stage that completes on one mbarrier; a producer warp keeps a ring of NS stages
//     full, consumer warps hand stages back through empty[] barriers (same protocol as sweep_tma);
//   * every consumer thread owns VX adjacent columns (16 bytes) and keeps a 3 x (VX + 2) register
//     window of rows y-1, y, y+1: per new row it reads ONE 128-bit centre vector and two halo scalars
//     from shared memory, shifts the window, evaluates the body functor for row y and writes 16 bytes.
//     Each input row is read from DRAM once (plus one halo row pair per row chunk).
// Semantics as sweep_tma: out = copy of input 0, body over `bounds` (points outside store the window's
// centre value); pointwise inputs stream with 128-bit __ldg.  Elements of a staged row that lie outside
// the box are never written by the copy engine and never feed an in-bounds point.
#pragma once
#include "sweep_tma.cuh"

namespace neptune {

template <int TX_, int RB_, int NS_, int MINB_> struct RowCfg {
  static constexpr int TX = TX_, RB = RB_, NS = NS_, MINB = MINB_;
};

template <class F, class C> struct RowGeom {
  typedef typename F::T T;
  static constexpr int S = sizeof(T);
  static constexpr int VX = 16 / S;
  static constexpr int RX = F::RX, RY = F::RY;
  static constexpr int HX = RX > 0 ? ((RX + VX - 1) / VX) * VX : 0;
  static constexpr int SXP = C::TX + 2 * HX;                 // staged row length (elements)
  static constexpr int ROW_BYTES = SXP * S;
  static constexpr int STAGE_BYTES = (C::RB * ROW_BYTES + 127) / 128 * 128;
  static constexpr int CONSUMERS = C::TX / VX;
  static constexpr int THREADS = CONSUMERS + 32;
  static constexpr int NS = C::NS, RB = C::RB;
  static constexpr int SMEM_BYTES = NS * STAGE_BYTES + 2 * NS * 8 + 128;
  static_assert(F::RANK == 2 && F::RZ == 0 && RY == 1 && RX <= 1, "row marching is for rank-2 radius-1 bodies");
  static_assert(C::TX % VX == 0 && CONSUMERS % 32 == 0 && THREADS <= 1024, "bad strip width");
  static_assert(ROW_BYTES % 16 == 0, "bulk copies move multiples of 16 bytes");
};

__device__ __forceinline__ void bulk_load_1d(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

// register window accessor: w[dy + 1][ix + dx + 1] for input 0, pw[][] for pointwise inputs
template <class F, int VX, int IX> struct RowAcc {
  typedef typename F::T T;
  static constexpr int NPW = F::NIN > 1 ? F::NIN - 1 : 1;
  const T (&w)[3][VX + 2];
  const T (&pw)[NPW][VX];
  template <int K, int DZ, int DY, int DX> __device__ __forceinline__ T tap() const {
    static_assert(DZ == 0, "rank-2 body");
    if constexpr (K == F::STAGED) {
      return w[DY + 1][IX + DX + 1];
    } else {
      static_assert(DY == 0 && DX == 0, "non-staged inputs are read at offset 0 only");
      return pw[K < F::STAGED ? K : K - 1][IX];
    }
  }
};

template <class F, class C, bool FMA>
__global__ void __launch_bounds__((RowGeom<F, C>::THREADS), C::MINB)
sweep_rows(const __grid_constant__ KParams<F> p) {
  typedef RowGeom<F, C> G;
  typedef typename F::T T;
  constexpr int VX = G::VX, HX = G::HX, NS = G::NS, RB = G::RB, S = G::S;
  constexpr int NPW = F::NIN > 1 ? F::NIN - 1 : 1;
  constexpr int NCW = G::CONSUMERS / 32;
  constexpr unsigned XFULL = (1u << VX) - 1u;

  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char *smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  const uint32_t full0 = smem_u32(smem + NS * G::STAGE_BYTES);
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

  // work item -> rows [yb, ye) of strip x0 .. x0 + TX
  auto geom = [&](int item, int &yb, int &ye, int &x0) {
    const int chunk = item / p.ncx, cx = item - chunk * p.ncx;
    yb = p.wlo[1] + chunk * p.zchunk + (chunk ? p.zgap : 0);
    ye = min(p.whi[1], yb + p.zchunk);
    x0 = (p.cxo + cx) * C::TX;
  };
  // Stage sequence of an item: a head stage with input rows yb-1 and yb (they fill the window, no
  // output), then stages of RB input rows yb+1+m*RB .. ; input row yb+1+q completes output row yb+q.

  if (tid >= G::CONSUMERS) {
    // ---- producer -----------------------------------------------------------------------------------
    if (tid == G::CONSUMERS) {
      const T *src0 = p.in[F::STAGED];
      const long long sy = p.in_sy[F::STAGED];
      unsigned n = 0;
      for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
        int yb, ye, x0;
        geom(item, yb, ye, x0);
        const int xs = max(0, x0 - HX), xe = min(p.nx, x0 + C::TX + HX);
        const uint32_t row_bytes = (uint32_t)(xe - xs) * S;
        const uint32_t doff = (uint32_t)(xs - (x0 - HX)) * S;
        // ya = first input row of the stage, cnt = rows it holds
        for (int ya = yb - 1, cnt = 2; ya <= ye; ya += cnt, cnt = min(RB, ye + 1 - ya), ++n) {
          const unsigned s = n % NS;
          mbar_wait(empty0 + 8 * s, ((n / NS) & 1) ^ 1);
          // rows that exist in the box (row -1 and row ny of the box's edge chunks do not)
          const int r_lo = max(0, -ya), r_hi = min(cnt, p.ny - ya);
          mbar_expect_tx(full0 + 8 * s, (uint32_t)max(0, r_hi - r_lo) * row_bytes);
          for (int r = r_lo; r < r_hi; ++r)
            bulk_load_1d(stage0 + s * G::STAGE_BYTES + r * G::ROW_BYTES + doff, src0 + (long long)(ya + r) * sy + xs, row_bytes,
                         full0 + 8 * s);
        }
      }
    }
    return;
  }

  // ---- consumers ------------------------------------------------------------------------------------
  const int lx = tid * VX;
  const bool lane0 = (tid & 31) == 0;
  const bool swap = ((tid >> 3) & 1) != 0;   // alternate left/right halo reads between lane groups (bank conflicts)
  unsigned n = 0;
  T w[3][VX + 2];   // rows y-1, y, y+1 of input 0 around the thread's VX columns
#pragma unroll
  for (int a = 0; a < 3; ++a)
#pragma unroll
    for (int b = 0; b < VX + 2; ++b) w[a][b] = T(0);

  // shift the window down and read one staged row into w[2]
  auto push_row = [&](const T *row) {
#pragma unroll
    for (int b = 0; b < VX + 2; ++b) { w[0][b] = w[1][b]; w[1][b] = w[2][b]; }
    T ctr[VX];
    lds_vec<T, VX>(ctr, row);
#pragma unroll
    for (int i = 0; i < VX; ++i) w[2][1 + i] = ctr[i];
    if constexpr (G::RX > 0) {
      const T v1 = row[swap ? VX : -1], v2 = row[swap ? -1 : VX];
      w[2][0] = swap ? v2 : v1;
      w[2][VX + 1] = swap ? v1 : v2;
    }
  };

  for (int item = blockIdx.x; item < p.nitems; item += gridDim.x) {
    int yb, ye, x0;
    geom(item, yb, ye, x0);
    const int gx = x0 + lx;
    const bool live = gx < p.whi[2] && gx + VX > p.wlo[2];
    const bool full_vec = gx >= p.wlo[2] && gx + VX <= p.whi[2];
    unsigned xmask = 0;   // bit ix: column inside the apply bounds
#pragma unroll
    for (int ix = 0; ix < VX; ++ix)
      if (gx + ix >= p.blo[2] && gx + ix < p.bhi[2]) xmask |= 1u << ix;
    const bool xfast = live && full_vec && xmask == XFULL;
    T *dst = p.out + (long long)yb * p.out_sy + gx;

    // head stage: rows yb-1, yb
    {
      const unsigned s = n % NS;
      mbar_wait(full0 + 8 * s, (n / NS) & 1);
      if (live) {
        const T *sp = reinterpret_cast<const T *>(smem + s * G::STAGE_BYTES) + HX + lx;
        push_row(sp);
        push_row(sp + G::SXP);
      }
      __syncwarp();
      if (lane0) mbar_arrive(empty0 + 8 * s);
      ++n;
    }

    for (int y0 = yb; y0 < ye; y0 += RB, ++n) {   // this stage completes output rows y0 .. y0+cnt-1
      const unsigned s = n % NS;
      const int cnt = min(RB, ye - y0);
      mbar_wait(full0 + 8 * s, (n / NS) & 1);
      const T *sp = reinterpret_cast<const T *>(smem + s * G::STAGE_BYTES) + HX + lx;
      if (xfast && cnt == RB && y0 >= p.blo[1] && y0 + RB <= p.bhi[1]) {
        // steady state: a whole stage strictly inside the apply bounds — no row, bounds or edge tests
        T pw[RB][NPW][VX];
        if constexpr (F::NIN > 1) {
#pragma unroll
          for (int r = 0; r < RB; ++r)
            static_for<F::NIN>([&](auto kc) {
              constexpr int K = decltype(kc)::value;
              if constexpr (K != F::STAGED)
                ldg_vec<T, VX>(pw[r][K < F::STAGED ? K : K - 1], p.in[K] + (long long)(y0 + r) * p.in_sy[K] + gx);
            });
        }
#pragma unroll
        for (int r = 0; r < RB; ++r) {
          push_row(sp + r * G::SXP);
          T res[VX];
          static_for<VX>([&](auto ixc) {
            constexpr int IX = decltype(ixc)::value;
            RowAcc<F, VX, IX> acc{w, pw[r]};
            res[IX] = F::template eval<FMA>(acc, p.c);
          });
          stg_vec<T, VX, false>(dst + r * p.out_sy, res);
        }
      } else if (live) {
        // edges: ragged last stage, rows or columns outside the apply bounds (copy-through), partial vectors
        for (int r = 0; r < cnt; ++r) {
          const int y = y0 + r;
          push_row(sp + r * G::SXP);
          const bool yin = y >= p.blo[1] && y < p.bhi[1];
          T pw[NPW][VX];
          if constexpr (F::NIN > 1) {
            if (yin && xmask) {
              static_for<F::NIN>([&](auto kc) {
                constexpr int K = decltype(kc)::value;
                if constexpr (K != F::STAGED)
                  ldg_vec<T, VX>(pw[K < F::STAGED ? K : K - 1], p.in[K] + (long long)y * p.in_sy[K] + gx);
              });
            }
          }
          T res[VX];
          static_for<VX>([&](auto ixc) {
            constexpr int IX = decltype(ixc)::value;
            T v = w[1][IX + 1];   // copy-through: input 0 at this point
            if (yin && ((xmask >> IX) & 1u)) {
              RowAcc<F, VX, IX> acc{w, pw};
              v = F::template eval<FMA>(acc, p.c);
            }
            res[IX] = v;
          });
          T *d = dst + r * p.out_sy;
          if (full_vec) {
            stg_vec<T, VX, false>(d, res);
          } else {
#pragma unroll
            for (int ix = 0; ix < VX; ++ix)
              if (gx + ix >= p.wlo[2] && gx + ix < p.whi[2]) d[ix] = res[ix];
          }
        }
      }
      // every row of this stage is in registers as far as this warp needs it
      __syncwarp();
      if (lane0) mbar_arrive(empty0 + 8 * s);
      dst += (long long)RB * p.out_sy;
    }
  }
}

}  // namespace neptune
