// Temporally blocked z-march, grouped and software-pipelined, with a lean steady state (sm_100a).  The default
// kernel of `neptune_rt_run_steps` on wide, deep single-input 3-D boxes (BASELINE configs[2]).
//
// Data movement (r01's sweep_tb2p, which this kernel replaces): geometry, TMA producer and z-star register windows of
// sweep_tb2m.cuh, but three planes per consumer barrier, and phase B of the PREVIOUS group riding in the same steps
// as phase A of the current one.  Step R of a group:
//     A : level-0 plane k arrives  -> level-1 plane k-1 (stored to the level-1 ring)
//     B': level-2 plane k-LAG        <- level-1 planes k-6, k-5, k-4 (all produced before the last barrier)
// so both levels' fp64 chains interleave in one instruction stream, with ONE consumer barrier per three planes.  The
// level-1 ring holds seven planes (three being written, four being read); an item runs three extra steps to drain.
//
// r01 ncu of sweep_tb2p on C3 (profiles/r01_c3_tb2p_128x14_ncu.txt): 42.5 thread instructions per point-update
// against 9 fp64 operations; the steady loop was 832 SASS instructions per three planes of which 216 DADD/DMUL,
// 96 FSEL (copy-through selects), 32 SEL, 45 ISETP, 22 BRA, ~250 IMAD/VIADD/LOP3/LEA/R2UR/LDC (masks, slot
// addresses, window rotation moves, shared-window base rebuilt from SR_CgaCtaId under register pressure).  Here
//   * the z windows of both levels live in THREE statically indexed register blocks, slot (step + q) mod 3, so a
//     group of three planes ends where it began and no register is moved;
//   * a group whose three steps are all in steady state AND whose planes all lie inside the apply's z bounds takes
//     the lean path: both levels are evaluated unconditionally in one basic block (eight independent fp64 chains
//     per thread), then ONE branch that only threads with a point outside the y/x bounds take applies the
//     copy-through selects, then the stores go out;
//   * shared memory is addressed by 32-bit shared-window addresses + immediate offsets (inline PTX), thread-constant
//     values are made opaque to the compiler so that they stay in registers instead of being recomputed per plane;
//   * row pitches are padded so that warps straddling two thread rows do not bank-conflict (Tb2qGeom);
//   * warps made of ring threads only run a copy of the loop without phase B';
//   * every other group (pipeline fill and drain, the first and last planes of the box) runs the fully masked step.
// Ring threads evaluate level 2 as well (its loads stay inside the guarded rings) and simply store nothing: a
// warp instruction costs the same with or without them, and the branch that used to skip them split the stream.
// Arithmetic is the body functor at both levels: bit-identical to two single sweeps.
// r02, C3 512^3 f64: 584 Gupd/s (sweep_tb2p: 466-476), 31.4 -> 29 thread instructions per update, shared wavefronts
// 95.9 M -> 80.2 M per launch (profiles/r02_c3_tb2q_ncu.txt).
#pragma once
#include "sweep_tb2m.cuh"

namespace neptune {

// GS = planes per barrier group.  Level-1 ring = 2*GS + 1 planes (GS being written, GS + 1 being read); phase B' lags
// phase A by GS + 1 planes.  GS = 3 needs no register rotation with three-slot z windows.
template <int TX_, int TY_, int VY_, int NS_, int MINB_, int GS_ = 3> struct Tb2pCfg : Tb2Cfg<TX_, TY_, VY_, NS_, MINB_> {
  static constexpr int GS = GS_;
};


// Shared-memory access by 32-bit shared-window address + immediate byte offset.  Generic pointers into dynamic
// shared memory made ptxas rebuild the window base (S2R SR_CgaCtaId + LEA) inside the plane loop when registers ran out.
template <int OFF> __device__ __forceinline__ void lds16(double (&d)[2], uint32_t addr) {
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2+%3];" : "=d"(d[0]), "=d"(d[1]) : "r"(addr), "n"(OFF));
}
template <int OFF> __device__ __forceinline__ void lds16(float (&d)[4], uint32_t addr) {
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+%5];" : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3]) : "r"(addr), "n"(OFF));
}
template <int OFF> __device__ __forceinline__ double lds1(uint32_t addr, double) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1+%2];" : "=d"(v) : "r"(addr), "n"(OFF));
  return v;
}
template <int OFF> __device__ __forceinline__ float lds1(uint32_t addr, float) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v) : "r"(addr), "n"(OFF));
  return v;
}
template <int OFF> __device__ __forceinline__ void sts16(uint32_t addr, const double (&d)[2]) {
  asm volatile("st.shared.v2.f64 [%0+%1], {%2, %3};" ::"r"(addr), "n"(OFF), "d"(d[0]), "d"(d[1]) : "memory");
}
template <int OFF> __device__ __forceinline__ void sts16(uint32_t addr, const float (&d)[4]) {
  asm volatile("st.shared.v4.f32 [%0+%1], {%2, %3, %4, %5};" ::"r"(addr), "n"(OFF), "f"(d[0]), "f"(d[1]), "f"(d[2]), "f"(d[3]) : "memory");
}
// one elected lane arrives, without a branch
__device__ __forceinline__ void mbar_arrive_if(uint32_t bar, bool pred) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %1, 0;\n\t@p mbarrier.arrive.shared::cta.b64 _, [%0];\n\t}" ::"r"(bar), "r"((uint32_t)pred) : "memory");
}

// in-plane halo of a VY x VX block at shared address `blk` (row pitch PB bytes) into the window c[][] (centre left to
// the caller); `lr` = blk + (swap ? 16 : -S), `rl` = blk + (swap ? -S : 16): lane groups of eight alternate the order of
// the left/right reads so that the 8-byte reads at a 16-byte lane stride do not bank-conflict (sweep_tb2m.cuh)
template <class T, int VX, int VY, int PB, bool CORNERS>
__device__ __forceinline__ void load_halo_s(T (&c)[VY + 2][VX + 2], uint32_t blk, uint32_t lr, uint32_t rl, bool swap) {
  T t[VX], b[VX];
  lds16<-PB>(t, blk);
  lds16<VY * PB>(b, blk);
#pragma unroll
  for (int i = 0; i < VX; ++i) { c[0][1 + i] = t[i]; c[VY + 1][1 + i] = b[i]; }
  if constexpr (CORNERS) {   // bodies with in-plane diagonal taps only
    constexpr int S = (int)sizeof(T);
    c[0][0] = lds1<-PB - S>(blk, T());
    c[0][VX + 1] = lds1<-PB + VX * S>(blk, T());
    c[VY + 1][0] = lds1<VY * PB - S>(blk, T());
    c[VY + 1][VX + 1] = lds1<VY * PB + VX * S>(blk, T());
  }
  static_for<VY>([&](auto rc) {
    constexpr int R = decltype(rc)::value;
    const T v1 = lds1<R * PB>(lr, T()), v2 = lds1<R * PB>(rl, T());
    c[1 + R][0] = swap ? v2 : v1;
    c[1 + R][VX + 1] = swap ? v1 : v2;
  });
}

// Row pitches.  A warp covers 32 consecutive thread blocks of the NBX-wide thread grid, so most warps straddle two
// thread rows (NBX = 66 for the 128-wide f64 tile); the 16-byte accesses of a quarter-warp then fall into two row
// segments VY shared-memory rows apart, and with the natural pitches those overlap in banks (r02 ncu of the first
// tb2q: 15.0 M of 95.9 M shared wavefronts per launch were bank-conflict replays, the shared pipe at 67 % being the
// busiest unit).  Padding the pitch so that VY * pitch = NBX * 16 (mod 128) makes the second row continue the bank
// sequence where the first one stopped: a straddling quarter-warp touches every bank once, as a contiguous one does.
constexpr int tb2q_pitch(int min_elems, int vx, int vy, int nbx, int s) {
  int p = min_elems;
  while (((vy * p * s - nbx * 16) % 128 + 128) % 128 != 0) p += vx;
  return p;
}

template <class F, class C> struct Tb2qGeom {
  typedef typename F::T T;
  typedef Tb2mGeom<F, C> B;
  static constexpr int S = sizeof(T);
  static constexpr int VX = B::VX, VY = B::VY, NBX = B::NBX, NBY = B::NBY;
  static constexpr int ACTIVE = B::ACTIVE, CONSUMERS = B::CONSUMERS, THREADS = B::THREADS, NS = B::NS;
  static constexpr int GS = C::GS, NL1 = 2 * GS + 1;
  static constexpr int P0 = tb2q_pitch(C::TX + 4 * VX, VX, VY, NBX, S);   // level-0 stage row pitch = TMA box width
  static constexpr int BOX_Y = C::TY + 4;
  static constexpr int STAGE_TX_BYTES = P0 * BOX_Y * S;
  static constexpr int STAGE_BYTES = (STAGE_TX_BYTES + 127) / 128 * 128;
  static constexpr int SLOT_BYTES = STAGE_BYTES;
  static constexpr int GUARD_BYTES = (P0 * S * (VY > 2 ? VY - 1 : 1) + 127) / 128 * 128;
  static constexpr int P1 = tb2q_pitch(NBX * VX, VX, VY, NBX, S), H1 = NBY * VY;   // level-1 plane
  static constexpr int L1_BYTES = (P1 * H1 * S + 127) / 128 * 128;
  static constexpr int L1_GUARD = (P1 * S + 127) / 128 * 128;
  static constexpr int SMEM_BYTES = GUARD_BYTES + NS * SLOT_BYTES + GUARD_BYTES + L1_GUARD + NL1 * L1_BYTES + L1_GUARD + 2 * NS * 8 + 128;
  static_assert(GS == 3, "three planes per barrier group");
  static_assert(F::NIN == 1, "single-input bodies");
  static_assert(P0 <= 256 && BOX_Y <= 256, "TMA box dimensions are limited to 256 elements");
  static_assert((P0 * S) % 16 == 0, "TMA box rows are multiples of 16 bytes");
  static_assert(SMEM_BYTES <= 227 * 1024, "rings do not fit in shared memory");
};

template <class F, class C, bool FMA>
__global__ void __launch_bounds__((Tb2qGeom<F, C>::THREADS), C::MINB)
sweep_tb2q(const __grid_constant__ KParams<F> p, const __grid_constant__ CUtensorMap tmap) {
  typedef Tb2qGeom<F, C> G;
  typedef typename F::T T;
  constexpr int VX = G::VX, VY = G::VY, NS = G::NS;
  constexpr int NCW = G::CONSUMERS / 32;
  constexpr int GS = G::GS, NL1 = G::NL1;
  static_assert(GS == 3, "the three-slot z windows need three planes per group");
  constexpr int LAG = GS + 2;
  constexpr unsigned ALL = (1u << (VY * VX)) - 1u;

  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char *smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  unsigned char *l0 = smem + G::GUARD_BYTES;
  unsigned char *l1 = l0 + NS * G::SLOT_BYTES + G::GUARD_BYTES + G::L1_GUARD;
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
  constexpr int S = (int)sizeof(T);
  constexpr int PB0 = G::P0 * S, PB1 = G::P1 * S;   // row pitches in bytes
  constexpr bool CORNERS = F::XYDIAG;
  const bool active = tid < G::ACTIVE;
  const int bx = active ? tid % G::NBX : 0, by = active ? tid / G::NBX : 0;
  const bool inner = active && bx >= 1 && bx <= G::NBX - 2 && by >= 1 && by <= G::NBY - 2;
  const bool lane0 = (tid & 31) == 0;
  const bool swap = ((tid >> 3) & 1) != 0;
  const bool ring_warp = __all_sync(0xffffffffu, !inner);   // warp-uniform
  // this thread's block inside level-0 slot 0 / level-1 plane 0, as shared-window byte addresses
  const uint32_t t0 = stage0 + (uint32_t)(((by * VY - VY + 2) * G::P0 + VX + bx * VX) * S);
  const uint32_t t1 = smem_u32(l1) + (uint32_t)(((by * VY) * G::P1 + bx * VX) * S);
  int o_lr = swap ? VX * S : -S, o_rl = swap ? -S : VX * S;   // first / second of the left-right halo reads
  uint32_t swp = swap ? 1u : 0u, ln0 = lane0 ? 1u : 0u, fullb = full0, emptyb = empty0;
  // opaque to the compiler from here on: under register pressure ptxas otherwise rebuilds these from SR_TID /
  // SR_CgaCtaId inside the plane loop (S2R + LOP3 + SEL + LEA per step)
  asm volatile("" : "+r"(o_lr), "+r"(o_rl), "+r"(swp), "+r"(ln0), "+r"(fullb), "+r"(emptyb));
  unsigned s = 0, ph = 0;   // level-0 ring slot / phase of the next plane to arrive
  unsigned c1 = 0;          // level-1 ring slot phase A writes at the current step
  // z windows, slot (R + q) mod 3 at step R of a group: a[.] = level-0 centre blocks of planes k-2, k-1, k;
  // b[.] = level-1 centre blocks of planes k-LAG-1, k-LAG, k-LAG+1
  T a[3][VY][VX], b[3][VY][VX];
#pragma unroll
  for (int q = 0; q < 3; ++q)
#pragma unroll
    for (int iy = 0; iy < VY; ++iy)
#pragma unroll
      for (int ix = 0; ix < VX; ++ix) a[q][iy][ix] = b[q][iy][ix] = T(0);
  T nopw[1][VY][VX];
#pragma unroll
  for (int iy = 0; iy < VY; ++iy)
#pragma unroll
    for (int ix = 0; ix < VX; ++ix) nopw[0][iy][ix] = T(0);

  auto slot = [](unsigned base, unsigned d) { const unsigned v = base + d; return v >= (unsigned)NL1 ? v - (unsigned)NL1 : v; };

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
    unsigned nb = ~bmask & ALL;          // points of this thread that copy through in (y, x)
    asm volatile("" : "+r"(nb), "+r"(wmask), "+r"(bmask));   // as integers, not as eight predicates recombined every step
    const int nslots = ze - zb + 4;
    // steps j whose phase-A plane (zb - 3 + j) and phase-B' plane (zb - 7 + j) both lie inside the z bounds
    // (and not the item's last arrival, whose stage is handed back at once)
    const int jz_lo = max(p.blo[0] - zb + 7, LAG + 2), jz_hi = min(p.bhi[0] - zb + 3, nslots - 1);
    const int g_lo = (jz_lo + GS - 1) / GS * GS;               // first lean group
    const int n_lean = jz_hi - g_lo >= GS ? (jz_hi - g_lo) / GS : 0;
    T *dst = p.out + (long long)zb * p.out_sz + (long long)gy * p.out_sy + gx;
    uint32_t prv = t0;        // this thread's block inside the stage of plane k-1
    unsigned s_prv = 0;

    for (int j0 = 0; j0 < nslots + GS; j0 += GS) {
      if (j0 == g_lo && n_lean > 0) {
        // ---- lean groups: three steady steps, every plane inside the z bounds ------------------------
        auto lean_step = [&](auto rc, auto with_b) {
          constexpr int R = decltype(rc)::value;
          constexpr bool WITH_B = decltype(with_b)::value != 0;
          constexpr int A0 = R % 3, A1 = (R + 1) % 3, A2 = (R + 2) % 3;
          T wa[VY + 2][VX + 2], wb[VY + 2][VX + 2];
          mbar_wait(fullb + 8 * s, ph);
          const uint32_t cur = t0 + s * (uint32_t)G::SLOT_BYTES;
          static_for<VY>([&](auto r) { lds16<decltype(r)::value * PB0>(a[A2][decltype(r)::value], cur); });
          load_halo_s<T, VX, VY, PB0, CORNERS>(wa, prv, prv + o_lr, prv + o_rl, swp != 0u);
          __syncwarp();
          mbar_arrive_if(emptyb + 8 * s_prv, ln0 != 0u);
          s_prv = s;
          prv = cur;
          if (++s == NS) { s = 0; ph ^= 1; }
          // level-1 reads of phase B' (planes produced before the last barrier).  They come AFTER the level-0 reads in
          // program order: issued ahead of the wait for the level-0 plane (they do not depend on it) they stay live
          // across it, and at 96 registers that spills (r02: 429 instead of 584 Gupd/s on C3)
          if constexpr (WITH_B) {
            const uint32_t pnew = t1 + slot(c1, GS + 1) * (uint32_t)G::L1_BYTES;   // plane k-LAG+1
            const uint32_t pctr = t1 + slot(c1, GS) * (uint32_t)G::L1_BYTES;       // plane k-LAG
            static_for<VY>([&](auto r) { lds16<decltype(r)::value * PB1>(b[A2][decltype(r)::value], pnew); });
            load_halo_s<T, VX, VY, PB1, CORNERS>(wb, pctr, pctr + o_lr, pctr + o_rl, swp != 0u);
          }
#pragma unroll
          for (int iy = 0; iy < VY; ++iy)
#pragma unroll
            for (int ix = 0; ix < VX; ++ix) {
              wa[1 + iy][1 + ix] = a[A1][iy][ix];
              if constexpr (WITH_B) wb[1 + iy][1 + ix] = b[A1][iy][ix];
            }
          // both levels, unconditionally, one basic block
          T ra[VY][VX], res[VY][VX];
          static_for<VY>([&](auto iyc) {
            constexpr int IY = decltype(iyc)::value;
            static_for<VX>([&](auto ixc) {
              constexpr int IX = decltype(ixc)::value;
              MarchAcc<F, VY, VX, IY, IX> acca{a[A0], wa, a[A2], nopw};
              ra[IY][IX] = F::template eval<FMA>(acca, p.c);
              if constexpr (WITH_B) {
                MarchAcc<F, VY, VX, IY, IX> accb{b[A0], wb, b[A2], nopw};
                res[IY][IX] = F::template eval<FMA>(accb, p.c);
              }
            });
          });
          if (nb != 0u) {   // threads on the y/x edge of the apply bounds only: copy-through of both levels
            asm volatile("" ::: "memory");   // keeps this a branch: if-converted it is 32 selects per step for every thread
#pragma unroll
            for (int iy = 0; iy < VY; ++iy)
#pragma unroll
              for (int ix = 0; ix < VX; ++ix)
                if ((nb >> (iy * VX + ix)) & 1u) {
                  ra[iy][ix] = a[A1][iy][ix];
                  if constexpr (WITH_B) res[iy][ix] = b[A1][iy][ix];
                }
          }
          const uint32_t l1p = t1 + c1 * (uint32_t)G::L1_BYTES;
          static_for<VY>([&](auto r) { sts16<decltype(r)::value * PB1>(l1p, ra[decltype(r)::value]); });
          if constexpr (WITH_B) {
            if (wmask == ALL) {
#pragma unroll
              for (int iy = 0; iy < VY; ++iy) stg_vec<T, VX, false>(dst + iy * p.out_sy, res[iy]);
            } else if (wmask != 0u) {   // tiles cut by the edge of the written box
#pragma unroll
              for (int iy = 0; iy < VY; ++iy)
#pragma unroll
                for (int ix = 0; ix < VX; ++ix)
                  if ((wmask >> (iy * VX + ix)) & 1u) dst[iy * p.out_sy + ix] = res[iy][ix];
            }
            dst += p.out_sz;
          }
          c1 = slot(c1, 1u);
          if constexpr (R == GS - 1) asm volatile("bar.sync 1, %0;" ::"n"(G::CONSUMERS) : "memory");
        };
        // warps made of ring threads only (the first and last thread rows: four of the twenty consumer warps of the
        // 128x14 tile) have no level-2 points: their own copy of the loop does phase A alone — no level-1 reads
        // (24 shared wavefronts per warp and step), no second evaluation
        if (ring_warp) {
          for (int g = n_lean; g > 0; --g)
            static_for<GS>([&](auto rc) { lean_step(rc, std::integral_constant<int, 0>{}); });
        } else {
          for (int g = n_lean; g > 0; --g)
            static_for<GS>([&](auto rc) { lean_step(rc, std::integral_constant<int, 1>{}); });
          // (dst advanced with the steps)
        }
        j0 += GS * n_lean;
      }
      // ---- general group: pipeline fill / drain and planes on the z edge of the bounds (fully masked) ----
      static_for<GS>([&](auto rc) {
        constexpr int R = decltype(rc)::value;
        constexpr int A0 = R % 3, A1 = (R + 1) % 3, A2 = (R + 2) % 3;
        const int j = j0 + R;
        const int k = zb - 2 + j;
        const bool arrive = j < nslots;                          // a level-0 plane arrives at this step
        const bool do_a = arrive && j >= 2;                      // ... and completes level-1 plane k-1
        const bool push_b = j >= LAG && j < nslots + GS;         // level-1 plane k-LAG+1 exists
        const bool do_b = j >= LAG + 2 && j < nslots + GS;       // level-2 plane k-LAG is wanted
        T wa[VY + 2][VX + 2], wb[VY + 2][VX + 2];
        if (arrive) {
          mbar_wait(fullb + 8 * s, ph);
          const uint32_t cur = t0 + s * (uint32_t)G::SLOT_BYTES;
          static_for<VY>([&](auto r) { lds16<decltype(r)::value * PB0>(a[A2][decltype(r)::value], cur); });
          if (do_a) load_halo_s<T, VX, VY, PB0, CORNERS>(wa, prv, prv + o_lr, prv + o_rl, swp != 0u);
          __syncwarp();
          mbar_arrive_if(emptyb + 8 * s_prv, ln0 != 0u && j >= 1);
          mbar_arrive_if(emptyb + 8 * s, ln0 != 0u && j == nslots - 1);
          s_prv = s;
          prv = cur;
          if (++s == NS) { s = 0; ph ^= 1; }
        }
        if (do_a) {
#pragma unroll
          for (int iy = 0; iy < VY; ++iy)
#pragma unroll
            for (int ix = 0; ix < VX; ++ix) wa[1 + iy][1 + ix] = a[A1][iy][ix];
          const unsigned m = (k - 1 >= p.blo[0] && k - 1 < p.bhi[0]) ? bmask : 0u;
          const uint32_t l1p = t1 + c1 * (uint32_t)G::L1_BYTES;
          static_for<VY>([&](auto iyc) {
            constexpr int IY = decltype(iyc)::value;
            T ra[VX];
            static_for<VX>([&](auto ixc) {
              constexpr int IX = decltype(ixc)::value;
              MarchAcc<F, VY, VX, IY, IX> acc{a[A0], wa, a[A2], nopw};
              const T v = F::template eval<FMA>(acc, p.c);
              ra[IX] = ((m >> (IY * VX + IX)) & 1u) ? v : a[A1][IY][IX];
            });
            sts16<IY * PB1>(l1p, ra);
          });
        }
        if (push_b && wmask != 0) {
          const uint32_t pnew = t1 + slot(c1, GS + 1) * (uint32_t)G::L1_BYTES;
          static_for<VY>([&](auto r) { lds16<decltype(r)::value * PB1>(b[A2][decltype(r)::value], pnew); });
          if (do_b) {
            const uint32_t pctr = t1 + slot(c1, GS) * (uint32_t)G::L1_BYTES;
            load_halo_s<T, VX, VY, PB1, CORNERS>(wb, pctr, pctr + o_lr, pctr + o_rl, swp != 0u);
          }
        }
        if (do_b) {
          if (wmask != 0) {
#pragma unroll
            for (int iy = 0; iy < VY; ++iy)
#pragma unroll
              for (int ix = 0; ix < VX; ++ix) wb[1 + iy][1 + ix] = b[A1][iy][ix];
            const unsigned m = (k - LAG >= p.blo[0] && k - LAG < p.bhi[0]) ? bmask : 0u;
            static_for<VY>([&](auto iyc) {
              constexpr int IY = decltype(iyc)::value;
              T res[VX];
              static_for<VX>([&](auto ixc) {
                constexpr int IX = decltype(ixc)::value;
                MarchAcc<F, VY, VX, IY, IX> acc{b[A0], wb, b[A2], nopw};
                const T v = F::template eval<FMA>(acc, p.c);
                res[IX] = ((m >> (IY * VX + IX)) & 1u) ? v : b[A1][IY][IX];
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
        if (j >= 2 && j < nslots + GS) c1 = slot(c1, 1u);
        if constexpr (R == GS - 1) asm volatile("bar.sync 1, %0;" ::"n"(G::CONSUMERS) : "memory");
      });
    }
  }
}

}  // namespace neptune
