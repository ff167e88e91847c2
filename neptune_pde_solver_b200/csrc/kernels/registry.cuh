// Launchers + registration of the sweep kernels for one body functor.
// Included only by the generated units (build/gen/body_*.cu).
#pragma once
#include <algorithm>

#include "registry.h"
#include "tiles.cuh"

namespace neptune {

template <class F> void fill_params(KParams<F> &p, const SweepArgs &a) {
  typedef typename F::T T;
  p.out = static_cast<T *>(a.out);
  p.copy_src = static_cast<const T *>(a.copy_src);
  for (int k = 0; k < F::NIN; ++k) {
    p.in[k] = static_cast<const T *>(a.in[k]);
    p.in_sz[k] = a.in_sz[k];
    p.in_sy[k] = a.in_sy[k];
  }
  p.out_sz = a.out_sz; p.out_sy = a.out_sy;
  p.nz = a.nz; p.ny = a.ny; p.nx = a.nx;
  for (int d = 0; d < 3; ++d) { p.blo[d] = a.blo[d]; p.bhi[d] = a.bhi[d]; p.wlo[d] = a.wlo[d]; p.whi[d] = a.whi[d]; }
  p.zchunk = p.nchunks = p.ncx = p.ncy = p.cxo = p.cyo = p.nitems = p.zgap = 0;
  for (int i = 0; i < F::NCONST; ++i) p.c[i] = (T)a.consts[i];
  if (F::NCONST == 0) p.c[0] = T(0);
}

template <class F, bool FMA> int launch_direct(const SweepArgs &a) {
  KParams<F> p;
  fill_params<F>(p, a);
  const int wx = a.whi[2] - a.wlo[2], wy = a.whi[1] - a.wlo[1], wz = a.whi[0] - a.wlo[0];
  if (wx <= 0 || wy <= 0 || wz <= 0) return 0;
  dim3 grid((wx + DIRECT_BX - 1) / DIRECT_BX, (wy + DIRECT_BY - 1) / DIRECT_BY, wz);
  if (grid.y > 65535u || grid.z > 65535u) return (int)cudaErrorInvalidConfiguration;
  sweep_direct<F, FMA><<<grid, dim3(DIRECT_BX, DIRECT_BY, 1), 0, a.stream>>>(p);
  return (int)cudaGetLastError();
}

// two boundary ranges [wlo, gap_lo) and [gap_hi, whi) of the marching dimension `dim` as chunks 0 and 1 of one launch
template <class KP> inline bool set_gap(KP &p, const SweepArgs &a, int dim) {
  if (a.gap_hi <= a.gap_lo) return true;
  const int w = a.gap_lo - a.wlo[dim];
  if (w <= 0 || a.whi[dim] - a.gap_hi != w) return false;
  p.zchunk = w;
  p.nchunks = 2;
  p.zgap = a.gap_hi - a.gap_lo;
  return true;
}

// ---- one launcher for every persistent-item kernel ------------------------------------------------------------
// A kernel family is described by a traits struct: the geometry (THREADS, SMEM_BYTES, tile), the marching
// dimension of its work items (0: (z-chunk, column) items; 1: (row-chunk, strip) items of the rank-2 kernels), the
// halo its items re-read in that dimension, and how the kernel is called.
template <class F, class C, bool FMA> struct TmaK {
  typedef TmaGeom<F, C> G;
  static constexpr int MARCH_DIM = 0, ZHALO = F::RZ, TX = C::TX, TY = C::TY, KIND = VARIANT_TMA, STEPS = 1, GAP_DIM = 0;
  static constexpr bool TMAP = true;
  static constexpr int BOX_X = G::SXP, BOX_Y = G::SY, HALO_X = G::HX, HALO_Y = G::RY;
  static void call(int grid, const KParams<F> &p, const SweepArgs &a) { sweep_tma<F, C, FMA><<<grid, G::THREADS, G::SMEM_BYTES, a.stream>>>(p, *a.tmap); }
  static const void *fn() { return (const void *)sweep_tma<F, C, FMA>; }
};
template <class F, class C, bool FMA> struct Tb2K {        // both levels through shared memory (narrow boxes)
  typedef Tb2Geom<F, C> G;
  static constexpr int MARCH_DIM = 0, ZHALO = 2, TX = C::TX, TY = C::TY, KIND = VARIANT_TB2, STEPS = 2, GAP_DIM = 0;
  static constexpr bool TMAP = true;
  static constexpr int BOX_X = G::P0, BOX_Y = G::BOX_Y, HALO_X = 2 * G::VX, HALO_Y = 2;
  static void call(int grid, const KParams<F> &p, const SweepArgs &a) { sweep_tb2<F, C, FMA><<<grid, G::THREADS, G::SMEM_BYTES, a.stream>>>(p, *a.tmap); }
  static const void *fn() { return (const void *)sweep_tb2<F, C, FMA>; }
};
template <class F, class C, bool FMA> struct Tb2mK {       // z-marching; thin launches and bodies with a pointwise input
  typedef Tb2mGeom<F, C> G;
  static constexpr int MARCH_DIM = 0, ZHALO = 2, TX = C::TX, TY = C::TY, KIND = VARIANT_TB2, STEPS = 2, GAP_DIM = 0;
  static constexpr bool TMAP = true;
  static constexpr int BOX_X = G::P0, BOX_Y = G::BOX_Y, HALO_X = 2 * G::VX, HALO_Y = 2;
  static void call(int grid, const KParams<F> &p, const SweepArgs &a) {
    sweep_tb2m<F, C, FMA><<<grid, G::THREADS, G::SMEM_BYTES, a.stream>>>(p, *a.tmap, a.tmap_pw ? *a.tmap_pw : *a.tmap);
  }
  static const void *fn() { return (const void *)sweep_tb2m<F, C, FMA>; }
};
template <class F, class C, bool FMA> struct Tb2qK {       // z-marching, grouped, software-pipelined, lean steady state
  typedef Tb2qGeom<F, C> G;
  static constexpr int MARCH_DIM = 0, ZHALO = 2, TX = C::TX, TY = C::TY, KIND = VARIANT_TB2, STEPS = 2, GAP_DIM = 0;
  static constexpr bool TMAP = true;
  static constexpr int BOX_X = G::P0, BOX_Y = G::BOX_Y, HALO_X = 2 * G::VX, HALO_Y = 2;
  static void call(int grid, const KParams<F> &p, const SweepArgs &a) { sweep_tb2q<F, C, FMA><<<grid, G::THREADS, G::SMEM_BYTES, a.stream>>>(p, *a.tmap); }
  static const void *fn() { return (const void *)sweep_tb2q<F, C, FMA>; }
};
template <class F, class C, bool FMA> struct RowsK {       // rank 2: strip of TX columns marched down the rows
  typedef RowGeom<F, C> G;
  static constexpr int MARCH_DIM = 1, ZHALO = 1, TX = C::TX, TY = C::RB, KIND = VARIANT_ROWS, STEPS = 1, GAP_DIM = 1;
  static constexpr bool TMAP = false;
  static constexpr int BOX_X = 0, BOX_Y = 0, HALO_X = G::HX, HALO_Y = 1;
  static void call(int grid, const KParams<F> &p, const SweepArgs &a) { sweep_rows<F, C, FMA><<<grid, G::THREADS, G::SMEM_BYTES, a.stream>>>(p); }
  static const void *fn() { return (const void *)sweep_rows<F, C, FMA>; }
};
template <class F, class C, bool FMA> struct Rows2K {      // rank 2, two time steps per launch
  typedef Rows2Geom<F, C> G;
  static constexpr int MARCH_DIM = 1, ZHALO = 2, TX = G::TX, TY = C::RB, KIND = VARIANT_TB2, STEPS = 2, GAP_DIM = 1;
  static constexpr bool TMAP = false;
  static constexpr int BOX_X = 0, BOX_Y = 0, HALO_X = 2 * G::VX, HALO_Y = 2;
  static void call(int grid, const KParams<F> &p, const SweepArgs &a) { sweep_rows2<F, C, FMA><<<grid, G::THREADS, G::SMEM_BYTES, a.stream>>>(p); }
  static const void *fn() { return (const void *)sweep_rows2<F, C, FMA>; }
};

template <class F, class K> int launch_items(const SweepArgs &a) {
  typedef typename K::G G;
  static int occ_cached[16] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 16) dev = 0;
  if (occ_cached[dev] == 0) {
    cudaError_t e = cudaFuncSetAttribute(K::fn(), cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, K::fn(), G::THREADS, G::SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    if (occ < 1) return (int)cudaErrorLaunchOutOfResources;
    occ_cached[dev] = occ;
  }
  if (K::TMAP && !a.tmap) return (int)cudaErrorInvalidValue;
  if (K::TMAP && F::NIN > 1 && K::KIND == VARIANT_TB2 && !a.tmap_pw) return (int)cudaErrorInvalidValue;
  KParams<F> p;
  fill_params<F>(p, a);
  constexpr int M = K::MARCH_DIM;
  const int planes = a.whi[M] - a.wlo[M];
  if (a.whi[0] <= a.wlo[0] || a.whi[1] <= a.wlo[1] || a.whi[2] <= a.wlo[2]) return 0;
  p.cxo = a.wlo[2] / K::TX;
  p.ncx = (a.whi[2] + K::TX - 1) / K::TX - p.cxo;
  p.cyo = 0;
  p.ncy = 1;
  if (M == 0) {
    p.cyo = a.wlo[1] / K::TY;
    p.ncy = (a.whi[1] + K::TY - 1) / K::TY - p.cyo;
  }
  const int ncols = p.ncx * p.ncy;
  const int slots = a.sm_count * occ_cached[dev];
  p.zchunk = a.zchunk > 0 ? std::min(a.zchunk, planes) : choose_zchunk(planes, ncols, slots, K::ZHALO);
  p.nchunks = (planes + p.zchunk - 1) / p.zchunk;
  if (!set_gap(p, a, M)) return (int)cudaErrorInvalidValue;
  const long long items = (long long)p.nchunks * ncols;
  if (items > 0x7fffffffLL) return (int)cudaErrorInvalidConfiguration;
  p.nitems = (int)items;
  K::call((int)std::min<long long>(items, slots), p, a);
  return (int)cudaGetLastError();
}

template <class F, class K, bool FMA> Variant item_variant(const char *name) {
  typedef typename K::G G;
  Variant v;
  v.name = name;
  v.kind = K::KIND;
  v.fma = FMA;
  v.steps = K::STEPS;
  v.needs_tmap = K::TMAP;
  v.tx = K::TX; v.ty = K::TY;
  v.box_x = K::BOX_X; v.box_y = K::BOX_Y;
  v.halo_x = K::HALO_X; v.halo_y = K::HALO_Y;
  v.threads = G::THREADS; v.smem = G::SMEM_BYTES;
  v.gap_dim = K::GAP_DIM;
  v.launch = &launch_items<F, K>;
  return v;
}

template <class F> struct BodyRegistrar {
  BodyRegistrar() {
    BodyEntry e;
    e.key = F::KEY; e.rank = F::RANK; e.dtype = sizeof(typename F::T) == 8 ? NEPTUNE_RT_F64 : NEPTUNE_RT_F32;
    e.nin = F::NIN; e.nconst = F::NCONST; e.staged = F::STAGED;
    e.rz = F::RZ; e.ry = F::RY; e.rx = F::RX; e.ntaps = F::NTAPS; e.flops = F::FLOPS;
    Variant d;
    d.name = "direct"; d.kind = VARIANT_DIRECT; d.fma = false; d.threads = DIRECT_BX * DIRECT_BY;
    d.launch = &launch_direct<F, false>;
    e.variants.push_back(d);
    d.fma = true; d.launch = &launch_direct<F, true>;
    e.variants.push_back(d);
    // the staged kernel copies through the staged centre value, so it needs input 0 staged.
    // One tile per kernel family and shape class: the r01 tile sweeps (profiles/r01_*) found these and nothing else
    // ever won a comparison; the losers are gone from the binary (r01: 388 kernels, 38 MB).
    if constexpr (F::STAGED == 0 && F::RANK >= 2 && F::RZ <= 1 && F::RY <= 1 && F::RX <= 1) {
      if constexpr (F::NIN > 4) {            // one coefficient field per tap (C4)
        e.variants.push_back(item_variant<F, TmaK<F, Tile3W, false>, false>("tma:64x8"));
        e.variants.push_back(item_variant<F, TmaK<F, Tile3W, true>, true>("tma:64x8"));
      } else if constexpr (F::RZ > 0) {
        if constexpr (F::NIN == 1) {   // ring read through generic pointers: measured faster for single-input bodies (tiles.cuh)
          e.variants.push_back(item_variant<F, TmaK<F, Tile3Ag, false>, false>("tma:64x16"));
          e.variants.push_back(item_variant<F, TmaK<F, Tile3Ag, true>, true>("tma:64x16"));
        } else {
          e.variants.push_back(item_variant<F, TmaK<F, Tile3A, false>, false>("tma:64x16"));
          e.variants.push_back(item_variant<F, TmaK<F, Tile3A, true>, true>("tma:64x16"));
        }
        if constexpr (F::NIN == 1 && F::RZ == 1) {
          e.variants.push_back(item_variant<F, Tb2K<F, Tb2A, false>, false>("tb2:64x16"));      // boxes narrower than 256
          e.variants.push_back(item_variant<F, Tb2K<F, Tb2A, true>, true>("tb2:64x16"));
          if constexpr (F::ZSTAR) {
            e.variants.push_back(item_variant<F, Tb2mK<F, Tb2mA, false>, false>("tb2m:128x14"));   // thin launches (slab boundary planes)
            e.variants.push_back(item_variant<F, Tb2mK<F, Tb2mA, true>, true>("tb2m:128x14"));
            e.variants.push_back(item_variant<F, Tb2qK<F, Tb2pA, false>, false>("tb2q:128x14"));   // the default on wide, deep boxes (C3)
            e.variants.push_back(item_variant<F, Tb2qK<F, Tb2pA, true>, true>("tb2q:128x14"));
          }
        } else if constexpr (F::NIN == 2 && F::RZ == 1 && F::ZSTAR) {
          // input 0 plus a pointwise input (Jacobi with a right-hand side, C5): the z-marching form only
          e.variants.push_back(item_variant<F, Tb2mK<F, Tb2mG, false>, false>("tb2m:64x24"));
          e.variants.push_back(item_variant<F, Tb2mK<F, Tb2mG, true>, true>("tb2m:64x24"));
          e.variants.push_back(item_variant<F, Tb2mK<F, Tb2mF, false>, false>("tb2m:128x14"));
        }
      } else {
        if constexpr (F::RANK == 2 && F::RY == 1 && F::NIN <= 4) {
          e.variants.push_back(item_variant<F, RowsK<F, Rows512, false>, false>("rows:512"));
          e.variants.push_back(item_variant<F, RowsK<F, Rows512, true>, true>("rows:512"));
          e.variants.push_back(item_variant<F, RowsK<F, Rows256, false>, false>("rows:256"));
          e.variants.push_back(item_variant<F, RowsK<F, Rows1024, false>, false>("rows:1024"));
          if constexpr (F::NIN == 1) {
            e.variants.push_back(item_variant<F, Rows2K<F, Rows2A, false>, false>("tb2r:256x3"));
            e.variants.push_back(item_variant<F, Rows2K<F, Rows2A, true>, true>("tb2r:256x3"));
          }
        }
        e.variants.push_back(item_variant<F, TmaK<F, Tile2A, false>, false>("tma:128x32"));
        e.variants.push_back(item_variant<F, TmaK<F, Tile2A, true>, true>("tma:128x32"));
      }
    }
    register_body(std::move(e));
  }
};

}  // namespace neptune

#define NEPTUNE_REGISTER_BODY(F) static neptune::BodyRegistrar<F> neptune_registrar_##F;
