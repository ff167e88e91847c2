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

template <class F, class C, bool FMA> int launch_tma(const SweepArgs &a) {
  typedef TmaGeom<F, C> G;
  auto kernel = sweep_tma<F, C, FMA>;
  static int occ_cached[16] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 16) dev = 0;
  if (occ_cached[dev] == 0) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, G::THREADS, G::SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    if (occ < 1) return (int)cudaErrorLaunchOutOfResources;
    occ_cached[dev] = occ;
  }
  if (!a.tmap) return (int)cudaErrorInvalidValue;
  KParams<F> p;
  fill_params<F>(p, a);
  const int wz = a.whi[0] - a.wlo[0];
  if (wz <= 0 || a.whi[1] <= a.wlo[1] || a.whi[2] <= a.wlo[2]) return 0;
  p.cxo = a.wlo[2] / C::TX;
  p.cyo = a.wlo[1] / C::TY;
  p.ncx = (a.whi[2] + C::TX - 1) / C::TX - p.cxo;
  p.ncy = (a.whi[1] + C::TY - 1) / C::TY - p.cyo;
  const int ncols = p.ncx * p.ncy;
  const int slots = a.sm_count * occ_cached[dev];
  p.zchunk = a.zchunk > 0 ? std::min(a.zchunk, wz) : choose_zchunk(wz, ncols, slots, F::RZ);
  p.nchunks = (wz + p.zchunk - 1) / p.zchunk;
  if (!set_gap(p, a, 0)) return (int)cudaErrorInvalidValue;
  const long long items = (long long)p.nchunks * ncols;
  if (items > 0x7fffffffLL) return (int)cudaErrorInvalidConfiguration;
  p.nitems = (int)items;
  const int grid = (int)std::min<long long>(items, slots);
  kernel<<<grid, G::THREADS, G::SMEM_BYTES, a.stream>>>(p, *a.tmap);
  return (int)cudaGetLastError();
}

// MARCH: 0 = sweep_tb2 (both levels through shared memory), 1 = sweep_tb2m (z-marching), 2 = sweep_tb2g (z-marching,
// grouped), 3 = sweep_tb2p (grouped, phase B software-pipelined into phase A), 4 = sweep_tb2q (tb2p with a lean steady state)
template <class F, class C, int MARCH> struct Tb2GeomOf { typedef Tb2Geom<F, C> type; };
template <class F, class C> struct Tb2GeomOf<F, C, 1> { typedef Tb2mGeom<F, C> type; };
template <class F, class C> struct Tb2GeomOf<F, C, 2> { typedef Tb2gGeom<F, C> type; };
template <class F, class C> struct Tb2GeomOf<F, C, 3> { typedef Tb2pGeom<F, C> type; };
template <class F, class C> struct Tb2GeomOf<F, C, 4> { typedef Tb2qGeom<F, C> type; };

template <class F, class C, bool FMA, int MARCH = 0> int launch_tb2(const SweepArgs &a) {
  typedef typename Tb2GeomOf<F, C, MARCH>::type G;
  auto kernel = [] {
    if constexpr (MARCH == 1) return sweep_tb2m<F, C, FMA>;
    else if constexpr (MARCH == 2) return sweep_tb2g<F, C, FMA>;
    else if constexpr (MARCH == 3) return sweep_tb2p<F, C, FMA>;
    else if constexpr (MARCH == 4) return sweep_tb2q<F, C, FMA>;
    else return sweep_tb2<F, C, FMA>;
  }();
  static int occ_cached[16] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 16) dev = 0;
  if (occ_cached[dev] == 0) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, G::THREADS, G::SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    if (occ < 1) return (int)cudaErrorLaunchOutOfResources;
    occ_cached[dev] = occ;
  }
  if (!a.tmap) return (int)cudaErrorInvalidValue;
  KParams<F> p;
  fill_params<F>(p, a);
  const int wz = a.whi[0] - a.wlo[0];
  if (wz <= 0 || a.whi[1] <= a.wlo[1] || a.whi[2] <= a.wlo[2]) return 0;
  p.cxo = a.wlo[2] / C::TX;
  p.cyo = a.wlo[1] / C::TY;
  p.ncx = (a.whi[2] + C::TX - 1) / C::TX - p.cxo;
  p.ncy = (a.whi[1] + C::TY - 1) / C::TY - p.cyo;
  const int ncols = p.ncx * p.ncy;
  const int slots = a.sm_count * occ_cached[dev];
  p.zchunk = a.zchunk > 0 ? std::min(a.zchunk, wz) : choose_zchunk(wz, ncols, slots, 2);
  p.nchunks = (wz + p.zchunk - 1) / p.zchunk;
  if (!set_gap(p, a, 0)) return (int)cudaErrorInvalidValue;
  const long long items = (long long)p.nchunks * ncols;
  if (items > 0x7fffffffLL) return (int)cudaErrorInvalidConfiguration;
  p.nitems = (int)items;
  const int grid = (int)std::min<long long>(items, slots);
  if constexpr (MARCH == 1) {
    if (F::NIN > 1 && !a.tmap_pw) return (int)cudaErrorInvalidValue;
    kernel<<<grid, G::THREADS, G::SMEM_BYTES, a.stream>>>(p, *a.tmap, a.tmap_pw ? *a.tmap_pw : *a.tmap);
  } else {
    kernel<<<grid, G::THREADS, G::SMEM_BYTES, a.stream>>>(p, *a.tmap);
  }
  return (int)cudaGetLastError();
}

// rank-2 row marching (sweep_rows.cuh): work item = (row chunk, strip of TX columns)
template <class F, class C, bool FMA> int launch_rows(const SweepArgs &a) {
  typedef RowGeom<F, C> G;
  auto kernel = sweep_rows<F, C, FMA>;
  static int occ_cached[16] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 16) dev = 0;
  if (occ_cached[dev] == 0) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, G::THREADS, G::SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    if (occ < 1) return (int)cudaErrorLaunchOutOfResources;
    occ_cached[dev] = occ;
  }
  KParams<F> p;
  fill_params<F>(p, a);
  const int wy = a.whi[1] - a.wlo[1];
  if (wy <= 0 || a.whi[2] <= a.wlo[2]) return 0;
  p.cxo = a.wlo[2] / C::TX;
  p.ncx = (a.whi[2] + C::TX - 1) / C::TX - p.cxo;
  p.ncy = 1;
  const int slots = a.sm_count * occ_cached[dev];
  p.zchunk = a.zchunk > 0 ? std::min(a.zchunk, wy) : choose_zchunk(wy, p.ncx, slots, 1);
  p.nchunks = (wy + p.zchunk - 1) / p.zchunk;
  if (!set_gap(p, a, 1)) return (int)cudaErrorInvalidValue;
  const long long items = (long long)p.nchunks * p.ncx;
  if (items > 0x7fffffffLL) return (int)cudaErrorInvalidConfiguration;
  p.nitems = (int)items;
  const int grid = (int)std::min<long long>(items, slots);
  kernel<<<grid, G::THREADS, G::SMEM_BYTES, a.stream>>>(p);
  return (int)cudaGetLastError();
}

template <class F, class C, bool FMA> Variant rows_variant(const char *name) {
  typedef RowGeom<F, C> G;
  Variant v;
  v.name = name;
  v.kind = VARIANT_ROWS;
  v.fma = FMA;
  v.tx = C::TX; v.ty = C::RB;
  v.halo_x = G::HX; v.halo_y = 1;
  v.threads = G::THREADS; v.smem = G::SMEM_BYTES;
  v.gap_dim = 1;
  v.launch = &launch_rows<F, C, FMA>;
  return v;
}

// rank-2 temporally blocked row marching (sweep_rows2.cuh)
template <class F, class C, bool FMA> int launch_rows2(const SweepArgs &a) {
  typedef Rows2Geom<F, C> G;
  auto kernel = sweep_rows2<F, C, FMA>;
  static int occ_cached[16] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 16) dev = 0;
  if (occ_cached[dev] == 0) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, G::THREADS, G::SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    if (occ < 1) return (int)cudaErrorLaunchOutOfResources;
    occ_cached[dev] = occ;
  }
  KParams<F> p;
  fill_params<F>(p, a);
  const int wy = a.whi[1] - a.wlo[1];
  if (wy <= 0 || a.whi[2] <= a.wlo[2]) return 0;
  p.cxo = a.wlo[2] / G::TX;
  p.ncx = (a.whi[2] + G::TX - 1) / G::TX - p.cxo;
  p.ncy = 1;
  const int slots = a.sm_count * occ_cached[dev];
  p.zchunk = a.zchunk > 0 ? std::min(a.zchunk, wy) : choose_zchunk(wy, p.ncx, slots, 2);
  p.nchunks = (wy + p.zchunk - 1) / p.zchunk;
  if (!set_gap(p, a, 1)) return (int)cudaErrorInvalidValue;
  const long long items = (long long)p.nchunks * p.ncx;
  if (items > 0x7fffffffLL) return (int)cudaErrorInvalidConfiguration;
  p.nitems = (int)items;
  const int grid = (int)std::min<long long>(items, slots);
  kernel<<<grid, G::THREADS, G::SMEM_BYTES, a.stream>>>(p);
  return (int)cudaGetLastError();
}

template <class F, class C, bool FMA> Variant rows2_variant(const char *name) {
  typedef Rows2Geom<F, C> G;
  Variant v;
  v.name = name;
  v.kind = VARIANT_TB2;
  v.needs_tmap = false;
  v.fma = FMA;
  v.steps = 2;
  v.tx = G::TX; v.ty = C::RB;
  v.halo_x = 2 * G::VX; v.halo_y = 2;
  v.threads = G::THREADS; v.smem = G::SMEM_BYTES;
  v.gap_dim = 1;
  v.launch = &launch_rows2<F, C, FMA>;
  return v;
}

template <class F, class C, bool FMA, int MARCH = 0> Variant tb2_variant(const char *name) {
  typedef typename Tb2GeomOf<F, C, MARCH>::type G;
  Variant v;
  v.name = name;
  v.kind = VARIANT_TB2;
  v.fma = FMA;
  v.steps = 2;
  v.tx = C::TX; v.ty = C::TY;
  v.box_x = G::P0; v.box_y = G::BOX_Y;
  v.halo_x = 2 * G::VX; v.halo_y = 2;
  v.threads = G::THREADS; v.smem = G::SMEM_BYTES;
  v.gap_dim = (MARCH <= 1 || MARCH == 4) ? 0 : -1;   // sweep_tb2, sweep_tb2m, sweep_tb2q
  v.launch = &launch_tb2<F, C, FMA, MARCH>;
  return v;
}

template <class F, class C, bool FMA> Variant tma_variant(const char *name) {
  typedef TmaGeom<F, C> G;
  Variant v;
  v.name = name;
  v.kind = VARIANT_TMA;
  v.fma = FMA;
  v.tx = C::TX; v.ty = C::TY;
  v.box_x = G::SXP; v.box_y = G::SY;
  v.halo_x = G::HX; v.halo_y = G::RY;
  v.threads = G::THREADS; v.smem = G::SMEM_BYTES;
  v.gap_dim = 0;
  v.launch = &launch_tma<F, C, FMA>;
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
    // the staged kernel copies through the staged centre value, so it needs input 0 staged
    if constexpr (F::STAGED == 0 && F::RANK >= 2 && F::RZ <= 1 && F::RY <= 1 && F::RX <= 1) {
      if constexpr (F::NIN > 4) {
        e.variants.push_back(tma_variant<F, Tile3W, false>("tma:64x8"));
        e.variants.push_back(tma_variant<F, Tile3W, true>("tma:64x8"));
        if constexpr (F::RZ > 0) {
          e.variants.push_back(tma_variant<F, Tile3W2, false>("tma:64x7"));
          e.variants.push_back(tma_variant<F, Tile3W3, false>("tma:64x14"));
        }
      } else if constexpr (F::RZ > 0) {
        e.variants.push_back(tma_variant<F, Tile3A, false>("tma:64x16"));
        e.variants.push_back(tma_variant<F, Tile3A, true>("tma:64x16"));
        e.variants.push_back(tma_variant<F, Tile3B, false>("tma:128x8"));
        e.variants.push_back(tma_variant<F, Tile3C, false>("tma:128x16"));
        if constexpr (F::NIN == 1 && F::RZ == 1) {
          e.variants.push_back(tb2_variant<F, Tb2A, false>("tb2:64x16"));
          e.variants.push_back(tb2_variant<F, Tb2A, true>("tb2:64x16"));
          e.variants.push_back(tb2_variant<F, Tb2B, false>("tb2:64x16n7"));
          e.variants.push_back(tb2_variant<F, Tb2C, false>("tb2:128x16"));
          e.variants.push_back(tb2_variant<F, Tb2D, false>("tb2:64x32"));
          if constexpr (F::ZSTAR) {
            e.variants.push_back(tb2_variant<F, Tb2mA, false, 1>("tb2m:128x14"));
            e.variants.push_back(tb2_variant<F, Tb2mA, true, 1>("tb2m:128x14"));
            e.variants.push_back(tb2_variant<F, Tb2mB, false, 1>("tb2m:64x16"));
            e.variants.push_back(tb2_variant<F, Tb2mC, false, 1>("tb2m:128x14n8"));
            e.variants.push_back(tb2_variant<F, Tb2mD, false, 1>("tb2m:64x24"));
            e.variants.push_back(tb2_variant<F, Tb2mA, false, 2>("tb2g:128x14"));
            e.variants.push_back(tb2_variant<F, Tb2mD, false, 2>("tb2g:64x24"));
            e.variants.push_back(tb2_variant<F, Tb2pA, false, 3>("tb2p:128x14"));
            e.variants.push_back(tb2_variant<F, Tb2pB, false, 3>("tb2p:64x24"));
            e.variants.push_back(tb2_variant<F, Tb2pA, true, 3>("tb2p:128x14"));
            e.variants.push_back(tb2_variant<F, Tb2pA, false, 4>("tb2q:128x14"));
            e.variants.push_back(tb2_variant<F, Tb2pA, true, 4>("tb2q:128x14"));
          }
        } else if constexpr (F::NIN == 2 && F::RZ == 1 && F::ZSTAR) {
          // input 0 plus pointwise inputs (Jacobi with a right-hand side): the z-marching form only
          e.variants.push_back(tb2_variant<F, Tb2mG, false, 1>("tb2m:64x24"));
          e.variants.push_back(tb2_variant<F, Tb2mG, true, 1>("tb2m:64x24"));
          e.variants.push_back(tb2_variant<F, Tb2mE, false, 1>("tb2m:128x10"));
          e.variants.push_back(tb2_variant<F, Tb2mF, false, 1>("tb2m:128x14"));
        }
      } else {
        if constexpr (F::RANK == 2 && F::RY == 1 && F::NIN <= 4) {
          e.variants.push_back(rows_variant<F, Rows512, false>("rows:512"));
          e.variants.push_back(rows_variant<F, Rows512, true>("rows:512"));
          e.variants.push_back(rows_variant<F, Rows256, false>("rows:256"));
          e.variants.push_back(rows_variant<F, Rows1024, false>("rows:1024"));
          if constexpr (F::NIN == 1) {
            e.variants.push_back(rows2_variant<F, Rows2A, false>("tb2r:256x3"));
            e.variants.push_back(rows2_variant<F, Rows2A, true>("tb2r:256x3"));
            e.variants.push_back(rows2_variant<F, Rows2B, false>("tb2r:256x6"));
            e.variants.push_back(rows2_variant<F, Rows2C, false>("tb2r:512x3"));
          }
        }
        e.variants.push_back(tma_variant<F, Tile2A, false>("tma:128x32"));
        e.variants.push_back(tma_variant<F, Tile2A, true>("tma:128x32"));
        e.variants.push_back(tma_variant<F, Tile2B, false>("tma:128x16"));
        e.variants.push_back(tma_variant<F, Tile2C, false>("tma:64x64"));
      }
    }
    register_body(std::move(e));
  }
};

}  // namespace neptune

#define NEPTUNE_REGISTER_BODY(F) static neptune::BodyRegistrar<F> neptune_registrar_##F;
