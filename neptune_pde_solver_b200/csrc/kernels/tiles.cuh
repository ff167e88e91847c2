// Tile menus of the sweep kernels: device-only header, shared by the host registry (registry.cuh) and the units
// runtime/jit.cpp compiles with NVRTC for bodies outside the prebuilt catalogue.
#pragma once
#include "sweep_direct.cuh"
#include "sweep_tma.cuh"
#include "sweep_tb2.cuh"
#include "sweep_tb2m.cuh"
#include "sweep_tb2q.cuh"
#include "sweep_rows.cuh"
#include "sweep_rows2.cuh"

namespace neptune {

// 3-D bodies march in z; bodies without z taps (2-D) take bigger tiles.  (TX, TY, VY, ring stages, min CTAs per SM)
typedef TileCfg<64, 16, 2, 8, 2> Tile3A;      // r01 sweep on C3: 0.87 of the measured copy peak
// The same tile with the ring read through generic pointers (LD.E) instead of LDS.  Measured A/B in one session (r02,
// profiles/r02_tma_lds_vs_generic.txt): single-input 7-point body on C3 344.5 (generic) vs 326.0 (LDS) Gupd/s; with a
// pointwise input (C5) 205.4 vs 210.1; 27 coefficient fields (C4) 24.3 vs 24.6.  Single-input bodies take the generic form.
typedef TileCfg<64, 16, 2, 8, 2, false, true> Tile3Ag;
typedef TileCfg<64, 8, 1, 6, 1> Tile3W;       // many-input bodies (one coefficient field per tap)
typedef TileCfg<128, 32, 8, 3, 2> Tile2A;
typedef Tb2Cfg<64, 16, 2, 5, 2> Tb2A;         // two time steps per launch, both levels through shared memory (sweep_tb2.cuh)
// z-marching forms.  Registers are granted per warp in units of 512: a CTA of 704 threads (128x16 tile) gets 80 per
// thread and spills; 128x14 is 640 threads -> 96 registers.  512 threads (64x24) get 128.
typedef Tb2Cfg<128, 14, 2, 6, 1> Tb2mA;
typedef Tb2Cfg<64, 24, 2, 5, 1> Tb2mG;        // two staged tiles per slot (input 0 + the pointwise input)
typedef Tb2Cfg<128, 14, 2, 4, 1> Tb2mF;
typedef Tb2pCfg<128, 14, 2, 4, 1, 3> Tb2pA;   // seven level-1 planes leave room for four level-0 slots
// strip of TX columns, RB rows per stage (a multiple of the 3-row window, so the unrolled stage loop needs no register
// moves), ring of NS stages, CTAs per SM
typedef RowCfg<512, 3, 5, 3> Rows512;
typedef RowCfg<256, 6, 4, 4> Rows256;
typedef RowCfg<1024, 3, 4, 1> Rows1024;
typedef Rows2Cfg<256, 3, 5, 2> Rows2A;        // 256 consumer threads (strip of 508 f64 columns), 3 rows per stage, two CTAs per SM

}  // namespace neptune
