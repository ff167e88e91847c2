// Tile menus of the sweep kernels: device-only header, shared by the host registry (registry.cuh) and the units
// runtime/jit.cpp compiles with NVRTC for bodies outside the prebuilt catalogue.
#pragma once
#include "sweep_direct.cuh"
#include "sweep_tma.cuh"
#include "sweep_tb2.cuh"
#include "sweep_tb2m.cuh"
#include "sweep_tb2g.cuh"
#include "sweep_tb2p.cuh"
#include "sweep_tb2q.cuh"
#include "sweep_rows.cuh"
#include "sweep_rows2.cuh"

namespace neptune {

// tile menus (elements).  3-D bodies march in z; bodies without z taps (2-D) take bigger tiles.
typedef TileCfg<64, 16, 2, 8, 2> Tile3A;    // r01 sweep on C3: 348 Gupd/s (0.864 of the measured copy peak)
typedef TileCfg<128, 8, 2, 8, 2> Tile3B;
typedef TileCfg<128, 16, 4, 5, 2> Tile3C;
typedef Tb2Cfg<64, 16, 2, 5, 2> Tb2A;       // two time steps per launch (sweep_tb2.cuh)
typedef Tb2Cfg<64, 16, 2, 7, 1> Tb2B;
typedef Tb2Cfg<128, 16, 2, 5, 1> Tb2C;
typedef Tb2Cfg<64, 32, 4, 5, 1> Tb2D;
// z-marching form (sweep_tb2m.cuh).  Registers are granted per warp in units of 512, so a CTA of 704
// threads (128x16 tile) gets 80 per thread and spills; 128x14 is 640 threads -> 96 registers.
typedef Tb2Cfg<128, 14, 2, 6, 1> Tb2mA;
typedef Tb2Cfg<64, 16, 2, 6, 2> Tb2mB;
typedef Tb2Cfg<128, 14, 2, 8, 1> Tb2mC;
// 512 threads = 4 warps per SM sub-partition -> up to 128 registers (5 warps per sub-partition: 96)
typedef Tb2Cfg<64, 24, 2, 8, 1> Tb2mD;
typedef Tb2Cfg<128, 10, 2, 6, 1> Tb2mE;
typedef Tb2Cfg<64, 24, 2, 5, 1> Tb2mG;
typedef Tb2pCfg<128, 14, 2, 4, 1, 3> Tb2pA;   // seven level-1 planes leave room for four level-0 slots
typedef Tb2pCfg<64, 24, 2, 6, 1, 3> Tb2pB;

typedef Tb2Cfg<128, 14, 2, 4, 1> Tb2mF;   // two staged tiles per slot: four slots fit beside the level-1 planes
typedef TileCfg<64, 8, 1, 6, 1> Tile3W;   // many-input bodies (one coefficient field per tap)
typedef TileCfg<64, 7, 1, 6, 2> Tile3W2;  // 256 threads, two CTAs per SM at 128 registers
typedef TileCfg<64, 14, 1, 5, 1> Tile3W3; // 480 threads, one CTA per SM at 128 registers
// strip of TX columns, RB rows per stage (a multiple of the 3-row window, so the unrolled stage loop needs
// no register moves), ring of NS stages, CTAs per SM
typedef RowCfg<512, 3, 5, 3> Rows512;
typedef RowCfg<256, 6, 4, 4> Rows256;
typedef RowCfg<1024, 3, 4, 1> Rows1024;
typedef Rows2Cfg<256, 3, 5, 2> Rows2A;     // 256 consumer threads (strip of 508 f64 columns), 3 rows per stage, two CTAs per SM
typedef Rows2Cfg<256, 6, 4, 1> Rows2B;
typedef Rows2Cfg<512, 3, 4, 1> Rows2C;
typedef TileCfg<128, 32, 8, 3, 2> Tile2A;
typedef TileCfg<128, 16, 4, 4, 2> Tile2B;
typedef TileCfg<64, 64, 8, 3, 2> Tile2C;

}  // namespace neptune
