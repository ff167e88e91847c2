// Host-side table of compiled sweep kernels, keyed by body structure (neptune_rt_body_key).
// Each generated translation unit (build/gen/body_*.cu) registers its functor's kernel variants
// here at load time; runtime/stencil.cpp looks them up in neptune_rt_stencil_create.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "neptune_rt.h"

namespace neptune {

// Type-erased launch arguments, canonical (z, y, x) coordinates (see kernels/common.cuh).
struct SweepArgs {
  void *out = nullptr;
  const void *copy_src = nullptr;   // input 0 storage origin (unshifted)
  const void *in[NEPTUNE_RT_MAX_INPUTS] = {};
  long long in_sz[NEPTUNE_RT_MAX_INPUTS] = {}, in_sy[NEPTUNE_RT_MAX_INPUTS] = {};
  long long out_sz = 0, out_sy = 0;
  int nz = 1, ny = 1, nx = 1;
  int blo[3] = {0, 0, 0}, bhi[3] = {0, 0, 0};
  int wlo[3] = {0, 0, 0}, whi[3] = {0, 0, 0};
  double consts[NEPTUNE_RT_MAX_CONSTS] = {};
  int zchunk = 0;        // planes per work item (0: choose)
  int gap_lo = 0, gap_hi = 0;   // planes [gap_lo, gap_hi) of the marching dimension are left out: the two equally wide
                                // ranges either side go out as the two chunks of ONE launch (slab boundary planes)
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  const CUtensorMap *tmap = nullptr;  // tensor map of the staged input (TMA variants)
  const CUtensorMap *tmap_pw = nullptr;  // tensor map of the pointwise input (two-input temporally blocked kernels)
};

enum VariantKind { VARIANT_DIRECT = 0, VARIANT_TMA = 1, VARIANT_TB2 = 2, VARIANT_ROWS = 3 };

struct Variant {
  std::string name;     // "direct", "tma:64x16", "tb2m:128x14", "rows:512", ...
  int kind = VARIANT_DIRECT;
  bool fma = false;
  int tx = 0, ty = 0;           // tile (TMA)
  int box_x = 0, box_y = 0;     // TMA box incl. halo, elements
  int halo_x = 0, halo_y = 0;   // box origin = tile origin - halo
  int threads = 0, smem = 0;
  int steps = 1;                // time steps one launch advances (2: temporally blocked)
  bool needs_tmap = true;       // TMA / TB2 kinds: the launch takes a tensor map of the staged input (row kernels use bulk copies)
  int gap_dim = -1;             // canonical dimension in which the launch honours SweepArgs::gap_lo/gap_hi (-1: it cannot)
  int (*launch)(const SweepArgs &) = nullptr;   // returns a cudaError_t
};

struct BodyEntry {
  uint64_t key = 0;
  int rank = 0, dtype = 0, nin = 0, nconst = 0, staged = -1, rz = 0, ry = 0, rx = 0, ntaps = 0, flops = 0;
  std::vector<Variant> variants;
};

void register_body(BodyEntry &&e);
const BodyEntry *find_body(uint64_t key);
const std::vector<BodyEntry> &all_bodies();

}  // namespace neptune
