// Host-side table of compiled sweep kernels, keyed by body structure (neptune_rt_body_key).
// Each generated translation unit (build/gen/body_*.cu) registers its functor's kernel variants
// here at load time; runtime/stencil.cpp looks them up in neptune_rt_stencil_create.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <deque>
#include <functional>
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
  std::function<int(const SweepArgs &)> launch;   // returns a cudaError_t (generated units: a launcher template; JIT bodies: a closure over the loaded kernel)
};

struct BodyEntry {
  uint64_t key = 0;
  int rank = 0, dtype = 0, nin = 0, nconst = 0, staged = -1, rz = 0, ry = 0, rx = 0, ntaps = 0, flops = 0;
  std::vector<Variant> variants;
};

// the table keeps addresses stable (stencils hold BodyEntry pointers while the JIT registers more bodies)
void register_body(BodyEntry &&e);
const BodyEntry *find_body(uint64_t key);
const std::deque<BodyEntry> &all_bodies();

// planes per work item: keep every CTA slot of the persistent grid busy for the same number of
// rounds (items / (rounds * slots)) while paying as little z-halo re-read per chunk as possible
inline int choose_zchunk(int planes, int ncols, int grid_slots, int rz) {
  if (rz == 0 || planes <= 1) return std::max(planes, 1);
  const int cand[] = {16, 24, 32, 48, 64, 96, 128, 192, 256, planes};
  double best = -1;
  int best_zc = planes;
  for (int zc : cand) {
    if (zc > planes) zc = planes;
    const long long chunks = (planes + zc - 1) / zc;
    const long long items = chunks * ncols;
    const long long rounds = (items + grid_slots - 1) / grid_slots;
    const double balance = (double)items / (double)(rounds * grid_slots);
    const double halo = (double)zc / (double)(zc + 2 * rz);
    const double score = balance * halo;
    if (score > best + 1e-9) { best = score; best_zc = zc; }
  }
  return best_zc;
}

// Bodies outside the prebuilt catalogue (runtime/jit.cpp): NVRTC-compile kernels/tiles.cuh + the generated functor for
// `desc`, load the cubin on `device` and register the variants; the cubin is cached on disk by key + header digest.
// Returns 0 or a NEPTUNE_RT_E* code with `err` filled. compile_only: stop after the cubin is built / found (no GPU needed).
int jit_body(const neptune_rt_stencil_desc &desc, int device, bool compile_only, std::string &err, double *seconds);

}  // namespace neptune
