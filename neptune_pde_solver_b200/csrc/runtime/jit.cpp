// Bodies outside the prebuilt kernel catalogue: compiled at run time with NVRTC.
//
// The reference lowers ANY apply body by cloning its ops into the loop nest (lib/Passes/DataflowLowering.cpp:413-423).
// The catalogue (build.py: catalogue/*.mlir -> one nvcc unit per body structure) covers the bodies of BASELINE.json's
// configs ahead of time; every other recognised body takes this path on first use:
//   1. frontend/codegen.cpp emits the point functor (the op list verbatim),
//   2. NVRTC compiles kernels/tiles.cuh + the functor for sm_100a with the hand-written sweep kernels instantiated
//      for it (direct, one TMA tile or the row marchers, and the temporally blocked forms where the body allows),
//   3. the cubin is cached on disk (key + digest of the kernel headers + NVRTC version) and loaded through the driver
//      API; the variants join the same registry the prebuilt bodies live in, each with a launch closure that packs
//      KParams<F> from the layout table the unit itself exports (sizeof / offsetof evaluated by the device compiler).
// A body the recogniser rejects still fails: this is not a fallback, it is the same kernels compiled later.
#include <cuda.h>
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <fcntl.h>
#include <nvrtc.h>
#include <sys/stat.h>
#include <unistd.h>

#include <array>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>

#include "frontend/ir.h"
#include "kernels/registry.h"

namespace neptune {
namespace {

// ---- NVRTC and the driver API, bound at run time (the library links neither) ----------------------------------
struct NvrtcApi {
  void *h = nullptr;
  nvrtcResult (*CreateProgram)(nvrtcProgram *, const char *, const char *, int, const char *const *, const char *const *) = nullptr;
  nvrtcResult (*DestroyProgram)(nvrtcProgram *) = nullptr;
  nvrtcResult (*CompileProgram)(nvrtcProgram, int, const char *const *) = nullptr;
  nvrtcResult (*GetCUBINSize)(nvrtcProgram, size_t *) = nullptr;
  nvrtcResult (*GetCUBIN)(nvrtcProgram, char *) = nullptr;
  nvrtcResult (*GetProgramLogSize)(nvrtcProgram, size_t *) = nullptr;
  nvrtcResult (*GetProgramLog)(nvrtcProgram, char *) = nullptr;
  nvrtcResult (*AddNameExpression)(nvrtcProgram, const char *) = nullptr;
  nvrtcResult (*GetLoweredName)(nvrtcProgram, const char *, const char **) = nullptr;
  nvrtcResult (*Version)(int *, int *) = nullptr;
  const char *(*GetErrorString)(nvrtcResult) = nullptr;
  bool ok = false;
};

NvrtcApi &nvrtc() {
  static NvrtcApi a;
  static bool tried = false;
  if (tried) return a;
  tried = true;
  const char *names[] = {"libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so"};
  for (const char *n : names)
    if ((a.h = dlopen(n, RTLD_NOW | RTLD_LOCAL))) break;
  if (!a.h) return a;
#define NB_SYM(f, s) a.f = reinterpret_cast<decltype(a.f)>(dlsym(a.h, s))
  NB_SYM(CreateProgram, "nvrtcCreateProgram");
  NB_SYM(DestroyProgram, "nvrtcDestroyProgram");
  NB_SYM(CompileProgram, "nvrtcCompileProgram");
  NB_SYM(GetCUBINSize, "nvrtcGetCUBINSize");
  NB_SYM(GetCUBIN, "nvrtcGetCUBIN");
  NB_SYM(GetProgramLogSize, "nvrtcGetProgramLogSize");
  NB_SYM(GetProgramLog, "nvrtcGetProgramLog");
  NB_SYM(AddNameExpression, "nvrtcAddNameExpression");
  NB_SYM(GetLoweredName, "nvrtcGetLoweredName");
  NB_SYM(Version, "nvrtcVersion");
  NB_SYM(GetErrorString, "nvrtcGetErrorString");
#undef NB_SYM
  a.ok = a.CreateProgram && a.DestroyProgram && a.CompileProgram && a.GetCUBINSize && a.GetCUBIN && a.GetProgramLogSize &&
         a.GetProgramLog && a.AddNameExpression && a.GetLoweredName && a.Version && a.GetErrorString;
  return a;
}

struct DriverApi {
  CUresult (*ModuleLoadData)(CUmodule *, const void *) = nullptr;
  CUresult (*ModuleGetFunction)(CUfunction *, CUmodule, const char *) = nullptr;
  CUresult (*ModuleGetGlobal)(CUdeviceptr *, size_t *, CUmodule, const char *) = nullptr;
  CUresult (*FuncSetAttribute)(CUfunction, CUfunction_attribute, int) = nullptr;
  CUresult (*OccupancyMaxActiveBlocksPerMultiprocessor)(int *, CUfunction, int, size_t) = nullptr;
  CUresult (*LaunchKernel)(CUfunction, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, CUstream, void **, void **) = nullptr;
  CUresult (*MemcpyDtoH)(void *, CUdeviceptr, size_t) = nullptr;
  bool ok = false;
};

DriverApi &driver() {
  static DriverApi d;
  static bool tried = false;
  if (tried) return d;
  tried = true;
  auto get = [](const char *name) -> void * {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
    return p;
  };
#define NB_DRV(f, s) d.f = reinterpret_cast<decltype(d.f)>(get(s))
  NB_DRV(ModuleLoadData, "cuModuleLoadData");
  NB_DRV(ModuleGetFunction, "cuModuleGetFunction");
  NB_DRV(ModuleGetGlobal, "cuModuleGetGlobal");
  NB_DRV(FuncSetAttribute, "cuFuncSetAttribute");
  NB_DRV(OccupancyMaxActiveBlocksPerMultiprocessor, "cuOccupancyMaxActiveBlocksPerMultiprocessor");
  NB_DRV(LaunchKernel, "cuLaunchKernel");
  NB_DRV(MemcpyDtoH, "cuMemcpyDtoH");
#undef NB_DRV
  d.ok = d.ModuleLoadData && d.ModuleGetFunction && d.ModuleGetGlobal && d.FuncSetAttribute &&
         d.OccupancyMaxActiveBlocksPerMultiprocessor && d.LaunchKernel && d.MemcpyDtoH;
  return d;
}

// ---- what gets compiled for a body -------------------------------------------------------------------------
enum JitLaunch { JL_DIRECT = 0, JL_TILE = 1, JL_ROWS = 2 };   // grid over points / (z-chunk, column) items / (row-chunk, strip) items

struct KernelSpec {
  std::string name;       // variant name ("tma:64x16")
  int kind;               // VARIANT_*
  bool fma;
  std::string expr;       // name expression of the instantiation
  std::string geom;       // C++ initialiser of its geometry row: threads, smem, tx, ty, box_x, box_y, halo_x, halo_y
  int launch;             // JitLaunch
  int steps = 1, gap_dim = -1, zhalo = 0;
  bool needs_tmap = false, two_maps = false;
};

std::vector<KernelSpec> kernel_menu(const neptune_rt_stencil_desc &d, const BodyInfo &bi, const std::string &F, bool zstar) {
  std::vector<KernelSpec> v;
  auto add = [&](const char *name, int kind, bool fma, const std::string &kernel, const std::string &geom_t, const std::string &cfg,
                 const char *geom_fmt, int launch, int steps, int gap_dim, int zhalo, bool tmap, bool two) {
    KernelSpec k;
    k.name = name; k.kind = kind; k.fma = fma; k.launch = launch; k.steps = steps; k.gap_dim = gap_dim; k.zhalo = zhalo;
    k.needs_tmap = tmap; k.two_maps = two;
    k.expr = "neptune::" + kernel + "<" + F + (cfg.empty() ? "" : ", neptune::" + cfg) + ", " + (fma ? "true" : "false") + ">";
    std::string g = geom_fmt;   // G = geometry type, C = config type
    const std::string G = geom_t.empty() ? "" : "neptune::" + geom_t + "<" + F + ", neptune::" + cfg + ">", C = "neptune::" + cfg;
    for (size_t p; (p = g.find("$G")) != std::string::npos;) g.replace(p, 2, G);
    for (size_t p; (p = g.find("$C")) != std::string::npos;) g.replace(p, 2, C);
    k.geom = g;
    v.push_back(k);
  };
  const char *G_TMA = "{$G::THREADS, $G::SMEM_BYTES, $C::TX, $C::TY, $G::SXP, $G::SY, $G::HX, $G::RY}";
  const char *G_TB2 = "{$G::THREADS, $G::SMEM_BYTES, $C::TX, $C::TY, $G::P0, $G::BOX_Y, 2 * $G::VX, 2}";
  const char *G_ROWS = "{$G::THREADS, $G::SMEM_BYTES, $C::TX, $C::RB, 0, 0, $G::HX, 1}";
  const char *G_ROWS2 = "{$G::THREADS, $G::SMEM_BYTES, $G::TX, $C::RB, 0, 0, 2 * $G::VX, 2}";
  add("direct", VARIANT_DIRECT, false, "sweep_direct", "", "", "{neptune::DIRECT_BX * neptune::DIRECT_BY, 0, 0, 0, 0, 0, 0, 0}", JL_DIRECT, 1, -1, 0, false, false);
  add("direct", VARIANT_DIRECT, true, "sweep_direct", "", "", "{neptune::DIRECT_BX * neptune::DIRECT_BY, 0, 0, 0, 0, 0, 0, 0}", JL_DIRECT, 1, -1, 0, false, false);
  int rz = 0, ry = 0, rx = 0;
  if (bi.staged >= 0) { rz = bi.radius[bi.staged][0]; ry = bi.radius[bi.staged][1]; rx = bi.radius[bi.staged][2]; }
  // the staged kernels copy through the staged centre value: input 0 must be the staged input (kernels/registry.cuh)
  if (bi.staged == 0 && d.rank >= 2 && rz <= 1 && ry <= 1 && rx <= 1) {
    if (d.n_inputs > 4) {
      for (bool fma : {false, true}) add("tma:64x8", VARIANT_TMA, fma, "sweep_tma", "TmaGeom", "Tile3W", G_TMA, JL_TILE, 1, 0, rz, true, false);
    } else if (rz > 0) {
      for (bool fma : {false, true})
        add("tma:64x16", VARIANT_TMA, fma, "sweep_tma", "TmaGeom", d.n_inputs == 1 ? "Tile3Ag" : "Tile3A", G_TMA, JL_TILE, 1, 0, rz, true, false);
      if (d.n_inputs == 1 && rz == 1) {
        add("tb2:64x16", VARIANT_TB2, false, "sweep_tb2", "Tb2Geom", "Tb2A", G_TB2, JL_TILE, 2, 0, 2, true, false);
        if (zstar) {
          add("tb2m:128x14", VARIANT_TB2, false, "sweep_tb2m", "Tb2mGeom", "Tb2mA", G_TB2, JL_TILE, 2, 0, 2, true, true);
          add("tb2q:128x14", VARIANT_TB2, false, "sweep_tb2q", "Tb2qGeom", "Tb2pA", G_TB2, JL_TILE, 2, 0, 2, true, false);
        }
      } else if (d.n_inputs == 2 && rz == 1 && zstar) {
        add("tb2m:64x24", VARIANT_TB2, false, "sweep_tb2m", "Tb2mGeom", "Tb2mG", G_TB2, JL_TILE, 2, 0, 2, true, true);
      }
    } else {
      if (d.rank == 2 && ry == 1 && d.n_inputs <= 4) {
        add("rows:512", VARIANT_ROWS, false, "sweep_rows", "RowGeom", "Rows512", G_ROWS, JL_ROWS, 1, 1, 1, false, false);
        add("rows:256", VARIANT_ROWS, false, "sweep_rows", "RowGeom", "Rows256", G_ROWS, JL_ROWS, 1, 1, 1, false, false);
        add("rows:1024", VARIANT_ROWS, false, "sweep_rows", "RowGeom", "Rows1024", G_ROWS, JL_ROWS, 1, 1, 1, false, false);
        if (d.n_inputs == 1) add("tb2r:256x3", VARIANT_TB2, false, "sweep_rows2", "Rows2Geom", "Rows2A", G_ROWS2, JL_ROWS, 2, 1, 2, false, false);
      }
      for (bool fma : {false, true}) add("tma:128x32", VARIANT_TMA, fma, "sweep_tma", "TmaGeom", "Tile2A", G_TMA, JL_TILE, 1, 0, rz, true, false);
    }
  }
  return v;
}

const char *LAYOUT_FIELDS[] = {"out", "copy_src", "in", "in_sz", "in_sy", "out_sz", "out_sy", "nz", "ny", "nx", "blo", "bhi", "wlo", "whi",
                               "zchunk", "nchunks", "ncx", "ncy", "cxo", "cyo", "nitems", "zgap", "c"};
constexpr int N_LAYOUT = 1 + (int)(sizeof(LAYOUT_FIELDS) / sizeof(LAYOUT_FIELDS[0]));   // sizeof first, then the offsets

std::string unit_source(const neptune_rt_stencil_desc &d, const std::vector<KernelSpec> &menu, const std::string &F) {
  std::ostringstream o;
  o << "// NVRTC unit of one apply body (runtime/jit.cpp)\n#include \"kernels/tiles.cuh\"\n\n" << generate_functor(d) << "\n";
  o << "extern \"C\" __device__ const int neptune_jit_geom[" << menu.size() << "][8] = {\n";
  for (auto &k : menu) o << "  " << k.geom << ",\n";
  o << "};\n";
  // sizeof / offsetof of KParams<F> as the device compiler lays it out (NVRTC has no offsetof: one thread measures it)
  o << "typedef neptune::KParams<" << F << "> neptune_jit_kp;\n";
  o << "extern \"C\" __global__ void neptune_jit_layout(int *o) {\n  const neptune_jit_kp *z = nullptr;\n  o[0] = (int)sizeof(neptune_jit_kp);\n";
  int idx = 1;
  for (const char *f : LAYOUT_FIELDS) o << "  o[" << idx++ << "] = (int)(unsigned long long)(const void *)&z->" << f << ";\n";
  o << "}\n";
  return o.str();
}

std::string csrc_dir() {   // .../neptune_pde_solver_b200/csrc, next to lib/libneptune_rt_b200.so; NEPTUNE_B200_CSRC overrides
  if (const char *e = std::getenv("NEPTUNE_B200_CSRC")) return e;
  Dl_info info;
  if (dladdr((void *)&csrc_dir, &info) && info.dli_fname) {
    std::string p = info.dli_fname;
    size_t s = p.rfind('/');
    if (s != std::string::npos) p = p.substr(0, s);            // .../lib
    s = p.rfind('/');
    if (s != std::string::npos) return p.substr(0, s) + "/csrc";
  }
  return "csrc";
}

uint64_t fnv(uint64_t h, const void *p, size_t n) {
  const unsigned char *b = (const unsigned char *)p;
  for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; }
  return h;
}

bool read_file(const std::string &path, std::string &out) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::ostringstream ss;
  ss << f.rdbuf();
  out = ss.str();
  return true;
}

uint64_t header_digest(const std::string &csrc) {
  static const char *files[] = {"common.cuh", "tiles.cuh", "sweep_direct.cuh", "sweep_tma.cuh", "sweep_tb2.cuh", "sweep_tb2m.cuh", "sweep_tb2q.cuh", "sweep_rows.cuh", "sweep_rows2.cuh"};
  uint64_t h = 1469598103934665603ull;
  for (const char *f : files) {
    std::string s;
    if (read_file(csrc + "/kernels/" + f, s)) h = fnv(h, s.data(), s.size());
  }
  int maj = 0, min = 0;
  if (nvrtc().ok) nvrtc().Version(&maj, &min);
  h = fnv(h, &maj, sizeof maj);
  h = fnv(h, &min, sizeof min);
  return h;
}

// The on-disk cache holds code that will be loaded into the GPU context, so it must be a directory only this user can
// write: $NEPTUNE_B200_JIT_CACHE, else $HOME/.cache/neptune_b200_jit, else /tmp/neptune_b200_jit-<uid>; created with
// mode 0700 and then checked (a real directory, owned by this user, no group/other write bit). If it is not, the cache
// is simply off (empty string): every process compiles for itself.
std::string cache_dir() {
  std::string d;
  if (const char *e = std::getenv("NEPTUNE_B200_JIT_CACHE")) d = e;
  else if (const char *h = std::getenv("HOME")) d = std::string(h) + "/.cache/neptune_b200_jit";
  else d = "/tmp/neptune_b200_jit-" + std::to_string((long)geteuid());
  while (d.size() > 1 && d.back() == '/') d.pop_back();
  std::string acc;
  for (size_t i = 0; i <= d.size(); ++i)
    if (i == d.size() || d[i] == '/') {
      if (!acc.empty()) mkdir(acc.c_str(), i == d.size() ? 0700 : 0755);
      if (i < d.size()) acc += '/';
    } else {
      acc += d[i];
    }
  struct stat st;
  if (lstat(d.c_str(), &st) != 0 || !S_ISDIR(st.st_mode) || st.st_uid != geteuid() || (st.st_mode & (S_IWGRP | S_IWOTH))) return "";
  return d;
}

// One cache file per body: magic, digest of the generated source + compile options, the lowered kernel names, the
// cubin, and a checksum of all of that. Names and code cannot be paired wrongly (one rename), a truncated or altered
// file fails its checksum, and a file made from other source or options (a body_key collision, a change to the option
// list below) fails the source digest: all of these read as a miss.
const char CACHE_MAGIC[8] = {'N', 'B', 'J', 'I', 'T', '0', '0', '1'};

bool cache_load(const std::string &path, uint64_t src_digest, size_t n_kernels, std::string &cubin, std::vector<std::string> &lowered) {
  std::string blob;
  if (!read_file(path, blob) || blob.size() < sizeof CACHE_MAGIC + 5 * 8) return false;
  if (std::memcmp(blob.data(), CACHE_MAGIC, sizeof CACHE_MAGIC) != 0) return false;
  uint64_t hdr[3], tail = 0;   // source digest, names bytes, cubin bytes
  std::memcpy(hdr, blob.data() + 8, sizeof hdr);
  const size_t body = 8 + sizeof hdr;
  if (hdr[0] != src_digest || hdr[1] > blob.size() || hdr[2] > blob.size() || body + hdr[1] + hdr[2] + 8 != blob.size()) return false;
  std::memcpy(&tail, blob.data() + blob.size() - 8, 8);
  if (tail != fnv(1469598103934665603ull, blob.data(), blob.size() - 8)) return false;
  std::istringstream is(blob.substr(body, (size_t)hdr[1]));
  std::string l;
  lowered.clear();
  while (std::getline(is, l))
    if (!l.empty()) lowered.push_back(l);
  if (lowered.size() != n_kernels || hdr[2] == 0) { lowered.clear(); return false; }
  cubin = blob.substr(body + (size_t)hdr[1], (size_t)hdr[2]);
  return true;
}

void cache_store(const std::string &path, uint64_t src_digest, const std::string &cubin, const std::vector<std::string> &lowered) {
  std::string names;
  for (auto &l : lowered) names += l + "\n";
  const uint64_t hdr[3] = {src_digest, (uint64_t)names.size(), (uint64_t)cubin.size()};
  std::string blob(CACHE_MAGIC, sizeof CACHE_MAGIC);
  blob.append((const char *)hdr, sizeof hdr);
  blob += names;
  blob += cubin;
  const uint64_t tail = fnv(1469598103934665603ull, blob.data(), blob.size());
  blob.append((const char *)&tail, 8);
  // best effort; written under a temporary name (created exclusively, mode 0600) and renamed, so that a concurrent
  // rank never reads half a file
  const std::string tmp = path + ".tmp" + std::to_string((long)getpid());
  const int fd = open(tmp.c_str(), O_WRONLY | O_CREAT | O_EXCL | O_NOFOLLOW, 0600);
  if (fd < 0) return;
  size_t done = 0;
  while (done < blob.size()) {
    const ssize_t w = write(fd, blob.data() + done, blob.size() - done);
    if (w <= 0) break;
    done += (size_t)w;
  }
  close(fd);
  if (done == blob.size()) std::rename(tmp.c_str(), path.c_str());
  else unlink(tmp.c_str());
}

// cubin + lowered kernel names, from the disk cache or from NVRTC
struct Compiled {
  std::string cubin;
  std::vector<std::string> lowered;
  bool from_cache = false;
};

int compile_unit(const neptune_rt_stencil_desc &d, const std::vector<KernelSpec> &menu, const std::string &F, uint64_t key,
                 Compiled &out, std::string &err) {
  const std::string csrc = csrc_dir();
  char tag[64];
  std::snprintf(tag, sizeof tag, "%016llx_%016llx", (unsigned long long)key, (unsigned long long)header_digest(csrc));
  const std::string dir = cache_dir();
  const std::string path = dir.empty() ? std::string() : dir + "/body_" + tag + ".nbjit";
  const std::string src = unit_source(d, menu, F);
  const std::string inc = "-I" + csrc;
  const char *opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo", "-default-device", inc.c_str()};
  uint64_t src_digest = fnv(1469598103934665603ull, src.data(), src.size());
  for (const char *o : opts) src_digest = fnv(src_digest, o, std::strlen(o) + 1);
  if (!path.empty() && cache_load(path, src_digest, menu.size(), out.cubin, out.lowered)) { out.from_cache = true; return NEPTUNE_RT_OK; }
  NvrtcApi &n = nvrtc();
  if (!n.ok) {
    err = "this apply body is not in the prebuilt kernel catalogue and libnvrtc.so.12 could not be loaded to compile it";
    return NEPTUNE_RT_EUNSUPPORTED;
  }
  nvrtcProgram prog;
  nvrtcResult r = n.CreateProgram(&prog, src.c_str(), ("body_" + std::string(tag) + ".cu").c_str(), 0, nullptr, nullptr);
  if (r != NVRTC_SUCCESS) { err = std::string("nvrtcCreateProgram: ") + n.GetErrorString(r); return NEPTUNE_RT_ECUDA; }
  for (auto &k : menu) n.AddNameExpression(prog, k.expr.c_str());
  r = n.CompileProgram(prog, (int)(sizeof opts / sizeof opts[0]), opts);
  if (r != NVRTC_SUCCESS) {
    size_t ls = 0;
    n.GetProgramLogSize(prog, &ls);
    std::string log(ls, '\0');
    if (ls) n.GetProgramLog(prog, &log[0]);
    err = std::string("NVRTC could not compile the body (") + n.GetErrorString(r) + "):\n" + log.substr(0, 3000);
    n.DestroyProgram(&prog);
    return NEPTUNE_RT_ECUDA;
  }
  for (auto &k : menu) {
    const char *low = nullptr;
    if (n.GetLoweredName(prog, k.expr.c_str(), &low) != NVRTC_SUCCESS || !low) {
      err = "nvrtcGetLoweredName failed for " + k.expr;
      n.DestroyProgram(&prog);
      return NEPTUNE_RT_ECUDA;
    }
    out.lowered.push_back(low);
  }
  size_t cs = 0;
  n.GetCUBINSize(prog, &cs);
  out.cubin.assign(cs, '\0');
  n.GetCUBIN(prog, &out.cubin[0]);
  n.DestroyProgram(&prog);
  if (!path.empty()) cache_store(path, src_digest, out.cubin, out.lowered);
  return NEPTUNE_RT_OK;
}

// ---- a loaded body: one module per device, launch closures over its kernels ------------------------------------
struct JitBody {
  uint64_t key = 0;
  int nin = 0, nconst = 0, elsize = 8, zradius = 0;
  std::vector<KernelSpec> menu;
  Compiled code;
  struct PerDevice {
    CUmodule mod = nullptr;
    std::vector<CUfunction> fn;
    std::vector<int> occ;                 // resident CTAs per SM of each kernel (0: not asked yet)
    std::vector<std::array<int, 8>> geom; // threads, smem, tx, ty, box_x, box_y, halo_x, halo_y
    int layout[N_LAYOUT] = {};
  };
  std::map<int, PerDevice> dev;
};

int load_on_device(JitBody &b, int device, std::string &err) {
  if (b.dev.count(device)) return NEPTUNE_RT_OK;
  DriverApi &dr = driver();
  if (!dr.ok) { err = "the CUDA driver entry points for module loading are not available"; return NEPTUNE_RT_ECUDA; }
  cudaSetDevice(device);
  cudaFree(nullptr);   // makes the primary context current
  JitBody::PerDevice pd;
  CUresult r = dr.ModuleLoadData(&pd.mod, b.code.cubin.data());
  if (r != CUDA_SUCCESS) { err = "cuModuleLoadData of the JIT-compiled body failed with CUresult " + std::to_string((int)r); return NEPTUNE_RT_ECUDA; }
  const size_t n = b.menu.size();
  pd.fn.resize(n); pd.occ.assign(n, 0); pd.geom.resize(n);
  for (size_t i = 0; i < n; ++i)
    if ((r = dr.ModuleGetFunction(&pd.fn[i], pd.mod, b.code.lowered[i].c_str())) != CUDA_SUCCESS) {
      err = "cuModuleGetFunction(" + b.menu[i].expr + ") failed with CUresult " + std::to_string((int)r);
      return NEPTUNE_RT_ECUDA;
    }
  CUdeviceptr gp = 0;
  size_t gs = 0;
  if ((r = dr.ModuleGetGlobal(&gp, &gs, pd.mod, "neptune_jit_geom")) != CUDA_SUCCESS || gs != n * 8 * sizeof(int) ||
      (r = dr.MemcpyDtoH(pd.geom.data(), gp, gs)) != CUDA_SUCCESS) {
    err = "the JIT unit's geometry table could not be read";
    return NEPTUNE_RT_ECUDA;
  }
  {
    CUfunction lf = nullptr;
    int *dl = nullptr;
    void *args[1] = {&dl};
    if (dr.ModuleGetFunction(&lf, pd.mod, "neptune_jit_layout") != CUDA_SUCCESS || cudaMalloc(&dl, sizeof pd.layout) != cudaSuccess ||
        dr.LaunchKernel(lf, 1, 1, 1, 1, 1, 1, 0, nullptr, args, nullptr) != CUDA_SUCCESS ||
        cudaMemcpy(pd.layout, dl, sizeof pd.layout, cudaMemcpyDeviceToHost) != cudaSuccess) {
      if (dl) cudaFree(dl);
      err = "the JIT unit's parameter layout could not be measured";
      return NEPTUNE_RT_ECUDA;
    }
    cudaFree(dl);
    if (pd.layout[0] <= 0 || pd.layout[0] > 4096) { err = "the JIT unit reports an impossible parameter size"; return NEPTUNE_RT_ECUDA; }
  }
  b.dev[device] = std::move(pd);
  return NEPTUNE_RT_OK;
}

// KParams<F>, packed from the unit's own sizeof / offsetof table
struct ParamPack {
  std::vector<unsigned char> buf;
  const int *lay;
  explicit ParamPack(const int *layout) : buf((size_t)layout[0], 0), lay(layout) {}
  template <class V> void put(int field, V v, int index = 0) { std::memcpy(buf.data() + lay[1 + field] + (size_t)index * sizeof(V), &v, sizeof(V)); }
};
enum { L_OUT, L_COPY, L_IN, L_INSZ, L_INSY, L_OUTSZ, L_OUTSY, L_NZ, L_NY, L_NX, L_BLO, L_BHI, L_WLO, L_WHI, L_ZCHUNK, L_NCHUNKS, L_NCX, L_NCY,
       L_CXO, L_CYO, L_NITEMS, L_ZGAP, L_C };

int jit_launch(JitBody &b, size_t ki, const SweepArgs &a) {
  int device = 0;
  cudaGetDevice(&device);
  std::string err;
  if (!b.dev.count(device))
    if (load_on_device(b, device, err) != NEPTUNE_RT_OK) return (int)cudaErrorInvalidDeviceFunction;
  JitBody::PerDevice &pd = b.dev[device];
  DriverApi &dr = driver();
  const KernelSpec &k = b.menu[ki];
  const int threads = pd.geom[ki][0], smem = pd.geom[ki][1], TX = pd.geom[ki][2], TY = pd.geom[ki][3];
  ParamPack p(pd.layout);
  p.put<void *>(L_OUT, a.out);
  p.put<const void *>(L_COPY, a.copy_src);
  for (int i = 0; i < b.nin; ++i) {
    p.put<const void *>(L_IN, a.in[i], i);
    p.put<long long>(L_INSZ, a.in_sz[i], i);
    p.put<long long>(L_INSY, a.in_sy[i], i);
  }
  p.put<long long>(L_OUTSZ, a.out_sz);
  p.put<long long>(L_OUTSY, a.out_sy);
  p.put<int>(L_NZ, a.nz); p.put<int>(L_NY, a.ny); p.put<int>(L_NX, a.nx);
  for (int x = 0; x < 3; ++x) {
    p.put<int>(L_BLO, a.blo[x], x); p.put<int>(L_BHI, a.bhi[x], x);
    p.put<int>(L_WLO, a.wlo[x], x); p.put<int>(L_WHI, a.whi[x], x);
  }
  for (int i = 0; i < std::max(b.nconst, 1); ++i) {
    const double v = i < b.nconst ? a.consts[i] : 0.0;
    if (b.elsize == 8) p.put<double>(L_C, v, i);
    else p.put<float>(L_C, (float)v, i);
  }
  const int wx = a.whi[2] - a.wlo[2], wy = a.whi[1] - a.wlo[1], wz = a.whi[0] - a.wlo[0];
  if (wx <= 0 || wy <= 0 || wz <= 0) return 0;
  unsigned gx = 1, gy = 1, gz = 1, bx = (unsigned)threads, by = 1;
  void *params[3] = {p.buf.data(), nullptr, nullptr};
  CUtensorMap tm0, tm1;
  if (k.launch == JL_DIRECT) {
    bx = 128; by = 2;   // DIRECT_BX x DIRECT_BY (kernels/sweep_direct.cuh)
    gx = (unsigned)((wx + 127) / 128); gy = (unsigned)((wy + 1) / 2); gz = (unsigned)wz;
    if (gy > 65535u || gz > 65535u) return (int)cudaErrorInvalidConfiguration;
  } else {
    if (pd.occ[ki] == 0) {
      if (dr.FuncSetAttribute(pd.fn[ki], CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, smem) != CUDA_SUCCESS) return (int)cudaErrorInvalidValue;
      int occ = 0;
      if (dr.OccupancyMaxActiveBlocksPerMultiprocessor(&occ, pd.fn[ki], threads, (size_t)smem) != CUDA_SUCCESS) return (int)cudaErrorInvalidValue;
      if (occ < 1) return (int)cudaErrorLaunchOutOfResources;
      pd.occ[ki] = occ;
    }
    const int slots = a.sm_count * pd.occ[ki];
    const int march = k.launch == JL_ROWS ? 1 : 0;   // canonical dimension the items march along
    const int planes = march == 0 ? wz : wy;
    int cxo = a.wlo[2] / TX, ncx = (a.whi[2] + TX - 1) / TX - cxo, cyo = 0, ncy = 1;
    if (k.launch == JL_TILE) { cyo = a.wlo[1] / TY; ncy = (a.whi[1] + TY - 1) / TY - cyo; }
    const int ncols = ncx * ncy;
    int zchunk = a.zchunk > 0 ? std::min(a.zchunk, planes) : choose_zchunk(planes, ncols, slots, k.zhalo);
    int nchunks = (planes + zchunk - 1) / zchunk, zgap = 0;
    if (a.gap_hi > a.gap_lo) {   // two equally wide boundary ranges as chunks 0 and 1 of one launch
      const int w = a.gap_lo - a.wlo[march];
      if (k.gap_dim != march || w <= 0 || a.whi[march] - a.gap_hi != w) return (int)cudaErrorInvalidValue;
      zchunk = w; nchunks = 2; zgap = a.gap_hi - a.gap_lo;
    }
    const long long items = (long long)nchunks * ncols;
    if (items > 0x7fffffffLL) return (int)cudaErrorInvalidConfiguration;
    p.put<int>(L_ZCHUNK, zchunk); p.put<int>(L_NCHUNKS, nchunks); p.put<int>(L_NCX, ncx); p.put<int>(L_NCY, ncy);
    p.put<int>(L_CXO, cxo); p.put<int>(L_CYO, cyo); p.put<int>(L_NITEMS, (int)items); p.put<int>(L_ZGAP, zgap);
    gx = (unsigned)std::min<long long>(items, slots);
    if (k.needs_tmap) {
      if (!a.tmap) return (int)cudaErrorInvalidValue;
      tm0 = *a.tmap;
      params[1] = &tm0;
      if (k.two_maps) {   // sweep_tb2m takes the pointwise input's map too (the staged one again when there is none)
        tm1 = a.tmap_pw ? *a.tmap_pw : *a.tmap;
        if (b.nin > 1 && !a.tmap_pw) return (int)cudaErrorInvalidValue;
        params[2] = &tm1;
      }
    }
  }
  params[0] = p.buf.data();
  const CUresult r = dr.LaunchKernel(pd.fn[ki], gx, gy, gz, bx, by, 1, k.launch == JL_DIRECT ? 0u : (unsigned)smem, (CUstream)a.stream, params, nullptr);
  return r == CUDA_SUCCESS ? 0 : (int)cudaErrorLaunchFailure;
}

std::vector<std::unique_ptr<JitBody>> &jit_bodies() {
  static std::vector<std::unique_ptr<JitBody>> v;
  return v;
}

}  // namespace

int jit_body(const neptune_rt_stencil_desc &d, int device, bool compile_only, std::string &err, double *seconds) {
  static std::mutex table_mutex;   // the JIT body table and its per-device modules: one compile / load at a time
  std::lock_guard<std::mutex> table_lock(table_mutex);
  const auto t_begin = std::chrono::steady_clock::now();
  const uint64_t key = body_key(d);
  const BodyInfo bi = analyze_body(d);
  const std::string F = functor_name(key);
  bool zstar = true;
  for (int i = 0; i < d.n_ops; ++i)
    if (d.ops[i].opcode == NEPTUNE_RT_OP_ACCESS) {
      int zyx[3];
      canonical_offset(d.rank, d.ops[i].offset, zyx);
      if (zyx[0] != 0 && (zyx[1] != 0 || zyx[2] != 0)) zstar = false;
    }
  JitBody *jb = nullptr;
  for (auto &b : jit_bodies()) if (b->key == key) jb = b.get();
  if (!jb) {
    std::unique_ptr<JitBody> nb(new JitBody());
    nb->key = key;
    nb->nin = d.n_inputs;
    nb->nconst = bi.n_consts;
    nb->elsize = d.dtype == NEPTUNE_RT_F64 ? 8 : 4;
    nb->menu = kernel_menu(d, bi, F, zstar);
    if (int rc = compile_unit(d, nb->menu, F, key, nb->code, err)) return rc;
    jit_bodies().push_back(std::move(nb));
    jb = jit_bodies().back().get();
  }
  if (compile_only) {
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
    return NEPTUNE_RT_OK;
  }
  if (int rc = load_on_device(*jb, device, err)) return rc;
  if (!find_body(key)) {
    BodyEntry e;
    e.key = key; e.rank = d.rank; e.dtype = d.dtype; e.nin = d.n_inputs; e.nconst = bi.n_consts; e.staged = bi.staged;
    if (bi.staged >= 0) { e.rz = bi.radius[bi.staged][0]; e.ry = bi.radius[bi.staged][1]; e.rx = bi.radius[bi.staged][2]; }
    const JitBody::PerDevice &pd = jb->dev[device];
    for (size_t i = 0; i < jb->menu.size(); ++i) {
      const KernelSpec &k = jb->menu[i];
      Variant v;
      v.name = k.name; v.kind = k.kind; v.fma = k.fma; v.steps = k.steps; v.gap_dim = k.gap_dim;
      v.needs_tmap = k.needs_tmap;
      v.threads = pd.geom[i][0]; v.smem = pd.geom[i][1]; v.tx = pd.geom[i][2]; v.ty = pd.geom[i][3];
      v.box_x = pd.geom[i][4]; v.box_y = pd.geom[i][5]; v.halo_x = pd.geom[i][6]; v.halo_y = pd.geom[i][7];
      v.launch = [jb, i](const SweepArgs &a) { return jit_launch(*jb, i, a); };
      e.variants.push_back(std::move(v));
    }
    register_body(std::move(e));
  }
  const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
  if (seconds) *seconds = sec;
  std::fprintf(stderr, "[NeptuneRT-B200] apply body 0x%016llx is not in the prebuilt catalogue: %zu sm_100a kernels %s in %.2f s\n",
               (unsigned long long)key, jb->menu.size(), jb->code.from_cache ? "loaded from the JIT cache" : "compiled with NVRTC", sec);
  return NEPTUNE_RT_OK;
}

}  // namespace neptune
