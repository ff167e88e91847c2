// neptune-b200-gen: build-time half of the `neptuneir-to-b200` pipeline.
// Reads NeptuneIR .mlir files (the kernel catalogue), recognises every neptune_ir.apply and
// writes one CUDA translation unit per distinct body structure:  <outdir>/body_<key>.cu
// (functor + NEPTUNE_REGISTER_BODY).  The build compiles those for sm_100a into the runtime.
//   usage: neptune-b200-gen <outdir> <file.mlir>...
#include <cstdio>
#include <fstream>
#include <iostream>
#include <map>
#include <sstream>

#include "frontend/ir.h"

int main(int argc, char **argv) {
  if (argc < 3) {
    std::fprintf(stderr, "usage: %s <outdir> <file.mlir>...\n", argv[0]);
    return 2;
  }
  const std::string outdir = argv[1];
  std::map<uint64_t, std::string> seen;  // key -> first file
  int rc = 0;
  for (int i = 2; i < argc; ++i) {
    std::ifstream in(argv[i]);
    if (!in) { std::fprintf(stderr, "cannot read %s\n", argv[i]); return 2; }
    std::stringstream ss;
    ss << in.rdbuf();
    try {
      neptune::Module m0 = neptune::parse_module(ss.str());
      // the bodies as written, then the bodies producer -> consumer fusion makes of them (fusion.cpp)
      neptune::Module m1 = m0;
      const int nf = neptune::fuse_applies(m1, false, nullptr);
      for (int pass = 0; pass < (nf ? 2 : 1); ++pass) {
      neptune::Module &m = pass ? m1 : m0;
      for (auto &fn : m.functions)
        for (auto &op : fn.ops) {
          if (op.kind != neptune::HostOp::ApplyOp) continue;
          neptune::StencilDesc d = neptune::recognize_apply(*op.apply, false);
          const uint64_t key = neptune::body_key(d.c);
          if (seen.count(key)) continue;
          seen[key] = argv[i];
          const std::string path = outdir + "/body_" + neptune::functor_name(key).substr(3) + ".cu";
          std::string text = "// from " + std::string(argv[i]) + " (" + fn.name + ")\n" +
                             neptune::generate_registration_unit({&d.c});
          // keep timestamps stable when nothing changed (incremental builds)
          std::ifstream old(path);
          std::stringstream os;
          if (old) os << old.rdbuf();
          if (!old || os.str() != text) {
            std::ofstream out(path);
            out << text;
          }
          std::printf("%s\n", path.c_str());
        }
      }
    } catch (const neptune::Diagnostic &e) {
      std::fprintf(stderr, "%s: %s\n", argv[i], e.what());
      rc = 1;
    }
  }
  return rc;
}
