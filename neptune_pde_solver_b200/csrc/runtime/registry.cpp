// Table of compiled sweep kernels (see kernels/registry.h). The generated units register
// themselves from static constructors, so the table lives behind a function-local static.
#include "kernels/registry.h"

#include <algorithm>
#include <mutex>

namespace neptune {

static std::deque<BodyEntry> &table() {
  static std::deque<BodyEntry> t;
  return t;
}

// registration can come from the JIT at any time (runtime/jit.cpp) while other threads look bodies up; entries are never
// removed and a deque does not move them, so the pointers find_body hands out stay valid
static std::mutex &table_mutex() {
  static std::mutex m;
  return m;
}

void register_body(BodyEntry &&e) {
  std::lock_guard<std::mutex> lock(table_mutex());
  for (auto &b : table())
    if (b.key == e.key) return;  // same structure reached from two catalogue files
  table().push_back(std::move(e));
}

const BodyEntry *find_body(uint64_t key) {
  std::lock_guard<std::mutex> lock(table_mutex());
  for (auto &b : table())
    if (b.key == key) return &b;
  return nullptr;
}

const std::deque<BodyEntry> &all_bodies() { return table(); }

}  // namespace neptune
