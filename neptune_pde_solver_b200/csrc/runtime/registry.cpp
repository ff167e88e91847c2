// Table of compiled sweep kernels (see kernels/registry.h). The generated units register
// themselves from static constructors, so the table lives behind a function-local static.
#include "kernels/registry.h"

#include <algorithm>

namespace neptune {

static std::deque<BodyEntry> &table() {
  static std::deque<BodyEntry> t;
  return t;
}

void register_body(BodyEntry &&e) {
  for (auto &b : table())
    if (b.key == e.key) return;  // same structure reached from two catalogue files
  table().push_back(std::move(e));
}

const BodyEntry *find_body(uint64_t key) {
  for (auto &b : table())
    if (b.key == key) return &b;
  return nullptr;
}

const std::deque<BodyEntry> &all_bodies() { return table(); }

}  // namespace neptune
