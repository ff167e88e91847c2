/* A host driver shaped like the one the reference's smoke test generates for its compiled @entry
 * (test/smoke_tests/smoke_apply.sh:39-80: build memref descriptors by hand, call the entry, print the result) —
 * here calling the B200 runtime's memref entry point the way code emitted by the `neptune-ir-apply-to-b200` pass
 * would: the stencil descriptor is data (desc.h, written by `neptune-opt-b200 --emit=c-desc`), the operands are MLIR
 * memref descriptors built on the stack.  Plain C, links against libneptune_rt_b200.so only.
 *
 *   usage: memref_driver <out.bin> <n_inputs> <in0.bin> [<in1.bin> ...]     (raw row-major doubles / floats)
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "neptune_rt.h"
#include "desc.h" /* neptune_apply0 */

#define PAD 3 /* elements before the data inside every buffer: exercises the descriptor's `offset` */

static void *read_file(const char *path, size_t bytes, size_t esz) {
  char *buf = (char *)malloc(bytes + PAD * esz);
  FILE *f = fopen(path, "rb");
  if (!f || fread(buf + PAD * esz, 1, bytes, f) != bytes) { fprintf(stderr, "cannot read %s\n", path); exit(2); }
  fclose(f);
  return buf;
}

int main(int argc, char **argv) {
  const neptune_rt_stencil_desc *d = &neptune_apply0;
  const size_t esz = d->dtype == NEPTUNE_RT_F64 ? 8 : 4;
  if (argc < 3 || atoi(argv[2]) != d->n_inputs || argc != 3 + d->n_inputs) { fprintf(stderr, "usage: %s out n_in in0 ...\n", argv[0]); return 2; }
  int64_t osz[3] = {1, 1, 1}, ostr[3];
  size_t onumel = 1;
  for (int k = 0; k < d->rank; ++k) { osz[k] = d->out_ub[k] - d->out_lb[k]; onumel *= (size_t)osz[k]; }
  void *out = malloc((onumel + PAD) * esz);
  memset(out, 0xff, (onumel + PAD) * esz);
  void *in[NEPTUNE_RT_MAX_INPUTS];
#define FILL(M, buf, lb, ub)                                                       \
  do {                                                                             \
    int64_t st_ = 1;                                                               \
    (M).allocated = (buf); (M).aligned = (buf); (M).offset = PAD;                  \
    for (int k_ = d->rank - 1; k_ >= 0; --k_) {                                    \
      (M).sizes[k_] = (ub)[k_] - (lb)[k_]; (M).strides[k_] = st_; st_ *= (M).sizes[k_]; \
    }                                                                              \
  } while (0)
  for (int i = 0; i < d->n_inputs; ++i) {
    size_t n = 1;
    for (int k = 0; k < d->rank; ++k) n *= (size_t)(d->in_ub[i * d->rank + k] - d->in_lb[i * d->rank + k]);
    in[i] = read_file(argv[3 + i], n * esz, esz);
  }
  (void)ostr;
  if (d->rank == 1) {
    neptune_rt_memref_1d mi[NEPTUNE_RT_MAX_INPUTS], mo;
    for (int i = 0; i < d->n_inputs; ++i) FILL(mi[i], in[i], d->in_lb + i * d->rank, d->in_ub + i * d->rank);
    FILL(mo, out, d->out_lb, d->out_ub);
    neptune_rt_apply_memref_1d(d, d->n_inputs, mi, &mo);
    neptune_rt_apply_memref_1d(d, d->n_inputs, mi, &mo); /* a second call reuses the cached stencil and device fields */
  } else if (d->rank == 2) {
    neptune_rt_memref_2d mi[NEPTUNE_RT_MAX_INPUTS], mo;
    for (int i = 0; i < d->n_inputs; ++i) FILL(mi[i], in[i], d->in_lb + i * d->rank, d->in_ub + i * d->rank);
    FILL(mo, out, d->out_lb, d->out_ub);
    neptune_rt_apply_memref_2d(d, d->n_inputs, mi, &mo);
    neptune_rt_apply_memref_2d(d, d->n_inputs, mi, &mo);
  } else {
    neptune_rt_memref_3d mi[NEPTUNE_RT_MAX_INPUTS], mo;
    for (int i = 0; i < d->n_inputs; ++i) FILL(mi[i], in[i], d->in_lb + i * d->rank, d->in_ub + i * d->rank);
    FILL(mo, out, d->out_lb, d->out_ub);
    if (getenv("NEPTUNE_DRIVER_BAD_STRIDE")) mi[0].strides[d->rank - 1] = 2; /* must abort with a message, not compute garbage */
    neptune_rt_apply_memref_3d(d, d->n_inputs, mi, &mo);
    neptune_rt_apply_memref_3d(d, d->n_inputs, mi, &mo);
  }
  FILE *f = fopen(argv[1], "wb");
  if (!f || fwrite((char *)out + PAD * esz, 1, onumel * esz, f) != onumel * esz) { fprintf(stderr, "cannot write %s\n", argv[1]); return 2; }
  fclose(f);
  neptune_rt_memref_cache_clear();
  printf("memref_driver ok: rank %d, %d input(s), %zu result elements\n", d->rank, d->n_inputs, onumel);
  return 0;
}
