/* Model importer: turns the generated C model files of a reference model tarball (lpcnet_data-*.tar.gz: nnet_data.c,
 * plc_data.c, optionally dred_rdovae_dec_data.c) into the weights_blob.bin that lpcnet_load_model /
 * lpcnet_plc_load_model / lpcnet_batch_create of liblpcnet_b200 (and the reference itself, -DUSE_WEIGHTS_FILE) read.
 *
 * Does the job of the reference's src/write_lpcnet_weights.c:47-78 (record = 64-byte head {"DNNw", version 0, type,
 * size, block_size, name[44]} + payload padded to 64 bytes; include/nnet.h:41-61), restricted to the tables this
 * library consumes, and -- unlike it -- without needing the DRED encoder's data file to be present.
 *
 * Build (the generated files include the reference's nnet.h, so its src/ is on the include path):
 *   gcc -O1 -DDOT_PROD -I<dir with nnet_data.[ch], plc_data.[ch]> -I$REF/src -I$REF/include \
 *       [-DWITH_PLC] [-DWITH_DRED_DEC] tools/write_weights_shim.c -o write_weights_shim
 *   ./write_weights_shim weights_blob.bin
 * DOT_PROD selects the int8 variant of the quantised arrays, the one the AVX2 / NEON builds and this library use
 * (src/vec_avx.h:36, dump_lpcnet.py:114-118).  USE_WEIGHTS_FILE must be off so that the arrays exist. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#ifdef USE_WEIGHTS_FILE
#undef USE_WEIGHTS_FILE
#endif
#define DUMP_BINARY_WEIGHTS 1 /* the generated init functions are not needed here (dump_lpcnet.py:406) */
#include "nnet.h"
#include "nnet_data.c"
#ifdef WITH_PLC
#include "plc_data.c"
#endif
#ifdef WITH_DRED_DEC
#include "dred_rdovae_dec_data.c"
#endif

static int put_table(const WeightArray *t, FILE *out) {
  static const unsigned char pad[WEIGHT_BLOCK_SIZE] = {0};
  int n = 0;
  for (; t->name != NULL; t++, n++) {
    WeightHead head;
    size_t len = strlen(t->name);
    if (sizeof(head) != WEIGHT_BLOCK_SIZE || len >= sizeof(head.name) || t->size <= 0) {
      fprintf(stderr, "write_weights_shim: cannot store array %s (%d bytes)\n", t->name, t->size);
      return -1;
    }
    memset(&head, 0, sizeof(head));
    memcpy(head.head, "DNNw", 4);
    head.version = WEIGHT_BLOB_VERSION;
    head.type = t->type;
    head.size = t->size;
    head.block_size = (t->size + WEIGHT_BLOCK_SIZE - 1) / WEIGHT_BLOCK_SIZE * WEIGHT_BLOCK_SIZE;
    memcpy(head.name, t->name, len);
    if (fwrite(&head, 1, sizeof(head), out) != sizeof(head)) return -1;
    if (fwrite(t->data, 1, (size_t)t->size, out) != (size_t)t->size) return -1;
    if (fwrite(pad, 1, (size_t)(head.block_size - t->size), out) != (size_t)(head.block_size - t->size)) return -1;
  }
  return n;
}

int main(int argc, char **argv) {
  const char *path = argc > 1 ? argv[1] : "weights_blob.bin";
  FILE *out = fopen(path, "wb");
  int n, total = 0;
  if (!out) { perror(path); return 1; }
  if ((n = put_table(lpcnet_arrays, out)) < 0) return 1;
  total += n;
#ifdef WITH_PLC
  if ((n = put_table(lpcnet_plc_arrays, out)) < 0) return 1;
  total += n;
#endif
#ifdef WITH_DRED_DEC
  if ((n = put_table(rdovae_dec_arrays, out)) < 0) return 1;
  total += n;
#endif
  if (fclose(out) != 0) { perror(path); return 1; }
  fprintf(stderr, "write_weights_shim: %d arrays -> %s\n", total, path);
  return 0;
}
