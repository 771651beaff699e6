/* TEST DOUBLE of the five libpgen_cuda.so entry points the JNI shim calls -- for tests/test_jni_shim.py only, so that
 * the shim's marshalling (byte[] in, byte[][][] out, error -> exception) can run on a machine without a GPU.  It decodes
 * nothing: poll returns every submitted row with the summary "summary of <uuid>".  Never linked into the product. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "pgen.h"

struct pgen { pgen_row_t rows[16]; size_t n; char err[64]; };

static char* dup_n(const char* p, size_t n) { char* q = (char*)malloc(n + 1); memcpy(q, p, n); q[n] = 0; return q; }

int pgen_create(const char* hyperparams, int device_ordinal, pgen_t** out) {
  *out = (pgen_t*)calloc(1, sizeof(struct pgen));
  return PGEN_OK;
}
int pgen_submit(pgen_t* h, const char* uuid, size_t nu, const char* art, size_t na, const char* ref, size_t nr) {
  if (h->n == 16) { strcpy(h->err, "fake: full"); return PGEN_ERR_STATE; }
  pgen_row_t* r = &h->rows[h->n++];
  char buf[128];
  r->uuid = dup_n(uuid, nu); r->uuid_len = nu;
  r->article = dup_n(art, na); r->article_len = na;
  r->reference = dup_n(ref, nr); r->reference_len = nr;
  snprintf(buf, sizeof buf, "summary of %s", r->uuid);
  r->summary = dup_n(buf, strlen(buf)); r->summary_len = strlen(buf);
  return PGEN_OK;
}
int pgen_poll(pgen_t* h, int flush, pgen_rows_t* out) {
  out->rows = 0; out->n_rows = 0; out->owner_ = 0;
  if (!flush) return PGEN_OK;
  out->rows = h->rows; out->n_rows = h->n;
  return PGEN_OK;
}
void pgen_rows_free(pgen_rows_t* rows) { rows->rows = 0; rows->n_rows = 0; }
const char* pgen_last_error(pgen_t* h) { return h ? h->err : "fake: create failed"; }
void pgen_destroy(pgen_t* h) { free(h); }
int pgen_device_count(void) { return 7; }
