// extern "C" boundary of libpgen_cuda.so (see include/pgen.h for the reference interfaces replaced).
#include <chrono>
#include <cstring>
#include <deque>
#include <memory>
#include <string>
#include <vector>

#include "../../include/pgen.h"
#include "engine.h"
#include "text.h"
#include "tfexample.h"
#include "handle_access.h"

using namespace pgn;

namespace {
thread_local std::string g_create_error;

struct PendingRow {
  std::string uuid, article, reference;
  EncodedArticle enc;
  std::chrono::steady_clock::time_point t_in;
};
struct DoneRow { std::string uuid, article, summary, reference; };
struct RowsOwner { std::vector<DoneRow> rows; std::vector<pgen_row_t> views; };

double ms_since(std::chrono::steady_clock::time_point t0) {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}
}  // namespace

struct InFlight { int slot; std::vector<PendingRow> rows; };

struct pgen {
  HyperParams hp;
  Vocab vocab;
  // One engine (weights replica + batch slot + stream) per slot: batch i+1 is tokenised, uploaded and
  // launched while batch i is still decoding, and their kernels overlap on the GPU.
  std::vector<std::unique_ptr<Engine>> engs;
  std::deque<InFlight> inflight;     // launched batches, oldest first (= submission order)
  int next_slot = 0, stage_slot = 0;
  volatile int where = 0;                  // host-side breadcrumb, read without locks by pgen_debug_where
  Engine& eng() { return *engs[stage_slot]; }
  std::string err;
  std::deque<PendingRow> pending;
  std::deque<DoneRow> done;
  EncodedArticle last_text;          // pgen_text_encode -> pgen_text_summary
  uint64_t rows_in = 0, rows_out = 0, tokens_in = 0, tokens_out = 0;
  double ms_host_text = 0;
  std::vector<int32_t> b_ids, b_ext, b_lens, b_noov, b_tok, b_len;
  std::vector<float> b_lp;
  std::vector<double> b_avg;
};

static int fail(pgen_t* h, int code, const std::string& msg) {
  if (h) h->err = msg; else g_create_error = msg;
  return code;
}

// finish the oldest in-flight batch: wait for it, detokenise, move its rows to `done`.  The batch leaves the
// in-flight list only once its results are on the host, so a failed download can be retried and loses no row; a row
// whose token ids cannot be mapped back to words is emitted with an empty summary and named in the error, the rest
// of its batch is emitted normally.
static int collect_oldest(pgen_t* h) {
  h->where = 30;
  InFlight& f = h->inflight.front();
  Engine& e = *h->engs[f.slot];
  const size_t n = f.rows.size();
  const int S = (int)h->hp.max_dec_steps;
  h->b_tok.resize(n * S); h->b_len.resize(n); h->b_lp.resize(n * S); h->b_avg.resize(n);
  std::string err = e.download((int)n, h->b_tok.data(), h->b_len.data(), h->b_lp.data(), h->b_avg.data());
  if (!err.empty()) return fail(h, PGEN_ERR_CUDA, err);
  auto t0 = std::chrono::steady_clock::now();
  std::string bad;
  for (size_t b = 0; b < n; ++b) {
    PendingRow& p = f.rows[b];
    DoneRow d;
    if (!summary_from_ids(h->vocab, p.enc.oovs, &h->b_tok[b * S], h->b_len[b], &d.summary)) {
      d.summary.clear();
      bad += (bad.empty() ? "" : ", ") + p.uuid;
    }
    h->tokens_out += h->b_len[b];
    d.uuid = std::move(p.uuid); d.article = std::move(p.article); d.reference = std::move(p.reference);
    h->done.push_back(std::move(d));
  }
  h->inflight.pop_front();
  h->ms_host_text += ms_since(t0);
  if (!bad.empty())
    return fail(h, PGEN_ERR_STATE, "decoder produced a token id outside the extended vocabulary of row(s) " + bad +
                                   " (emitted with an empty summary)");
  return PGEN_OK;
}

// launch the first n pending rows as one device batch on the next slot (asynchronous)
static int launch_pending(pgen_t* h, size_t n) {
  h->where = 40;
  const int slot = h->next_slot;
  while (!h->inflight.empty()) {                   // the slot must be free; batches complete in FIFO order
    bool busy = false;
    for (auto& f : h->inflight) busy |= (f.slot == slot);
    if (!busy) break;
    int rc = collect_oldest(h);
    if (rc != PGEN_OK) return rc;
  }
  const int T = (int)h->hp.max_enc_steps;
  h->b_ids.assign(n * T, 0); h->b_ext.assign(n * T, 0); h->b_lens.assign(n, 0); h->b_noov.assign(n, 0);
  for (size_t b = 0; b < n; ++b) {
    const EncodedArticle& e = h->pending[b].enc;
    h->b_lens[b] = (int)e.enc_ids.size();
    h->b_noov[b] = (int)e.oovs.size();
    if (!e.enc_ids.empty()) {                                  // (an empty vector's data() may be NULL: no memcpy from it)
      memcpy(&h->b_ids[b * T], e.enc_ids.data(), e.enc_ids.size() * 4);
      memcpy(&h->b_ext[b * T], e.ext_ids.data(), e.ext_ids.size() * 4);
    }
  }
  Engine& e = *h->engs[slot];
  h->where = 41;
  std::string err = e.upload((int)n, h->b_ids.data(), h->b_ext.data(), T, h->b_lens.data(), h->b_noov.data());
  h->where = 42;
  if (err.empty()) err = e.run();
  h->where = 43;
  if (!err.empty()) return fail(h, PGEN_ERR_CUDA, err);        // the rows stay in `pending`: nothing is lost
  InFlight f;
  f.slot = slot;
  f.rows.reserve(n);
  for (size_t b = 0; b < n; ++b) {
    f.rows.push_back(std::move(h->pending.front()));
    h->pending.pop_front();
  }
  h->inflight.push_back(std::move(f));
  h->next_slot = (slot + 1) % (int)h->engs.size();
  return PGEN_OK;
}

// The stage / tensor-level entry points drive one slot's engine directly: refused while a batch submitted through
// pgen_submit is in flight on that slot (its buffers would be overwritten under it).
static bool slot_busy(pgen_t* h) {
  for (auto& f : h->inflight) if (f.slot == h->stage_slot) return true;
  return false;
}
static int busy_error(pgen_t* h, const char* what) {
  return fail(h, PGEN_ERR_STATE, std::string(what) + ": slot " + std::to_string(h->stage_slot) +
                                 " holds a batch of pgen_submit rows in flight; pgen_poll(flush) first or choose another slot");
}

pgn::Engine* pgen_internal_engine(pgen_t* h) { return &h->eng(); }
bool pgen_internal_coverage(pgen_t* h) { return h->hp.coverage; }
int pgen_internal_fail(pgen_t* h, int code, const std::string& msg) { return fail(h, code, msg); }

extern "C" {

int pgen_create(const char* hyperparams, int device_ordinal, pgen_t** out) {
  if (!out) return fail(nullptr, PGEN_ERR_ARG, "out is NULL");
  *out = nullptr;
  if (!hyperparams) return fail(nullptr, PGEN_ERR_ARG, "hyperparams is NULL");
  std::unique_ptr<pgen> h(new pgen());
  std::string err = parse_hyperparams(hyperparams, &h->hp);
  if (!err.empty()) return fail(nullptr, PGEN_ERR_ARG, err);
  const HyperParams& hp = h->hp;
  if (hp.mode != "decode") {
    if (hp.mode == "train" || hp.mode == "eval")
      return fail(nullptr, PGEN_ERR_UNSUPPORTED, "mode=" + hp.mode + ": libpgen_cuda serves mode=decode only");
    return fail(nullptr, PGEN_ERR_ARG, "The 'mode' flag must be one of train/eval/decode");   // run_summarization.py:426
  }
  err = h->vocab.load(hp.vocab_path, hp.vocab_size);                                          // :307
  if (!err.empty()) return fail(nullptr, PGEN_ERR_IO, err);
  if (hp.slots < 1 || hp.slots > 8) return fail(nullptr, PGEN_ERR_ARG, "slots must be in 1..8");
  for (int64_t i = 0; i < hp.slots; ++i) {
    int code = PGEN_OK;
    h->engs.emplace_back(new Engine());
    err = h->engs.back()->init(hp, h->vocab.size(), device_ordinal, &code, i == 0 ? nullptr : h->engs[0].get());
    if (!err.empty()) return fail(nullptr, code, err);
  }
  *out = h.release();
  return PGEN_OK;
}

void pgen_destroy(pgen_t* h) { delete h; }

int pgen_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

const char* pgen_last_error(pgen_t* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int64_t pgen_flag_int(pgen_t* h, const char* name) { return (h && name) ? flag_int(h->hp, name) : -1; }

int pgen_submit(pgen_t* h, const char* uuid, size_t uuid_len, const char* article, size_t article_len,
                const char* reference, size_t reference_len) {
  if (!h) return PGEN_ERR_ARG;
  if ((!uuid && uuid_len) || (!article && article_len) || (!reference && reference_len))
    return fail(h, PGEN_ERR_ARG, "NULL buffer with non-zero length");
  auto t0 = std::chrono::steady_clock::now();
  PendingRow p;
  p.uuid.assign(uuid ? uuid : "", uuid_len);
  p.article.assign(article ? article : "", article_len);
  p.reference.assign(reference ? reference : "", reference_len);
  encode_article(h->vocab, p.article, (int)h->hp.max_enc_steps, &p.enc);
  p.t_in = t0;
  h->tokens_in += p.enc.enc_ids.size();
  ++h->rows_in;
  h->pending.push_back(std::move(p));
  h->ms_host_text += ms_since(t0);
  if ((int64_t)h->pending.size() >= h->hp.articles_per_batch) return launch_pending(h, (size_t)h->hp.articles_per_batch);
  return PGEN_OK;
}

int pgen_poll(pgen_t* h, int flush, pgen_rows_t* out) {
  if (!h || !out) return PGEN_ERR_ARG;
  out->rows = nullptr; out->n_rows = 0; out->owner_ = nullptr;
  if (!h->pending.empty()) {
    const bool late = h->hp.flush_ms >= 0 && ms_since(h->pending.front().t_in) >= (double)h->hp.flush_ms;
    if (flush || late) {
      while (!h->pending.empty()) {
        const size_t n = std::min<size_t>(h->pending.size(), (size_t)h->hp.articles_per_batch);
        int rc = launch_pending(h, n);
        if (rc != PGEN_OK) return rc;
      }
    }
  }
  // finished batches, oldest first; with flush wait for everything in flight
  while (!h->inflight.empty() && (flush || h->engs[h->inflight.front().slot]->done())) {
    int rc = collect_oldest(h);
    if (rc != PGEN_OK) return rc;
  }
  if (h->done.empty()) return PGEN_OK;
  auto* own = new RowsOwner();
  own->rows.assign(std::make_move_iterator(h->done.begin()), std::make_move_iterator(h->done.end()));
  h->done.clear();
  own->views.resize(own->rows.size());
  for (size_t i = 0; i < own->rows.size(); ++i) {
    const DoneRow& d = own->rows[i];
    own->views[i] = {d.uuid.c_str(), d.uuid.size(), d.article.c_str(), d.article.size(),
                     d.summary.c_str(), d.summary.size(), d.reference.c_str(), d.reference.size()};
  }
  h->rows_out += own->rows.size();
  out->rows = own->views.data(); out->n_rows = own->views.size(); out->owner_ = own;
  return PGEN_OK;
}

void pgen_rows_free(pgen_rows_t* rows) {
  if (!rows) return;
  delete static_cast<RowsOwner*>(rows->owner_);
  rows->rows = nullptr; rows->n_rows = 0; rows->owner_ = nullptr;
}

int pgen_stats(pgen_t* h, pgen_stats_t* out) {
  if (!h || !out) return PGEN_ERR_ARG;
  memset(out, 0, sizeof *out);
  EngineStats s;
  for (auto& e : h->engs) {
    const EngineStats& x = e->stats;
    s.batches += x.batches; s.decode_steps += x.decode_steps; s.article_steps += x.article_steps;
    s.early_stops += x.early_stops; s.kernel_launches += x.kernel_launches;
    s.selection_widened += x.selection_widened; s.selection_fallbacks += x.selection_fallbacks;
  }
  const EngineStats& last = h->eng().stats;          // timings: the slot the stage API addresses
  out->rows_in = h->rows_in; out->rows_out = h->rows_out; out->batches = s.batches;
  out->tokens_in = h->tokens_in; out->tokens_out = h->tokens_out; out->decode_steps = s.decode_steps;
  out->article_steps = s.article_steps; out->early_stops = s.early_stops; out->kernel_launches = s.kernel_launches;
  out->ms_host_text = h->ms_host_text; out->ms_h2d = last.ms_h2d; out->ms_encoder = last.ms_encoder;
  out->ms_decoder = last.ms_decoder; out->ms_d2h = last.ms_d2h;
  out->ms_attention = last.ms_attention; out->ms_vocab = last.ms_vocab; out->ms_select = last.ms_beam;
  out->prof_steps = (double)last.prof_steps;
  out->selection_widened = s.selection_widened; out->selection_fallbacks = s.selection_fallbacks;
  out->ms_enc_recurrence = last.ms_enc_recurrence; out->ms_enc_features = last.ms_enc_features;
  return PGEN_OK;
}

int pgen_text_encode(pgen_t* h, const char* article, size_t article_len, int32_t* enc_ids, int32_t* ext_ids,
                     int cap, int32_t* enc_len, int32_t* n_oov) {
  if (!h || (!article && article_len)) return PGEN_ERR_ARG;
  encode_article(h->vocab, std::string_view(article ? article : "", article_len), (int)h->hp.max_enc_steps, &h->last_text);
  const int n = (int)h->last_text.enc_ids.size();
  if (enc_len) *enc_len = n;
  if (n_oov) *n_oov = (int)h->last_text.oovs.size();
  if (n > cap) return fail(h, PGEN_ERR_ARG, "pgen_text_encode: cap smaller than the encoded length");
  if (enc_ids && n) memcpy(enc_ids, h->last_text.enc_ids.data(), (size_t)n * 4);
  if (ext_ids && n) memcpy(ext_ids, h->last_text.ext_ids.data(), (size_t)n * 4);
  return PGEN_OK;
}

int pgen_text_summary(pgen_t* h, const int32_t* tokens, int n_tokens, char* out, size_t cap, size_t* len) {
  if (!h || (!tokens && n_tokens) || n_tokens < 0) return PGEN_ERR_ARG;
  std::string s;
  if (!summary_from_ids(h->vocab, h->last_text.oovs, tokens, n_tokens, &s))
    return fail(h, PGEN_ERR_ARG, "token id outside the extended vocabulary of the last encoded article");
  if (len) *len = s.size();
  if (out && cap) memcpy(out, s.data(), std::min(cap, s.size()));
  return PGEN_OK;
}

int pgen_parse_flag(const char* hyperparams, const char* name, int64_t* value) {
  if (!hyperparams || !name || !value) return fail(nullptr, PGEN_ERR_ARG, "NULL argument");
  HyperParams hp;
  std::string err = parse_hyperparams(hyperparams, &hp);
  if (!err.empty()) return fail(nullptr, PGEN_ERR_ARG, err);
  const int64_t v = flag_int(hp, name);
  if (v == -1 && std::string(name) != "flush_ms") return fail(nullptr, PGEN_ERR_ARG, std::string("not an integer or boolean flag: ") + name);
  *value = v;
  return PGEN_OK;
}

// ---- the host text path on its own handle: no device involved ----
struct pgen_text {
  Vocab vocab;
  int max_enc_steps = 400;
  EncodedArticle last;
};

int pgen_text_open(const char* vocab_path, int64_t vocab_size, int max_enc_steps, pgen_text_t** out) {
  if (!out) return fail(nullptr, PGEN_ERR_ARG, "out is NULL");
  *out = nullptr;
  if (!vocab_path) return fail(nullptr, PGEN_ERR_ARG, "vocab_path is NULL");
  if (vocab_size < 0 || max_enc_steps < 1) return fail(nullptr, PGEN_ERR_ARG, "vocab_size must be >= 0 and max_enc_steps positive");
  std::unique_ptr<pgen_text> t(new pgen_text());
  std::string err = t->vocab.load(vocab_path, vocab_size);                                  // run_summarization.py:307
  if (!err.empty()) return fail(nullptr, PGEN_ERR_IO, err);
  t->max_enc_steps = max_enc_steps;
  *out = t.release();
  return PGEN_OK;
}

int pgen_text_vocab_size(pgen_text_t* t) { return t ? t->vocab.size() : -1; }

int pgen_text_encode_article(pgen_text_t* t, const char* article, size_t article_len, int32_t* enc_ids, int32_t* ext_ids,
                             int cap, int32_t* enc_len, int32_t* n_oov) {
  if (!t || (!article && article_len)) return PGEN_ERR_ARG;
  encode_article(t->vocab, std::string_view(article ? article : "", article_len), t->max_enc_steps, &t->last);
  const int n = (int)t->last.enc_ids.size();
  if (enc_len) *enc_len = n;
  if (n_oov) *n_oov = (int)t->last.oovs.size();
  if (n > cap) return fail(nullptr, PGEN_ERR_ARG, "pgen_text_encode_article: cap smaller than the encoded length");
  if (enc_ids && n) memcpy(enc_ids, t->last.enc_ids.data(), (size_t)n * 4);
  if (ext_ids && n) memcpy(ext_ids, t->last.ext_ids.data(), (size_t)n * 4);
  return PGEN_OK;
}

int pgen_text_decode_summary(pgen_text_t* t, const int32_t* tokens, int n_tokens, char* out, size_t cap, size_t* len) {
  if (!t || (!tokens && n_tokens) || n_tokens < 0) return PGEN_ERR_ARG;
  std::string s;
  if (!summary_from_ids(t->vocab, t->last.oovs, tokens, n_tokens, &s))
    return fail(nullptr, PGEN_ERR_ARG, "token id outside the extended vocabulary of the last encoded article");
  if (len) *len = s.size();
  if (out && cap) memcpy(out, s.data(), std::min(cap, s.size()));
  return PGEN_OK;
}

void pgen_text_close(pgen_text_t* t) { delete t; }

int pgen_stage_upload(pgen_t* h, int n_articles, const int32_t* enc_ids, const int32_t* enc_ext_ids, int ld,
                      const int32_t* enc_lens, const int32_t* n_oovs) {
  if (!h || n_articles < 0 || (n_articles && (!enc_ids || !enc_ext_ids || !enc_lens))) return h ? fail(h, PGEN_ERR_ARG, "bad arguments") : PGEN_ERR_ARG;
  if (slot_busy(h)) return busy_error(h, "pgen_stage_upload");
  std::string err = h->eng().upload(n_articles, enc_ids, enc_ext_ids, ld, enc_lens, n_oovs);
  return err.empty() ? PGEN_OK : fail(h, PGEN_ERR_ARG, err);
}
int pgen_stage_run(pgen_t* h) {
  if (!h) return PGEN_ERR_ARG;
  if (slot_busy(h)) return busy_error(h, "pgen_stage_run");
  std::string err = h->eng().run();
  return err.empty() ? PGEN_OK : fail(h, err.find("before") != std::string::npos ? PGEN_ERR_STATE : PGEN_ERR_CUDA, err);
}
int pgen_stage_sync(pgen_t* h) {
  if (!h) return PGEN_ERR_ARG;
  std::string err = h->eng().sync();
  return err.empty() ? PGEN_OK : fail(h, PGEN_ERR_CUDA, err);
}
int pgen_stage_download(pgen_t* h, int32_t* out_tokens, int32_t* out_lens, float* out_logprobs, double* out_avg) {
  if (!h) return PGEN_ERR_ARG;
  if (slot_busy(h)) return busy_error(h, "pgen_stage_download");
  std::string err = h->eng().download(-1, out_tokens, out_lens, out_logprobs, out_avg);
  return err.empty() ? PGEN_OK : fail(h, err.find("before") != std::string::npos ? PGEN_ERR_STATE : PGEN_ERR_CUDA, err);
}
void* pgen_stream(pgen_t* h) { return h ? (void*)h->eng().stream() : nullptr; }

int pgen_stage_slot(pgen_t* h, int slot) {
  if (!h) return PGEN_ERR_ARG;
  if (slot < 0 || slot >= (int)h->engs.size()) return fail(h, PGEN_ERR_ARG, "slot out of range");
  h->stage_slot = slot;
  return PGEN_OK;
}

int pgen_decode_ids(pgen_t* h, int n_articles, const int32_t* enc_ids, const int32_t* enc_ext_ids, int ld,
                    const int32_t* enc_lens, const int32_t* n_oovs, int32_t* out_tokens, int32_t* out_lens,
                    float* out_logprobs, double* out_avg) {
  if (!h) return PGEN_ERR_ARG;
  if (n_articles < 0 || (n_articles && (!enc_ids || !enc_ext_ids || !enc_lens))) return fail(h, PGEN_ERR_ARG, "bad arguments");
  const int cap = h->eng().capacity(), S = (int)h->hp.max_dec_steps;
  for (int b0 = 0; b0 < n_articles; b0 += cap) {
    const int n = std::min(cap, n_articles - b0);
    int rc = pgen_stage_upload(h, n, enc_ids + (size_t)b0 * ld, enc_ext_ids + (size_t)b0 * ld, ld, enc_lens + b0,
                               n_oovs ? n_oovs + b0 : nullptr);
    if (rc == PGEN_OK) rc = pgen_stage_run(h);
    if (rc == PGEN_OK) rc = pgen_stage_sync(h);
    if (rc == PGEN_OK)
      rc = pgen_stage_download(h, out_tokens ? out_tokens + (size_t)b0 * S : nullptr, out_lens ? out_lens + b0 : nullptr,
                               out_logprobs ? out_logprobs + (size_t)b0 * S : nullptr, out_avg ? out_avg + b0 : nullptr);
    if (rc != PGEN_OK) return rc;
  }
  return PGEN_OK;
}

int pgen_attention(pgen_t* h, int n_articles, const float* enc_feat, const float* enc_states, const int32_t* lens,
                   const float* dfeat, const float* cov, float* attn, float* ctx) {
  if (!h) return PGEN_ERR_ARG;
  if (n_articles < 0 || (n_articles && (!enc_feat || !enc_states || !lens || !dfeat || !attn || !ctx)))
    return fail(h, PGEN_ERR_ARG, "pgen_attention: NULL device pointer");
  if (n_articles && h->hp.coverage && !cov) return fail(h, PGEN_ERR_ARG, "pgen_attention: coverage=1 needs cov");
  if (slot_busy(h)) return busy_error(h, "pgen_attention");
  std::string err = h->eng().attention_op(n_articles, enc_feat, enc_states, lens, dfeat, cov, attn, ctx);
  return err.empty() ? PGEN_OK : fail(h, PGEN_ERR_CUDA, err);
}

// ---- tf.Example wire format of the rows (stateless; no GPU involved) ----
static const pgn::BytesFeature* find_feature(const std::vector<pgn::BytesFeature>& fs, const char* name) {
  for (auto& f : fs) if (f.name == name) return &f;
  return nullptr;
}

int pgen_example_parse(const void* example, size_t len, pgen_row_t* out) {
  if ((!example && len) || !out) return PGEN_ERR_ARG;
  std::vector<pgn::BytesFeature> fs;
  std::string err = pgn::parse_example(static_cast<const uint8_t*>(example), len, &fs);
  if (!err.empty()) { g_create_error = err; return PGEN_ERR_ARG; }
  memset(out, 0, sizeof *out);
  const char* need[3] = {"uuid", "article", "reference"};                 // batcher.py:554-557, FixedLenFeature([], string)
  const pgn::BytesFeature* got[3];
  for (int i = 0; i < 3; ++i) {
    got[i] = find_feature(fs, need[i]);
    if (!got[i]) { g_create_error = std::string("tf.Example lacks the bytes feature '") + need[i] + "'"; return PGEN_ERR_ARG; }
  }
  out->uuid = got[0]->value.data(); out->uuid_len = got[0]->value.size();
  out->article = got[1]->value.data(); out->article_len = got[1]->value.size();
  out->reference = got[2]->value.data(); out->reference_len = got[2]->value.size();
  if (const pgn::BytesFeature* s = find_feature(fs, "summary")) { out->summary = s->value.data(); out->summary_len = s->value.size(); }
  return PGEN_OK;
}

int pgen_example_write(const pgen_row_t* row, void* out, size_t cap, size_t* len) {
  if (!row || !len) return PGEN_ERR_ARG;
  auto sv = [](const char* p, size_t n) { return std::string_view(p ? p : "", n); };
  std::vector<pgn::BytesFeature> fs = {{"uuid", sv(row->uuid, row->uuid_len)}, {"article", sv(row->article, row->article_len)},
                                       {"summary", sv(row->summary, row->summary_len)},
                                       {"reference", sv(row->reference, row->reference_len)}};   // flink_writer.py:26-33
  std::string s;
  pgn::write_example(fs, &s);
  *len = s.size();
  if (out && cap >= s.size()) memcpy(out, s.data(), s.size());
  else if (out) { g_create_error = "pgen_example_write: buffer too small"; return PGEN_ERR_ARG; }
  return PGEN_OK;
}

int pgen_submit_example(pgen_t* h, const void* example, size_t len) {
  if (!h) return PGEN_ERR_ARG;
  pgen_row_t r;
  int rc = pgen_example_parse(example, len, &r);
  if (rc != PGEN_OK) return fail(h, rc, g_create_error);
  return pgen_submit(h, r.uuid, r.uuid_len, r.article, r.article_len, r.reference, r.reference_len);
}

int pgen_debug_vocab_logits(pgen_t* h, int rows, const float* out_dev, float* logits_dev, int use_tc) {
  if (!h || !out_dev || !logits_dev) return PGEN_ERR_ARG;
  if (slot_busy(h)) return busy_error(h, "pgen_debug_vocab_logits");
  std::string err = h->eng().vocab_logits_op(rows, out_dev, logits_dev, use_tc);
  return err.empty() ? PGEN_OK : fail(h, PGEN_ERR_CUDA, err);
}

// Where is the handle's thread?  Callable from ANOTHER thread while a call on `h` is blocked (reads breadcrumbs without
// locks; device values through a stream of its own): out[0] handle breadcrumb, out[1] next slot, out[2] batches in
// flight, then per slot (breadcrumb, live articles, steps done by article 0, attention ticket).  Diagnostic only.
int pgen_debug_where(pgen_t* h, long long* out, int cap) {
  if (!h || !out || cap < 3) return PGEN_ERR_ARG;
  out[0] = h->where; out[1] = h->next_slot; out[2] = (long long)h->inflight.size();
  int n = 3;
  for (size_t i = 0; i < h->engs.size() && n + 4 <= cap; ++i, n += 4) {
    out[n] = h->engs[i]->where_;
    h->engs[i]->debug_progress(out + n + 1);
  }
  if (!h->engs.empty() && n + 48 <= cap) { h->engs[0]->debug_live(out + n); n += 48; }
  return n;
}

int pgen_debug_fetch(pgen_t* h, const char* name, void* out, size_t cap_bytes, size_t* n) {
  if (!h || !name || !out || !n) return PGEN_ERR_ARG;
  std::string err = h->eng().fetch(name, out, cap_bytes, n);
  return err.empty() ? PGEN_OK : fail(h, PGEN_ERR_ARG, err);
}

int pgen_set_option(pgen_t* h, const char* key_value) {
  if (!h || !key_value) return PGEN_ERR_ARG;
  if (!h->inflight.empty())
    return fail(h, PGEN_ERR_STATE, "pgen_set_option: batches of pgen_submit rows are in flight; pgen_poll(flush) first");
  for (auto& e : h->engs) {
    std::string err = e->set_option(key_value);
    if (!err.empty()) return fail(h, PGEN_ERR_ARG, err);
  }
  return PGEN_OK;
}

}  // extern "C"
