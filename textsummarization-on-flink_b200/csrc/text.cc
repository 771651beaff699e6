#include "text.h"

#include <algorithm>
#include <cerrno>
#include <cstdlib>
#include <cstring>
#include <fstream>

namespace pgn {
namespace {

struct SpaceTable {                        // bytes.split(): ASCII whitespace = space, \t, \n, \r, \v, \f
  bool is[256] = {};
  constexpr SpaceTable() { is[(int)' '] = is[(int)'\t'] = is[(int)'\n'] = is[(int)'\r'] = is[(int)'\v'] = is[(int)'\f'] = true; }
};
constexpr SpaceTable kSpace;
inline bool is_space(unsigned char c) { return kSpace.is[c]; }

template <class F>
void for_each_word(std::string_view s, F&& f) {
  size_t i = 0, n = s.size();
  while (i < n) {
    while (i < n && is_space((unsigned char)s[i])) ++i;
    size_t j = i;
    while (j < n && !is_space((unsigned char)s[j])) ++j;
    if (j > i && !f(s.substr(i, j - i))) return;
    i = j;
  }
}

bool parse_bool(const std::string& v, bool* out) {
  std::string l;
  for (char c : v) l.push_back((char)tolower((unsigned char)c));
  if (l == "1" || l == "true" || l == "t" || l == "yes" || l == "y") { *out = true; return true; }
  if (l == "0" || l == "false" || l == "f" || l == "no" || l == "n") { *out = false; return true; }
  return false;
}

bool parse_int(const std::string& v, int64_t* out) {
  if (v.empty()) return false;
  errno = 0;
  char* end = nullptr;
  long long x = strtoll(v.c_str(), &end, 10);
  if (errno || *end) return false;
  *out = x;
  return true;
}

struct FlagRef { const char* name; int kind; void* p; };   // kind 0 str, 1 int, 2 bool, 3 float(ignored)

std::vector<FlagRef> flag_table(HyperParams* h) {
  return {
      {"data_path", 0, &h->data_path}, {"vocab_path", 0, &h->vocab_path}, {"mode", 0, &h->mode},
      {"num_steps", 1, &h->num_steps}, {"single_pass", 2, &h->single_pass}, {"inference", 2, &h->inference},
      {"log_root", 0, &h->log_root}, {"exp_name", 0, &h->exp_name}, {"hidden_dim", 1, &h->hidden_dim},
      {"emb_dim", 1, &h->emb_dim}, {"batch_size", 1, &h->batch_size}, {"max_enc_steps", 1, &h->max_enc_steps},
      {"max_dec_steps", 1, &h->max_dec_steps}, {"beam_size", 1, &h->beam_size},
      {"min_dec_steps", 1, &h->min_dec_steps}, {"vocab_size", 1, &h->vocab_size},
      {"lr", 3, nullptr}, {"adagrad_init_acc", 3, nullptr}, {"rand_unif_init_mag", 3, nullptr},
      {"trunc_norm_init_std", 3, nullptr}, {"max_grad_norm", 3, nullptr}, {"cov_loss_wt", 3, nullptr},
      {"pointer_gen", 2, &h->pointer_gen}, {"coverage", 2, &h->coverage},
      {"convert_to_coverage_model", 2, &h->convert_to_coverage_model},
      {"restore_best_model", 2, &h->restore_best_model}, {"debug", 2, &h->debug},
      {"articles_per_batch", 1, &h->articles_per_batch}, {"flush_ms", 1, &h->flush_ms}, {"slots", 1, &h->slots}, {"encoder_table", 1, &h->encoder_table},
      {"weights_path", 0, &h->weights_path}, {"decoder", 0, &h->decoder},
  };
}

}  // namespace

std::string parse_hyperparams(const std::string& s, HyperParams* hp) {
  auto table = flag_table(hp);
  // tf_hyperparameter.split(' ') (run_summarization.py:420): split on single spaces, token 0 dropped.
  std::vector<std::string> toks;
  size_t i = 0;
  while (true) {
    size_t j = s.find(' ', i);
    toks.push_back(s.substr(i, j == std::string::npos ? std::string::npos : j - i));
    if (j == std::string::npos) break;
    i = j + 1;
  }
  for (size_t t = 1; t < toks.size(); ++t) {
    const std::string& tok = toks[t];
    if (tok.size() < 3 || tok[0] != '-' || tok[1] != '-') continue;
    std::string body = tok.substr(2), name, val;
    bool has_val = false;
    size_t eq = body.find('=');
    if (eq == std::string::npos) name = body;
    else { name = body.substr(0, eq); val = body.substr(eq + 1); has_val = true; }
    const FlagRef* f = nullptr;
    for (auto& e : table) if (name == e.name) f = &e;
    if (!f && !has_val && name.rfind("no", 0) == 0) {        // --noflag for booleans
      for (auto& e : table) if (e.kind == 2 && name.substr(2) == e.name) { *(bool*)e.p = false; f = &e; }
      if (f) continue;
    }
    if (!f) continue;                                        // known_only=True
    switch (f->kind) {
      case 0:
        if (!has_val) return "flag --" + name + " needs a value";
        *(std::string*)f->p = val;
        break;
      case 1:
        if (!has_val || !parse_int(val, (int64_t*)f->p)) return "flag --" + name + ": bad integer value '" + val + "'";
        break;
      case 2:
        if (!has_val) *(bool*)f->p = true;
        else if (!parse_bool(val, (bool*)f->p)) return "flag --" + name + ": bad boolean value '" + val + "'";
        break;
      default:
        break;                                               // float flags: training only
    }
  }
  if (hp->decoder != "auto" && hp->decoder != "fused" && hp->decoder != "split") return "flag --decoder must be auto, fused or split";
  if (hp->mode == "decode") hp->batch_size = hp->beam_size;  // run_summarization.py:312-313
  return "";
}

int64_t flag_int(const HyperParams& hp, const std::string& name) {
  HyperParams copy = hp;
  for (auto& e : flag_table(&copy)) {
    if (name != e.name) continue;
    if (e.kind == 1) return *(int64_t*)e.p;
    if (e.kind == 2) return *(bool*)e.p ? 1 : 0;
  }
  return -1;
}

bool Vocab::contains(std::string_view w) const {
  if (slots_.empty()) return false;
  const uint64_t pre = prefix8(w);
  for (size_t i = (size_t)(mix(pre, w.size()) & mask_);; i = (i + 1) & mask_) {
    const Slot& s = slots_[i];
    if (s.id < 0) return false;
    if (s.prefix == pre && s.len == (uint32_t)w.size() && words_[s.id] == w) return true;
  }
}

void Vocab::insert(int id) {
  const std::string& w = words_[id];
  const uint64_t pre = prefix8(w);
  size_t i = (size_t)(mix(pre, w.size()) & mask_);
  while (slots_[i].id >= 0) i = (i + 1) & mask_;
  slots_[i] = Slot{pre, (uint32_t)w.size(), id};
}

void Vocab::rehash(size_t n_slots) {
  slots_.assign(n_slots, Slot{0, 0, -1});
  mask_ = n_slots - 1;
  for (int id = 0; id < (int)words_.size(); ++id) insert(id);
}

std::string Vocab::load(const std::string& path, int64_t max_size) {
  words_.clear();
  warnings_ = 0;
  rehash(1024);
  auto add = [&](std::string w) {
    words_.push_back(std::move(w));
    if (words_.size() * 2 > slots_.size()) rehash(slots_.size() * 4);       // load factor <= 1/2
    else insert((int)words_.size() - 1);
  };
  for (const char* w : {"[UNK]", "[PAD]", "[START]", "[STOP]"}) add(w);        // data.py:50-54
  std::ifstream f(path, std::ios::binary);
  if (!f) return "cannot open vocab file '" + path + "'";
  std::string line;
  while (std::getline(f, line)) {
    std::string_view pieces[3];
    int n = 0;
    for_each_word(line, [&](std::string_view w) { if (n < 3) pieces[n] = w; ++n; return n <= 2; });
    if (n != 2) { ++warnings_; continue; }                                      // data.py:59-62
    std::string_view w = pieces[0];
    for (const char* bad : {"<s>", "</s>", "[UNK]", "[PAD]", "[START]", "[STOP]"})
      if (w == bad) return "<s>, </s>, [UNK], [PAD], [START] and [STOP] shouldn't be in the vocab file, but " +
                           std::string(w) + " is";                              // data.py:64-65
    if (contains(w)) return "Duplicated word in vocabulary file: " + std::string(w);   // data.py:66-67
    add(std::string(w));
    if (max_size != 0 && (int64_t)words_.size() >= max_size) break;             // data.py:71-73
  }
  return "";
}

void encode_article(const Vocab& v, std::string_view article, int max_enc_steps, EncodedArticle* out) {
  out->enc_ids.clear();
  out->ext_ids.clear();
  out->oovs.clear();
  const size_t guess = std::min<size_t>((size_t)max_enc_steps, article.size() / 2 + 1);
  out->enc_ids.reserve(guess);
  out->ext_ids.reserve(guess);
  // In-article OOVs in first-appearance order (data.py:156-164).  Few of them: a linear scan over the list; the
  // reference's own `oovs.index(w)` is linear too.  Many (BASELINE configs[3]: 200 per article): an index on the side.
  std::vector<std::string_view> seen;                      // views into `article`, alive for this call
  std::unordered_map<std::string_view, int> oov_index;
  constexpr size_t kLinear = 12;
  const int V = v.size();
  for_each_word(article, [&](std::string_view w) {
    if ((int)out->enc_ids.size() >= max_enc_steps) return false;               // batcher.py:52-54
    int id = v.word2id(w);
    out->enc_ids.push_back(id);                                                // batcher.py:56
    if (id == kUnkId) {                                                        // data.py:159-164
      int j = -1;
      if (seen.size() <= kLinear) {
        for (size_t q = 0; q < seen.size(); ++q)
          if (seen[q] == w) { j = (int)q; break; }
      } else {
        auto it = oov_index.find(w);
        if (it != oov_index.end()) j = it->second;
      }
      if (j < 0) {
        j = (int)seen.size();
        seen.push_back(w);
        out->oovs.emplace_back(w);
        if (seen.size() > kLinear) {
          if (oov_index.empty())                                               // crossing the threshold: index what is there
            for (size_t q = 0; q + 1 < seen.size(); ++q) oov_index.emplace(seen[q], (int)q);
          oov_index.emplace(w, j);
        }
      }
      out->ext_ids.push_back(V + j);
    } else {
      out->ext_ids.push_back(id);
    }
    return true;
  });
}

bool summary_from_ids(const Vocab& v, const std::vector<std::string>& oovs, const int32_t* tokens,
                      int n_tokens, std::string* out) {
  out->clear();
  const int V = v.size();
  bool sent_start = true, first = true;
  for (int i = 0; i < n_tokens; ++i) {
    int id = tokens[i];
    if (id == kStopId) break;                                                  // decode.py:113-118
    const std::string* w;
    if (id >= 0 && id < V) w = &v.id2word(id);                                 // data.py:211-212
    else if (id >= V && id - V < (int)oovs.size()) w = &oovs[id - V];          // data.py:213-215
    else return false;
    // decode.py:167-182: words of a sentence joined by ' ', sentences (ending after ".") by "  ".
    if (!first) out->append(sent_start ? "  " : " ");
    out->append(*w);
    first = false;
    sent_start = (*w == ".");
  }
  return true;
}

}  // namespace pgn
