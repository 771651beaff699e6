// Host-side text path of the decode operator: flag parsing, vocabulary, article -> ids,
// ids -> summary string.  Integer / byte work only; mirrors data.py, batcher.py, decode.py and the
// flag table of run_summarization.py (paths relative to
// /root/reference/src/main/python/pointer-generator/).
#pragma once
#include <cstdint>
#include <cstring>
#include <string>
#include <string_view>
#include <unordered_map>
#include <vector>
#include <deque>

namespace pgn {

constexpr int kUnkId = 0, kPadId = 1, kStartId = 2, kStopId = 3;   // data.py:50-54

struct HyperParams {                       // run_summarization.py:48-88 (names and defaults)
  std::string data_path, vocab_path, mode = "train", log_root, exp_name;
  int64_t num_steps = 0, hidden_dim = 256, emb_dim = 128, batch_size = 16, max_enc_steps = 400,
          max_dec_steps = 100, beam_size = 4, min_dec_steps = 35, vocab_size = 50000;
  bool single_pass = false, inference = false, pointer_gen = true, coverage = false,
       convert_to_coverage_model = false, restore_best_model = false, debug = false;
  // Flags that only this library knows; the reference's parser ignores them (known_only=True).
  int64_t encoder_table = 1;               // 1: keep x.Kx + b of every vocabulary word on the device (8 V H bytes x 4)
  int64_t articles_per_batch = 64;         // device-level article batching
  int64_t flush_ms = 50;                   // deadline after which poll() decodes a partial batch
  int64_t slots = 4;                       // device batches in flight per handle (each with its own stream)
  std::string weights_path;                // overrides <log_root>/<exp_name>/train/model.pgenw
  std::string decoder = "auto";            // fused | split | auto (fused with one batch in flight, split with several)
};

// Returns "" on success, else the error message.  Token 0 is a dummy program name
// (TensorFlowTest.java:41); unknown flags are ignored; mode=decode forces batch_size := beam_size
// (run_summarization.py:312-313).
std::string parse_hyperparams(const std::string& s, HyperParams* hp);
// -1 when the name is not an integer/boolean flag.
int64_t flag_int(const HyperParams& hp, const std::string& name);

class Vocab {                              // data.py:37-91
 public:
  // Returns "" on success.  max_size == 0 reads the whole file.
  std::string load(const std::string& path, int64_t max_size);
  // data.py:77-81.  Open addressing over 16-byte slots {first 8 bytes of the word, length, id}: a word of up to 8 bytes
  // (nearly all of a vocabulary) is found with one cache line and no string comparison; longer ones compare their tail
  // against the stored word.  The tokeniser calls this once per article word: it is the host path's inner loop.
  int word2id(std::string_view w) const {
    if (slots_.empty()) return kUnkId;
    const uint64_t pre = prefix8(w);
    size_t i = (size_t)(mix(pre, w.size()) & mask_);
    for (;;) {
      const Slot& s = slots_[i];
      if (s.id < 0) return kUnkId;
      if (s.prefix == pre && s.len == (uint32_t)w.size() && (w.size() <= 8 || memcmp(words_[s.id].data() + 8, w.data() + 8, w.size() - 8) == 0))
        return s.id;
      i = (i + 1) & mask_;
    }
  }
  const std::string& id2word(int id) const { return words_[id]; }
  int size() const { return (int)words_.size(); }
  int warnings() const { return warnings_; }

 private:
  struct Slot { uint64_t prefix; uint32_t len; int32_t id; };
  static uint64_t prefix8(std::string_view w) {
    uint64_t p = 0;
    if (!w.empty()) memcpy(&p, w.data(), w.size() < 8 ? w.size() : 8);
    return p;
  }
  static uint64_t mix(uint64_t p, size_t len) {        // (words longer than 8 bytes that share a prefix and a length
    uint64_t x = (p ^ (len * 0x9E3779B97F4A7C15ull)) * 0xD6E8FEB86659FD93ull;   //  land in one probe sequence: rare)
    return x ^ (x >> 32);
  }
  bool contains(std::string_view w) const;
  void insert(int id);
  void rehash(size_t n_slots);
  std::deque<std::string> words_;
  std::vector<Slot> slots_;
  size_t mask_ = 0;
  int warnings_ = 0;
};

struct EncodedArticle {                    // Example.__init__ (batcher.py:36-81), encoder side
  std::vector<int32_t> enc_ids;            // OOV -> [UNK]              (batcher.py:56)
  std::vector<int32_t> ext_ids;            // j-th distinct OOV -> V+j  (data.py:156-168)
  std::vector<std::string> oovs;           // first-appearance order
};

void encode_article(const Vocab& v, std::string_view article, int max_enc_steps, EncodedArticle* out);

// tokens after [START] -> summary (data.py:196-219, decode.py:109-118, 167-182).  Returns false
// when a token id is outside the extended vocabulary of this article.
bool summary_from_ids(const Vocab& v, const std::vector<std::string>& oovs, const int32_t* tokens,
                      int n_tokens, std::string* out);

}  // namespace pgn
