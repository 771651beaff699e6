/* libpgen_cuda.so -- C ABI of the B200-native pointer-generator beam-search decode.
 *
 * This is the drop-in boundary for the hot path behind TFModel.transform
 * (reference: src/main/java/org/apache/flink/table/ml/lib/tensorflow/TFModel.java:55-81).
 * In the reference, TFUtils.inference (TFModel.java:76) adds a flatMap operator whose body
 * hands each row to a Python/TensorFlow worker process (run_summarization.py:402-426).  Here
 * the operator body calls these entry points instead (JNI / cgo / ctypes stubs are shown in
 * INTEGRATION.md).  Plain C types only: pointers, sizes, ints.  No CPU fallback exists:
 * every entry point that needs the GPU fails with PGEN_ERR_CUDA when there is none.
 *
 * Threading: a handle is single-owner -- all calls on one handle come from one thread (the
 * Flink subtask's task thread).  Different handles are independent.
 *
 * Paths in the citations are relative to /root/reference/src/main/python/pointer-generator/
 * unless they start with "src/".
 */
#ifndef PGEN_H_
#define PGEN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define PGEN_API __attribute__((visibility("default")))
#else
#define PGEN_API
#endif

#define PGEN_OK            0
#define PGEN_ERR_ARG       1   /* bad argument / bad flag value                           */
#define PGEN_ERR_IO        2   /* vocab / log dir / weights archive missing or malformed   */
#define PGEN_ERR_CUDA      3   /* no device, CUDA runtime error                            */
#define PGEN_ERR_STATE     4   /* call sequence error (e.g. download before run)           */
#define PGEN_ERR_UNSUPPORTED 5 /* mode != decode, dims the kernels do not cover            */

typedef struct pgen pgen_t;

/* One output row {uuid, article, summary, reference}: the schema FlinkWriter.write_result emits
 * (flink_writer.py:25-34; column order from src/test/java/.../TensorFlowTest.java:232-233).
 * Byte strings, not NUL-terminated contract-wise (a trailing NUL is present for convenience). */
typedef struct {
  const char* uuid;      size_t uuid_len;
  const char* article;   size_t article_len;
  const char* summary;   size_t summary_len;
  const char* reference; size_t reference_len;
} pgen_row_t;

typedef struct {
  pgen_row_t* rows;
  size_t n_rows;
  void* owner_;          /* library-private; released by pgen_rows_free */
} pgen_rows_t;

/* Counters in place of the reference's tf.logging lines (decode.py:185, batcher.py:574-575). */
typedef struct {
  uint64_t rows_in, rows_out, batches, tokens_in, tokens_out;
  uint64_t decode_steps;       /* beam-search steps launched (per batch, summed)             */
  uint64_t article_steps;      /* sum over articles of the steps each stayed active         */
  uint64_t early_stops;        /* articles that ended with beam_size finished hypotheses     */
  uint64_t kernel_launches;    /* CUDA kernels launched by this handle                       */
  double ms_host_text;         /* tokenise + id maps + detokenise                            */
  double ms_h2d, ms_encoder, ms_decoder, ms_d2h;   /* CUDA-event times of the last batch    */
  double ms_attention, ms_vocab;   /* summed over the steps of the last batch run with option profile=1 */
  double ms_select;                /* ... same for what follows the vocabulary GEMM: re-scoring, merge, top-2K */
  double prof_steps;               /* number of steps those two sums cover */
  /* tf.nn.top_k (model.py:284) is exact; the tensor-core pass picks candidates from tf32 scores.  Every decoder row
   * is checked against an error bound after its exact re-scoring: rows that needed extra candidate groups, and rows
   * that had to score the whole vocabulary in fp32 (flat or tied distributions).  Either way the row's top-2K is exact. */
  uint64_t selection_widened, selection_fallbacks;
  double ms_enc_recurrence, ms_enc_features;   /* parts of ms_encoder of the last batch: BiLSTM recurrence, W_h.h_i product */
} pgen_stats_t;

/* Replaces main_on_flink/inference_on_flink/default_setup (run_summarization.py:402-426,
 * 384-399, 295-330) and BeamSearchDecoder.__init__ (decode.py:42-86): parses the space-joined
 * hyper-parameter string TFModel stores under its hyper-params key (TFModel.java:47; token 0 is a
 * dummy program name, unknown flags are ignored like known_only=True), loads the vocabulary
 * (data.py:40-75) and the weights from <log_root>/<exp_name>/train (util.py:29-41; here
 * model.pgenw, no retry loop: a missing file is an immediate PGEN_ERR_IO), and binds the handle to
 * CUDA device `device_ordinal` (one GPU per Flink subtask).  On failure *out is NULL and
 * pgen_last_error(NULL) holds the message for the calling thread. */
PGEN_API int pgen_create(const char* hyperparams, int device_ordinal, pgen_t** out);

/* Replaces MLMapFunction.flatMap's write of one row into the Python input queue and
 * FlinkInferenceBatcher.next_batch (batcher.py:565-585) + Example/Batch construction
 * (batcher.py:36-81, 141-187).  Inputs are borrowed for the call only.  When
 * `articles_per_batch` rows are buffered the batch is decoded. */
PGEN_API int pgen_submit(pgen_t* h, const char* uuid, size_t uuid_len, const char* article, size_t article_len,
                const char* reference, size_t reference_len);

/* Replaces draining the Python output queue (FlinkWriter.write_result, flink_writer.py:25-34;
 * decode_one_batch + write_for_flink, decode.py:95-129, 159-185).  Returns finished rows in
 * submission order.  flush != 0 (or the flush deadline having passed) decodes a partial batch,
 * so a lone row is never held waiting for the next one (the reference's known defect,
 * doc/Flink-AI-Extended Integration Report.md:341-376). */
PGEN_API int pgen_poll(pgen_t* h, int flush, pgen_rows_t* out);
PGEN_API void pgen_rows_free(pgen_rows_t* rows);

PGEN_API int pgen_stats(pgen_t* h, pgen_stats_t* out);
/* Message of the last failure on `h`; with h == NULL, of the last failed pgen_create on this thread. */
PGEN_API const char* pgen_last_error(pgen_t* h);
/* Replaces the end of BeamSearchDecoder.decode when the stream is exhausted (decode.py:139-142). */
PGEN_API void pgen_destroy(pgen_t* h);

/* Number of CUDA devices this process can use (0: none, the library then serves nothing).  The operator maps subtask
 * index -> device ordinal = index % pgen_device_count(): one GPU per Flink subtask, worker_num parallel subtasks
 * (src/main/java/org/apache/flink/table/ml/lib/tensorflow/param/HasClusterConfig.java:20-24).  No counterpart in
 * the reference, which runs on CPU (model.py:313). */
PGEN_API int pgen_device_count(void);

/* Parsed hyper-parameters (run_summarization.py:48-88 names).  Returns the value of an integer or
 * boolean flag after the decode-mode overrides (batch_size := beam_size, :312-313), or -1. */
PGEN_API int64_t pgen_flag_int(pgen_t* h, const char* name);

/* Host text path alone, for parity tests against data.py:144-168 / batcher.py:51-56:
 * article bytes -> enc ids (OOV -> [UNK]) and extended ids (j-th distinct OOV -> vocab_size+j). */
PGEN_API int pgen_text_encode(pgen_t* h, const char* article, size_t article_len, int32_t* enc_ids,
                     int32_t* ext_ids, int cap, int32_t* enc_len, int32_t* n_oov);
/* ids -> summary string (data.py:196-219, decode.py:109-118, 167-182) for the article last passed
 * to pgen_text_encode on this handle.  Writes at most cap bytes, returns the full length in *len. */
PGEN_API int pgen_text_summary(pgen_t* h, const int32_t* tokens, int n_tokens, char* out, size_t cap, size_t* len);

/* The flag parser alone (no handle, no device): parses `hyperparams` the way pgen_create does -- the flag table of
 * run_summarization.py:48-88 (names, types, defaults), token 0 a dummy program name, unknown flags ignored
 * (known_only=True, :420), --noflag for booleans, batch_size := beam_size in decode mode (:312-313) -- and stores the
 * value of the integer or boolean flag `name` in *value.  PGEN_ERR_ARG for a malformed string (message via
 * pgen_last_error(NULL)) or a name that is not an integer / boolean flag. */
PGEN_API int pgen_parse_flag(const char* hyperparams, const char* name, int64_t* value);

/* The host text path on a handle of its own, WITHOUT a device: vocabulary (data.py:40-91), article -> ids
 * (Example.__init__, batcher.py:51-56, 66-68; article2ids, data.py:144-168) and ids -> summary string
 * (outputids2words, data.py:196-219; decode.py:109-118, 167-182).  For operators that tokenise upstream of the GPU
 * stage, and for parity tests that need no GPU: tests/test_reference_golden.py holds these calls to what the
 * reference's own data.py / batcher.py / decode.py produce.  Same code as pgen_text_encode / pgen_text_summary
 * (csrc/text.cc).  pgen_text_decode_summary refers to the article last passed to pgen_text_encode_article on `t`.
 * On a failed open *out is NULL and pgen_last_error(NULL) holds the message. */
typedef struct pgen_text pgen_text_t;
PGEN_API int pgen_text_open(const char* vocab_path, int64_t vocab_size, int max_enc_steps, pgen_text_t** out);
PGEN_API int pgen_text_vocab_size(pgen_text_t* t);
PGEN_API int pgen_text_encode_article(pgen_text_t* t, const char* article, size_t article_len, int32_t* enc_ids,
                             int32_t* ext_ids, int cap, int32_t* enc_len, int32_t* n_oov);
PGEN_API int pgen_text_decode_summary(pgen_text_t* t, const int32_t* tokens, int n_tokens, char* out, size_t cap, size_t* len);
PGEN_API void pgen_text_close(pgen_text_t* t);

/* Tensor-level decode = run_beam_search (beam_search.py:82-168) over run_encoder / decode_onestep
 * (model.py:347-443) for n_articles independent articles, host buffers in and out:
 *   enc_ids, enc_ext_ids  [n_articles, ld]  (row b valid for enc_lens[b] <= max_enc_steps entries)
 *   out_tokens, out_logprobs [n_articles, max_dec_steps]  tokens after [START]; a finished
 *                            hypothesis ends with [STOP]; out_lens[b] entries are valid
 *   out_avg [n_articles]  avg_log_prob of the returned hypothesis (beam_search.py:77-79)
 * Articles are processed in device batches of at most articles_per_batch. */
PGEN_API int pgen_decode_ids(pgen_t* h, int n_articles, const int32_t* enc_ids, const int32_t* enc_ext_ids, int ld,
                    const int32_t* enc_lens, const int32_t* n_oovs, int32_t* out_tokens, int32_t* out_lens,
                    float* out_logprobs, double* out_avg);

/* The same path split at the PCIe boundary, so that a harness can time the device-resident part
 * alone: upload (H2D of ids), run (encoder + beam search, asynchronous on the handle's stream),
 * sync, download (D2H of the best hypotheses).  n_articles <= articles_per_batch.
 * These calls (and pgen_decode_ids, pgen_attention, pgen_debug_*) drive the engine of the slot chosen with
 * pgen_stage_slot directly; while a batch of pgen_submit rows is in flight on that slot they fail with
 * PGEN_ERR_STATE -- pgen_poll(flush) first, or keep the row API and the stage API on different handles. */
PGEN_API int pgen_stage_upload(pgen_t* h, int n_articles, const int32_t* enc_ids, const int32_t* enc_ext_ids, int ld,
                      const int32_t* enc_lens, const int32_t* n_oovs);
PGEN_API int pgen_stage_run(pgen_t* h);
PGEN_API int pgen_stage_sync(pgen_t* h);
PGEN_API int pgen_stage_download(pgen_t* h, int32_t* out_tokens, int32_t* out_lens, float* out_logprobs, double* out_avg);
/* A handle keeps `--slots` (default 4) device batches in flight, each on its own stream; the stage calls above
 * and pgen_stream / pgen_debug_fetch address the slot chosen here (default 0). */
PGEN_API int pgen_stage_slot(pgen_t* h, int slot);
/* The handle's cudaStream_t (as void*), so that a harness can record CUDA events on it. */
PGEN_API void* pgen_stream(pgen_t* h);

/* tf.Example wire format of the rows: what the reference moves through its Java<->Python queues.
 * pgen_example_parse replaces tf.parse_single_example with the three FixedLenFeature([], tf.string) features of
 * FlinkInferenceBatcher (batcher.py:554-557): it fills uuid / article / reference (and summary when present) of *out
 * with views INTO `example`; a missing feature, or one that is not a single bytes value, is PGEN_ERR_ARG (message via
 * pgen_last_error(NULL)).  pgen_example_write replaces FlinkWriter.write_result's serialisation
 * (flink_writer.py:26-34): features uuid, article, summary, reference; *len receives the size, call with out == NULL
 * to query it.  Both are stateless and need no GPU.  pgen_submit_example = parse + pgen_submit. */
PGEN_API int pgen_example_parse(const void* example, size_t len, pgen_row_t* out);
PGEN_API int pgen_example_write(const pgen_row_t* row, void* out, size_t cap, size_t* len);
PGEN_API int pgen_submit_example(pgen_t* h, const void* example, size_t len);

/* Kernel-level entry: one attention pass (attention_decoder.py:79-129 -- energies with coverage,
 * masked softmax, re-normalisation, context vector) over n_articles x beam_size decoder rows, using the
 * handle's attention weights v / w_c.  ALL pointers are DEVICE memory on the handle's device:
 *   enc_feat, enc_states [n_articles, max_enc_steps, 2H]   W_h.h_i and h_i
 *   lens [n_articles] (int32)   dfeat [n_articles*beam_size, 2H] = W_s s_t + b_attn
 *   cov [n_articles*beam_size, max_enc_steps] (ignored when coverage=0)
 *   attn [n_articles*beam_size, max_enc_steps]   ctx [n_articles*beam_size, 2H]      (outputs)
 * attn[r, t] is written for every t < max_enc_steps: positions at or past lens[b] hold exactly 0, like the
 * reference's masked distribution (attention_decoder.py:99); articles with lens[b] == 0 are left untouched.
 * With hidden_dim 256 and beam_size 4 or 8 this is the step loop's one-launch kernel (csrc/attn_fused.cu), otherwise
 * the energy + context kernel pair.  The scratch (energies, per-segment partials) is cudaMalloc'ed for the call.
 * Synchronous: returns after the kernels have finished. */
PGEN_API int pgen_attention(pgen_t* h, int n_articles, const float* enc_feat, const float* enc_states,
                   const int32_t* lens, const float* dfeat, const float* cov, float* attn, float* ctx);

/* No counterpart in the reference (it never touches device memory).  Same computation, arguments and
 * results as pgen_attention, for callers that are not allowed to allocate device memory: an operator
 * embedded in a process whose GPU memory belongs entirely to another allocator or to a fixed budget, or a
 * service loop that cannot afford the device-wide stall and driver lock of cudaMalloc/cudaFree.  The
 * energies scratch (plus one flag per article) lives in a fixed device array inside the library,
 * PGEN_ATTENTION_STATIC_SCRATCH_BYTES large, which every GPU that loads the library holds until the
 * process exits.  The call makes no cudaMalloc/cudaFree and calls no allocating library routine.
 *   Limit: round_up(4*n_articles, 256) + 4*n_articles*beam_size*max_enc_steps bytes must fit in the array
 *   (e.g. 163 articles at beam 4 x 400 positions); larger requests fail with PGEN_ERR_ARG.
 *   Set-up: the first call on a device looks up the array's device address (once per device); make that
 *   first call before starting any stream capture.
 *   Sharing: all calls on a device use the same array, which keeps the previous call's contents (the
 *   call rewrites what it needs); two calls whose GPU work overlaps in time race on it -- serialise them. */
#define PGEN_ATTENTION_STATIC_SCRATCH_BYTES (1u << 20)
PGEN_API int pgen_attention_static(pgen_t* h, int n_articles, const float* enc_feat, const float* enc_states,
                          const int32_t* lens, const float* dfeat, const float* cov, float* attn, float* ctx);

/* Test hooks: copy a named device buffer of the last staged batch to the host.  Names:
 * "enc_states" [B,T,2H], "enc_feat" [B,T,2H], "dec_c0" / "dec_h0" [B,H], "trace_tokens" /
 * "trace_parents" [B,S,K] (int32), "trace_logp" [B,S,K].  Returns the element count in *n. */
PGEN_API int pgen_debug_fetch(pgen_t* h, const char* name, void* out, size_t cap_bytes, size_t* n);
/* Diagnostic for a call that does not return: the ONE entry point that may be called from another thread while a call
 * on `h` is blocked.  out[0] = which stage the handle's thread is in (30 collecting the oldest batch, 40..43 launching
 * one), out[1] = next slot, out[2] = batches in flight, then per slot four values: the engine's stage (1 upload, 10..19
 * inside run, 20 waiting for the stream, 21 after), live articles, steps done by article 0, the attention ticket
 * counter -- the last three read from the device on a stream of their own.  Returns the number of values written. */
PGEN_API int pgen_debug_where(pgen_t* h, long long* out, int cap);
/* Test hook: the vocabulary scores out . w + v (model.py:236) for `rows` decoder outputs, DEVICE pointers
 * out_dev [rows, H] -> logits_dev [rows, V]; use_tc != 0 runs the tcgen05 tf32 kernel (scores carry tf32
 * rounding), 0 the fp32 SIMT kernel. */
PGEN_API int pgen_debug_vocab_logits(pgen_t* h, int rows, const float* out_dev, float* logits_dev, int use_tc);
/* Run-time switches (tests, profiling, A/B measurements), "key=value":
 *   vocab=tc|simt        tcgen05 tf32 vocabulary pass + exact re-scoring, or the all-fp32 dense path
 *   features=tc|simt     W_h.h_i by the 3xTF32 tcgen05 GEMM, or the fp32 SIMT GEMM
 *   encoder=persistent|steps   one cluster kernel for the whole recurrence, or one launch per time step
 *   decoder=fused|split  one-launch decoder cell + fused out projection / p_gen (csrc/dec_cell.cu), or gather + four
 *                        split-K GEMM launches + point-wise kernels
 *   attention=fused|split   one-launch attention (csrc/attn_fused.cu), or the energy + context kernel pair
 *   graph=1|0, graph_steps=N   replay N (even) decode steps per CUDA graph, or launch every kernel
 *   check_every=N        poll the device every N steps after min_dec_steps and stop early when every article is done
 *   vocab_guard=auto|widen|dense|off   exactness guard of the tensor-core candidate selection: auto (default), or force
 *                        the widened / dense re-scoring on every row (tests), or switch the check off (A-B timing)
 *   profile=1|0|N        per-step CUDA events for pgen_stats (plain launches); N > 1: sums over the first N steps only
 *                        (no article can finish before min_dec_steps, so those steps have every article live)
 *   measure=N            replay the attention pass and the tensor-core pass of the last decoded batch N times each
 *                        from a graph between two events, with an L2 flush between replays (its time subtracted); sums
 *                        land in ms_attention / ms_vocab, prof_steps = N
 * Fails with PGEN_ERR_STATE while pgen_submit rows are in flight.  No counterpart in the reference. */
PGEN_API int pgen_set_option(pgen_t* h, const char* key_value);

#ifdef __cplusplus
}
#endif
#endif  /* PGEN_H_ */
