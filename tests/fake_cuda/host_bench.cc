// Host cost of the row path per device batch (tools/host_bench.sh): submit + poll for every row exactly like the operator
// does, against the fake CUDA runtime -- what the CPU spends per batch of 64 articles on tokenising, id maps, copy lists,
// enqueueing and detokenising, with the GPU taken out of the picture.
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

#include "pgen.h"

extern "C" void fake_cuda_skip_async_memset(int);

int main(int argc, char** argv) {
  if (argc < 3) { fprintf(stderr, "usage: host_bench <hyperparams> <articles file, one per line> [batches]\n"); return 2; }
  std::vector<std::string> arts;
  std::ifstream f(argv[2]);
  for (std::string line; std::getline(f, line);) arts.push_back(line);
  const int batches = argc > 3 ? atoi(argv[3]) : 40;
  fake_cuda_skip_async_memset(1);
  pgen_t* h = nullptr;
  if (pgen_create(argv[1], 0, &h) != PGEN_OK) { fprintf(stderr, "create: %s\n", pgen_last_error(nullptr)); return 1; }
  const int per = (int)pgen_flag_int(h, "articles_per_batch");
  size_t rows = 0, bytes = 0;
  auto run = [&](int nb) {
    for (int i = 0; i < nb * per; ++i) {
      const std::string& a = arts[i % arts.size()];
      char uuid[32];
      const int nu = snprintf(uuid, sizeof uuid, "uuid-%d", i);
      if (pgen_submit(h, uuid, nu, a.data(), a.size(), "reference .", 11) != PGEN_OK) { fprintf(stderr, "submit: %s\n", pgen_last_error(h)); exit(1); }
      pgen_rows_t out;
      pgen_poll(h, 0, &out);
      rows += out.n_rows;
      for (size_t r = 0; r < out.n_rows; ++r) bytes += out.rows[r].summary_len + out.rows[r].article_len;
      pgen_rows_free(&out);
    }
    pgen_rows_t out;
    pgen_poll(h, 1, &out);
    rows += out.n_rows;
    pgen_rows_free(&out);
  };
  run(4);
  const auto t0 = std::chrono::steady_clock::now();
  run(batches);
  const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  pgen_stats_t st;
  pgen_stats(h, &st);
  printf("{\"host_ms_per_batch\": %.3f, \"articles_per_batch\": %d, \"articles_per_s_one_thread\": %.0f, \"batches\": %d, \"rows\": %zu}\n",
         ms / batches, per, 1e3 * batches * per / ms, batches, rows);
  pgen_destroy(h);
  return 0;
}
