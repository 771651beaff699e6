// ThreadSanitizer driver (tools/host_tsan.sh): N threads, one handle each on device k mod devices, all created at the same
// time (the per-device attribute caches and the lazily set function attributes are the shared state), then rows through
// submit / poll / flush -- what the subtasks of one Flink TaskManager do.  Runs against the fake CUDA runtime.
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "pgen.h"

int main(int argc, char** argv) {
  if (argc < 3) { fprintf(stderr, "usage: threads_driver <hyperparams> <n_threads>\n"); return 2; }
  const std::string hyper = argv[1];
  const int n = atoi(argv[2]);
  std::vector<int> rows_out(n, 0), failed(n, 0);
  std::vector<std::thread> ts;
  for (int k = 0; k < n; ++k)
    ts.emplace_back([&, k] {
      pgen_t* h = nullptr;
      if (pgen_create(hyper.c_str(), k % pgen_device_count(), &h) != PGEN_OK) { fprintf(stderr, "create: %s\n", pgen_last_error(nullptr)); failed[k] = 1; return; }
      for (int i = 0; i < 200; ++i) {
        char uuid[32];
        snprintf(uuid, sizeof uuid, "t%d-%d", k, i);
        const char* art = (i % 7) ? "w7 w8 . w9 zz w10 ." : "";
        if (pgen_submit(h, uuid, strlen(uuid), art, strlen(art), "r", 1) != PGEN_OK) failed[k] = 1;
        pgen_rows_t rows;
        if (pgen_poll(h, i == 199, &rows) != PGEN_OK) failed[k] = 1;
        rows_out[k] += (int)rows.n_rows;
        pgen_rows_free(&rows);
      }
      pgen_stats_t st;
      pgen_stats(h, &st);
      if (st.rows_in != st.rows_out) failed[k] = 1;
      pgen_destroy(h);
    });
  for (auto& t : ts) t.join();
  int bad = 0;
  for (int k = 0; k < n; ++k) bad += failed[k] || rows_out[k] != 200;
  printf("%d threads, %d rows each: %s\n", n, 200, bad ? "FAILED" : "ok");
  return bad ? 1 : 0;
}
