// Latency and concurrent throughput of da_store_topk, the call the Go decorator makes once per /api/query
// (cmd/query/main.go:101), measured from C++ so that no interpreter sits between the clock and the C ABI.
//
//   g++ -O2 -std=c++17 -I include tools/topk_latency.cpp -o build/topk_latency doc_agents_b200/libdocagents_b200.so \
//       -Wl,-rpath,$PWD/doc_agents_b200 -lpthread
//   build/topk_latency [rows] [docs_in_filter] [k] [calls] [threads]
//
// BASELINE configs[0]: rows = 200, 1 document, k = 5.  configs[1]: rows = 1 000 000, 16 documents, k = 10.
// Prints one JSON line: median / p99 latency of single calls (host buffers in and out), and calls/s with `threads`
// callers at once (each thread asks for its own documents).
#include "docagents_b200.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

static void die(const char *what)
{
    fprintf(stderr, "%s: %s\n", what, da_last_error());
    exit(1);
}

int main(int argc, char **argv)
{
    const uint64_t rows = argc > 1 ? strtoull(argv[1], nullptr, 10) : 200;
    const int n_filter = argc > 2 ? atoi(argv[2]) : 1;
    const int k = argc > 3 ? atoi(argv[3]) : 5;
    const int calls = argc > 4 ? atoi(argv[4]) : 2000;
    const int threads = argc > 5 ? atoi(argv[5]) : 16;
    const int dim = 3072;

    da_config cfg{};
    cfg.struct_size = sizeof cfg;
    cfg.dim = dim; cfg.n_devices = 1; cfg.device_ids[0] = 0;
    cfg.capacity_rows = rows; cfg.min_similarity = DA_DEFAULT_MIN_SIMILARITY;
    da_store *st = nullptr;
    if (da_store_create(&cfg, &st) != DA_OK) die("create");
    da_synth_params sp{};
    sp.struct_size = sizeof sp;
    sp.seed_data = 0xD0C5EED; sp.seed_centroid = 0xC0FFEE; sp.seed_ids = 0x1D5EED;
    sp.chunks_per_doc = 200; sp.n_clusters = 1024; sp.cmul = 2; sp.nmul = 1;
    sp.row_lo = 0; sp.n_rows = rows;
    if (da_store_generate_synthetic(st, &sp) != DA_OK) die("synthetic");
    const uint64_t n_docs = (rows + 199) / 200;

    // per caller: a query aimed at cluster c and n_filter documents, half of them from that cluster
    auto make_request = [&](int who, std::vector<float> &q, std::vector<uint8_t> &docs) {
        const uint32_t cluster = (uint32_t)((who * 37) % 1024) % (uint32_t)std::min<uint64_t>(n_docs, 1024);
        q.resize(dim);
        if (da_synth_query(&sp, 0xBEEF, (uint64_t)who, cluster, dim, q.data()) != DA_OK) die("query");
        docs.resize((size_t)n_filter * 16);
        for (int j = 0; j < n_filter; j++) {
            const uint64_t d = j < (n_filter + 1) / 2 ? (cluster + (uint64_t)j * 1024) % n_docs
                                                      : (7 + (uint64_t)j * (n_docs / (uint64_t)n_filter + 1) + (uint64_t)who) % n_docs;
            if (da_synth_ids(&sp, 1, d, 1, docs.data() + 16 * j) != DA_OK) die("ids");
        }
    };

    std::vector<float> q;
    std::vector<uint8_t> docs;
    make_request(0, q, docs);
    std::vector<uint8_t> out_ids((size_t)k * 16);
    std::vector<float> out_scores(k);
    int32_t n = 0;
    for (int i = 0; i < 50; i++)
        if (da_store_topk(st, docs.data(), (size_t)n_filter, q.data(), dim, k, out_ids.data(), out_scores.data(), &n) != DA_OK) die("topk");
    const int hits = n;
    std::vector<double> us(calls);
    for (int i = 0; i < calls; i++) {
        const auto t0 = std::chrono::steady_clock::now();
        if (da_store_topk(st, docs.data(), (size_t)n_filter, q.data(), dim, k, out_ids.data(), out_scores.data(), &n) != DA_OK) die("topk");
        us[i] = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
    }
    std::sort(us.begin(), us.end());
    da_stats s0{};
    da_store_stats(st, &s0);

    // concurrent callers
    std::atomic<int> go{0};
    std::atomic<long> done{0};
    std::vector<std::thread> pool;
    const int per_thread = std::max(1, calls / 4);
    for (int t = 0; t < threads; t++)
        pool.emplace_back([&, t] {
            std::vector<float> tq;
            std::vector<uint8_t> td, ids((size_t)k * 16);
            std::vector<float> sc(k);
            make_request(t, tq, td);
            int32_t tn = 0;
            da_store_topk(st, td.data(), (size_t)n_filter, tq.data(), dim, k, ids.data(), sc.data(), &tn);   // warm (slot creation)
            while (!go.load()) std::this_thread::yield();
            for (int i = 0; i < per_thread; i++) {
                if (da_store_topk(st, td.data(), (size_t)n_filter, tq.data(), dim, k, ids.data(), sc.data(), &tn) != DA_OK) die("topk (thread)");
                done++;
            }
        });
    std::this_thread::sleep_for(std::chrono::milliseconds(200));
    const auto t0 = std::chrono::steady_clock::now();
    go = 1;
    for (auto &th : pool) th.join();
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    da_stats s1{};
    da_store_stats(st, &s1);

    printf("{\"workload\": \"da_store_topk from C++: %llu x %d rows, %d document(s) in the filter, top_k=%d, floor 0.7, host buffers in and out\", "
           "\"hits\": %d, \"calls\": %d, \"latency_us\": {\"median\": %.2f, \"p10\": %.2f, \"p90\": %.2f, \"p99\": %.2f, \"min\": %.2f}, "
           "\"rows_read\": %llu, \"small_search_path\": %s, "
           "\"concurrent\": {\"threads\": %d, \"calls\": %ld, \"calls_per_s\": %.0f, \"mean_latency_us_per_thread\": %.2f}}\n",
           (unsigned long long)rows, dim, n_filter, k, hits, calls, us[calls / 2], us[calls / 10], us[calls * 9 / 10], us[calls * 99 / 100], us[0],
           (unsigned long long)s0.last_rows_read, s0.small_searches > 0 ? "true" : "false",
           threads, done.load(), done.load() / sec, sec * 1e6 * threads / (double)done.load());
    (void)s1;
    da_store_destroy(st);
    return 0;
}
