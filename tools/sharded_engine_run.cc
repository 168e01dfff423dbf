// sharded_engine_run.cc — the in-process N-engine host driver (include/igs.hpp, igs::ShardedEngine) on real GPUs:
// one process, one engine + host thread + pinned staging per device — the shape the Go service has
// (services/risk/cmd/main.go:72-164 is one process serving every request).  Loads a synthetic player table, event
// history and a 500 x depth-8 ensemble into every shard, scores a batch of GLOBAL account ids through
// ShardedEngine::ScoreBatch, checks the results against ONE engine that holds everything (a result depends on its own
// account only, so the two must agree bit for bit) and prints one JSON line with the whole-call wall time.
//
//   g++ -std=c++17 -O2 -pthread -I include tools/sharded_engine_run.cc -L igaming-platform_b200 -ligs
//       -Wl,-rpath,$PWD/igaming-platform_b200 -o sharded_engine_run;  ./sharded_engine_run <n_devices> [n_tx]
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "igs.hpp"

static uint64_t g_state = 0x9E3779B97F4A7C15ull;
static inline uint64_t rnd() {
    g_state ^= g_state << 13; g_state ^= g_state >> 7; g_state ^= g_state << 17;
    return g_state;
}

int main(int argc, char** argv) {
    const int n_dev = argc > 1 ? std::atoi(argv[1]) : 1;
    const uint64_t n_tx = argc > 2 ? std::strtoull(argv[2], nullptr, 10) : 4000000ull;
    const uint32_t n_players = 1000000, n_events = 5000000, n_trees = 500;
    const int64_t now = 1750000000;
    try {
        std::vector<igs_player_row> pl(n_players);
        for (auto& p : pl) {
            p = igs_player_row{};
            p.account_created_at = now - (int64_t)(rnd() % 600) * 86400;
            p.total_deposits = (int64_t)(rnd() % 300000); p.total_withdrawals = (int64_t)(rnd() % 300000);
            p.deposit_count = (int32_t)(rnd() % 9); p.bet_count = (int32_t)(rnd() % 500); p.win_count = p.bet_count / 3;
            p.bonus_claim_count = (int32_t)(rnd() % 6);
        }
        std::vector<uint32_t> a(n_events), t(n_events); std::vector<int64_t> m(n_events); std::vector<uint8_t> y(n_events);
        for (uint32_t i = 0; i < n_events; ++i) {
            a[i] = (uint32_t)(rnd() % n_players); t[i] = (uint32_t)(now - 86400 + (int64_t)(rnd() % 86400));
            m[i] = 100 * (int64_t)(1 + rnd() % 500); y[i] = (uint8_t)(rnd() % 3);
        }
        std::vector<uint8_t> feat((size_t)n_trees * 255); std::vector<float> thr((size_t)n_trees * 255), leaf((size_t)n_trees * 256);
        for (auto& f : feat) f = (uint8_t)(rnd() % 42);
        for (auto& x : thr) x = (float)(rnd() % 1000) / 1000.0f;
        for (auto& x : leaf) x = ((float)(rnd() % 2001) - 1000.0f) / 50000.0f;
        igs_gbdt_desc gd{};
        gd.n_trees = n_trees; gd.n_features = IGS_NUM_FEATURES; gd.post_transform = IGS_POST_LOGISTIC; gd.base_score = 0.f;
        gd.feat = feat.data(); gd.thr = thr.data(); gd.leaf = leaf.data();
        std::vector<igs_tx> tx(n_tx);
        for (auto& x : tx) {
            x = igs_tx{};
            x.account = (uint32_t)(rnd() % n_players);
            x.ts = (uint32_t)(now - 600 + (int64_t)(rnd() % 600)); x.seq_limit = 0xFFFFFFFFu;
            x.amount = (int64_t)(rnd() % 400000); x.tx_type = (uint8_t)(rnd() % 3);
            x.unique_devices_24h = (uint16_t)(rnd() % 6); x.unique_ips_24h = (uint16_t)(rnd() % 8);
            x.bits = IGS_BIT_VALID;
        }
        igs_config cfg = igs::DefaultConfig();
        cfg.model_kind = IGS_MODEL_GBDT;

        igs::ShardedEngine sharded(n_dev, n_dev);                  // shard s on device s
        sharded.LoadGbdt(gd);
        sharded.LoadPlayers(pl.data(), n_players);
        sharded.LoadEvents(a.data(), t.data(), m.data(), y.data(), n_events);
        std::vector<igs_result> got(n_tx);
        sharded.ScoreBatch(cfg, tx.data(), n_tx, got.data());      // warm-up: staging buffers, first launches
        double best = 1e30;
        for (int rep = 0; rep < 3; ++rep) {
            const auto t0 = std::chrono::steady_clock::now();
            sharded.ScoreBatch(cfg, tx.data(), n_tx, got.data());
            best = std::min(best, std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
        }

        // one engine with everything, a sample of the batch
        const uint64_t n_chk = std::min<uint64_t>(n_tx, 400000);
        igs_engine* one = nullptr;
        if (igs_engine_create(0, &one) != IGS_OK) throw igs::Error(IGS_ERR_CUDA, igs_last_error(nullptr));
        std::vector<igs_result> want(n_chk);
        const igs_event_cols cols{a.data(), t.data(), m.data(), y.data()};
        int rc = igs_load_gbdt(one, &gd);
        if (rc == IGS_OK) rc = igs_load_players(one, pl.data(), n_players);
        if (rc == IGS_OK) rc = igs_load_events(one, &cols, n_events, n_players);
        if (rc == IGS_OK) rc = igs_score_batch(one, &cfg, tx.data(), n_chk, want.data(), nullptr, nullptr);
        if (rc != IGS_OK) throw igs::Error(rc, igs_last_error(one));
        igs_engine_destroy(one);
        const bool same = std::memcmp(got.data(), want.data(), n_chk * sizeof(igs_result)) == 0;
        std::printf("{\"tool\": \"sharded_engine_run\", \"devices\": %d, \"shards\": %d, \"tx\": %llu, \"players\": %u, "
                    "\"events\": %u, \"call_wall_ms\": %.3f, \"tx_per_s\": %.4g, \"checked_rows\": %llu, "
                    "\"bit_identical_to_one_engine\": %s, \"note\": \"ShardedEngine::ScoreBatch wall clock: routing by "
                    "igs_shard_of and the gather into per-shard pinned staging on the host, one host thread per shard "
                    "calling igs_score_batch, scatter of the results\"}\n",
                    n_dev, sharded.n_shards(), (unsigned long long)n_tx, n_players, n_events, best * 1e3,
                    (double)n_tx / best, (unsigned long long)n_chk, same ? "true" : "false");
        return same ? 0 : 1;
    } catch (const igs::Error& e) {
        std::fprintf(stderr, "sharded_engine_run: %s\n", e.what());
        return 2;
    }
}
