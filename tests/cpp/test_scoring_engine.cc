// C++ host mirror (include/igs.hpp) driven the way a test of scoring.ScoringEngine would read, checked
// against the oracle's per-request pipeline (the oracle is linked here as the checker only).
//   usage: test_scoring_engine            -> needs a B200; exit 0 on success
//          test_scoring_engine --compile-only-check  -> returns 0 without touching the GPU
#include <cinttypes>
#include <cstdio>
#include <cstring>
#include <random>

#include "igs.hpp"
#include "../../oracle/igs_oracle.h"

#define REQUIRE(c)                                                         \
    do {                                                                   \
        if (!(c)) { std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } \
    } while (0)

static igs::AccountID make_id(uint32_t k) {
    igs::AccountID id{};
    std::memcpy(id.data() + 12, &k, 4);
    id[0] = 0xAB;
    return id;
}

int main(int argc, char** argv) {
    if (argc > 1 && std::strcmp(argv[1], "--compile-only-check") == 0) {
        igs_config c = igs::DefaultConfig();
        REQUIRE(c.block_threshold == 80 && c.review_threshold == 50);
        return 0;
    }
    const int n_players = 200, n_events = 6000;
    const int64_t now = 1700003600;
    std::mt19937_64 rng(7);
    std::vector<igs_player_row> players(n_players);
    igs::PlayerDirectory dir;
    for (int i = 0; i < n_players; ++i) {
        igs_player_row p{};
        p.account_created_at = now - (int64_t)(rng() % 900) * 86400 - (int64_t)(rng() % 86400);
        p.total_deposits = (int64_t)(rng() % 200000);
        p.total_withdrawals = (int64_t)(rng() % 250000);
        p.deposit_count = (int32_t)(rng() % 10);
        p.withdraw_count = (int32_t)(rng() % 5);
        p.bet_count = (int32_t)(rng() % 1000);
        p.win_count = p.bet_count / 2;
        p.avg_bet_size = 10.0 + (double)(rng() % 5000) / 7.0;
        p.bonus_claim_count = (int32_t)(rng() % 7);
        p.bonus_wager_complete = (double)(rng() % 100) / 100.0;
        players[i] = p;
        dir.Add(make_id(i));
    }
    std::vector<uint32_t> ea(n_events), et(n_events);
    std::vector<int64_t> em(n_events);
    std::vector<uint8_t> ey(n_events);
    for (int i = 0; i < n_events; ++i) {
        ea[i] = (uint32_t)(rng() % 40);                 // hot accounts: velocity rules fire
        // accounts 0..9 burst inside the last minute (HIGH_VELOCITY), the rest spread over the hour
        et[i] = (uint32_t)(now - 1 - (int64_t)(rng() % (ea[i] < 10 ? 60 : 3600)));
        em[i] = (int64_t)(100 + rng() % 100000);
        ey[i] = (uint8_t)(rng() % 3);
    }
    igs_event_cols ev{ea.data(), et.data(), em.data(), ey.data()};

    igs_config cfg = igs::DefaultConfig();               // model_kind = mock: the only model the reference can run
    igs::ScoringEngine engine(
        0, cfg, &dir, [](const std::string& ip) { return std::make_tuple(ip == "10.1.1.1", false, ip == "tor-exit"); },
        [](const std::string& dev, const std::string&, const std::string&) { return dev == "stolen-phone"; },
        [](const igs::AccountID& id) { return std::make_pair(id[12] % 7, id[12] % 9); });
    engine.LoadPlayers(players.data(), players.size());
    engine.LoadEvents(ev, n_events, n_players);

    std::vector<igs::ScoreRequest> reqs;
    const char* types[3] = {"bet", "deposit", "withdraw"};
    for (int i = 0; i < 3000; ++i) {
        igs::ScoreRequest r;
        r.account_id = make_id((uint32_t)(rng() % (n_players + 5)));     // a few unknown accounts
        r.amount = (int64_t)(rng() % 300000);
        r.tx_type = types[rng() % 3];
        r.ip = (rng() % 20 == 0) ? "10.1.1.1" : (rng() % 50 == 0) ? "tor-exit" : "8.8.8.8";
        r.device_id = (rng() % 100 == 0) ? "stolen-phone" : "phone";
        r.timestamp = now;
        reqs.push_back(r);
    }
    const std::vector<igs::ScoreResponse> res = engine.ScoreBatch(reqs);
    REQUIRE(res.size() == reqs.size());

    // the same batch through the oracle
    const std::vector<igs_tx> tx = engine.Marshal(reqs);
    std::vector<igs_result> want(tx.size());
    orc_score_batch(&cfg, nullptr, players.data(), players.size(), &ev, n_events, tx.data(), tx.size(), want.data(), nullptr,
                    nullptr);
    uint32_t seen = 0;
    for (size_t i = 0; i < res.size(); ++i) {
        REQUIRE(res[i].score == want[i].score);
        REQUIRE(res[i].action == igs_action_name(want[i].action));
        REQUIRE((float)res[i].ml_score == want[i].ml_score);
        REQUIRE(res[i].rule_score == igs_rule_score_from_flags(&cfg, want[i].flags));
        const char* names[IGS_NUM_REASONS];
        const int k = igs_reason_codes(want[i].flags, names);
        REQUIRE((int)res[i].reason_codes.size() == k);
        for (int j = 0; j < k; ++j) REQUIRE(res[i].reason_codes[j] == names[j]);
        seen |= want[i].flags;
    }
    REQUIRE(seen & IGS_R_HIGH_VELOCITY);
    REQUIRE(seen & IGS_R_VPN_DETECTED);
    REQUIRE(seen & IGS_R_KNOWN_FRAUDSTER);
    REQUIRE(seen & IGS_R_ML_HIGH_RISK);

    // Score == ScoreBatch of one; SetThresholds moves the action, not the score (engine.go:498-504)
    const igs::ScoreResponse one = engine.Score(reqs[0]);
    REQUIRE(one.score == res[0].score && one.action == res[0].action);
    REQUIRE(engine.GetThresholds() == std::make_pair(80, 50));
    engine.SetThresholds(101, 101);
    for (const auto& r : engine.ScoreBatch(reqs)) REQUIRE(r.action == "approve");
    engine.SetThresholds(0, 0);
    for (const auto& r : engine.ScoreBatch(reqs)) REQUIRE(r.action == "block");
    engine.SetThresholds(80, 50);
    REQUIRE(engine.ScoreWithExplanation(reqs[0]).find("Final Score:") != std::string::npos);

    // errors surface as exceptions with the library's message (no fallback)
    bool threw = false;
    try {
        igs_config g = cfg;
        g.model_kind = IGS_MODEL_GBDT;                   // no ensemble loaded
        igs::ScoringEngine e2(0, g, &dir);
        e2.ScoreBatch(reqs);
    } catch (const igs::Error& e) {
        threw = e.status == IGS_ERR_STATE;
    }
    REQUIRE(threw);

    // Identifiers on the device: the reference's real-time loop — Score, then UpdateFeatures — with the unique
    // device / IP counts coming from the event index and rule 8 from AddToBlacklist, checked against the oracle's
    // command-level Redis emulation of the same stream.
    {
        igs_config c0 = cfg;
        c0.model_kind = IGS_MODEL_NONE;
        igs::ScoringEngine e3(0, c0, &dir);
        e3.LoadPlayers(players.data(), players.size());
        const igs_event_cols none{nullptr, nullptr, nullptr, nullptr};
        const igs_event_ids no_ids{nullptr, nullptr};
        e3.LoadEvents(none, no_ids, 0, n_players);
        e3.AddToBlacklist("ip", "203.0.113.66");
        bool bad_type = false;
        try { e3.AddToBlacklist("email", "x"); } catch (const igs::Error& e) { bad_type = e.status == IGS_ERR_INVALID; }
        REQUIRE(bad_type);
        orc_redis_account* redis = orc_redis_new();
        uint32_t fired = 0;
        for (int step = 0; step < 8; ++step) {
            igs::ScoreRequest r;
            r.account_id = make_id(3);
            r.amount = 700; r.tx_type = "bet"; r.timestamp = now + 5 * step;
            r.device_id = "dev-" + std::to_string(step % 5);
            r.ip = step == 7 ? "203.0.113.66" : "198.51.100." + std::to_string(step);
            const igs::ScoreResponse got = e3.Score(r);
            orc_realtime rt;
            orc_redis_get(redis, r.timestamp, &rt);
            std::vector<std::string> want_codes;
            if (rt.unique_devices_24h > 3) want_codes.push_back("MULTIPLE_DEVICES");
            if (rt.unique_ips_24h > 5) want_codes.push_back("IP_COUNTRY_MISMATCH");
            if (step == 7) want_codes.push_back("KNOWN_FRAUDSTER");
            std::vector<std::string> got_codes;
            for (const std::string& code : got.reason_codes)
                if (code == "MULTIPLE_DEVICES" || code == "IP_COUNTRY_MISMATCH" || code == "KNOWN_FRAUDSTER") got_codes.push_back(code);
            REQUIRE(got_codes == want_codes);
            fired |= (uint32_t)want_codes.size();
            igs::ScoringEngine::TransactionEvent te;
            te.account_id = r.account_id; te.amount = r.amount; te.ip = r.ip; te.device_id = r.device_id; te.timestamp = r.timestamp;
            e3.UpdateFeatures({te});
            orc_redis_update_ids(redis, r.timestamp, r.amount, orc_hash_id(r.device_id.data(), r.device_id.size()),
                                 orc_hash_id(r.ip.data(), r.ip.size()));
        }
        orc_redis_free(redis);
        REQUIRE(fired == 3);      // some step fired one rule, the last one all three

        // RiskChecker.CheckBonusAbuse: accounts 3 and 7 share "dev-2" -> MULTI_ACCOUNT for both, checked against
        // the oracle's table over the same history
        igs::ScoringEngine::TransactionEvent te;
        te.account_id = make_id(7); te.amount = 5; te.device_id = "dev-2"; te.ip = "198.51.100.77"; te.timestamp = now + 100;
        e3.UpdateFeatures({te});
        e3.BuildLinks();
        std::vector<uint64_t> hd, ha;
        for (int step = 0; step < 8; ++step) {
            const std::string dv = "dev-" + std::to_string(step % 5);
            hd.push_back(orc_hash_id(dv.data(), dv.size())); ha.push_back(3);
        }
        hd.push_back(orc_hash_id("dev-2", 5)); ha.push_back(7);
        const uint64_t q[3] = {3, 7, 9};
        igs_links_row want_rows[3];
        orc_links(hd.data(), ha.data(), hd.size(), q, 3, want_rows);
        for (int k = 0; k < 3; ++k) {
            igs_abuse got{}, want_abuse{};
            const bool is_abuser = e3.CheckBonusAbuse(make_id((uint32_t)q[k]), &got);
            orc_check_bonus_abuse(&players[q[k]], &want_rows[k], &want_abuse);
            REQUIRE(got.n_linked == want_abuse.n_linked && got.signals == want_abuse.signals);
            REQUIRE(got.abuse_score == want_abuse.abuse_score && is_abuser == (want_abuse.is_abuser != 0));
        }
        REQUIRE(want_rows[0].n_linked == 1 && want_rows[0].linked[0] == 7 && want_rows[2].n_linked == 0);
    }
    {
        // ShardedEngine: three shards (engines) in this process, one host thread each, against one engine holding
        // everything — a result depends on its own account only, so the two must agree bit for bit
        const uint32_t np = 500, ne = 20000, nt = 30000;
        std::vector<igs_player_row> pl(np);
        for (uint32_t i = 0; i < np; ++i) {
            igs_player_row p{};
            p.account_created_at = now - (int64_t)(rng() % 600) * 86400;
            p.total_deposits = (int64_t)(rng() % 300000); p.total_withdrawals = (int64_t)(rng() % 300000);
            p.deposit_count = (int32_t)(rng() % 9); p.bet_count = (int32_t)(rng() % 500); p.win_count = p.bet_count / 3;
            p.bonus_claim_count = (int32_t)(rng() % 6);
            pl[i] = p;
        }
        std::vector<uint32_t> a(ne), t(ne); std::vector<int64_t> m(ne); std::vector<uint8_t> y(ne);
        for (uint32_t i = 0; i < ne; ++i) {
            a[i] = (uint32_t)(rng() % np); t[i] = (uint32_t)(now - 7200 + (int64_t)(rng() % 7200));
            m[i] = 100 * (int64_t)(1 + rng() % 50); y[i] = (uint8_t)(rng() % 3);
        }
        std::vector<igs_tx> tx(nt);
        for (uint32_t i = 0; i < nt; ++i) {
            igs_tx x{};
            x.account = (uint32_t)(rng() % (np + 3));               // a few unknown accounts
            x.ts = (uint32_t)(now - 600 + (int64_t)(rng() % 600)); x.seq_limit = 0xFFFFFFFFu;
            x.amount = (int64_t)(rng() % 400000); x.tx_type = (uint8_t)(rng() % 3);
            x.unique_devices_24h = (uint16_t)(rng() % 6); x.unique_ips_24h = (uint16_t)(rng() % 8);
            x.bits = IGS_BIT_VALID | (uint8_t)(rng() % 100 < 3 ? IGS_BIT_VPN : 0);
            tx[i] = x;
        }
        igs_config cfg = igs::DefaultConfig();                      // mock model: needs no loaded ensemble
        igs::ShardedEngine sharded(3, 1);
        sharded.LoadPlayers(pl.data(), np);
        sharded.LoadEvents(a.data(), t.data(), m.data(), y.data(), ne);
        std::vector<igs_result> got(nt), want(nt);
        sharded.ScoreBatch(cfg, tx.data(), nt, got.data());
        igs_engine* one = nullptr;
        REQUIRE(igs_engine_create(0, &one) == IGS_OK);
        REQUIRE(igs_load_players(one, pl.data(), np) == IGS_OK);
        const igs_event_cols cols{a.data(), t.data(), m.data(), y.data()};
        REQUIRE(igs_load_events(one, &cols, ne, np) == IGS_OK);
        for (igs_tx& x : tx) if (x.account >= np) x.account = 0xFFFFFFFEu;
        REQUIRE(igs_score_batch(one, &cfg, tx.data(), nt, want.data(), nullptr, nullptr) == IGS_OK);
        igs_engine_destroy(one);
        REQUIRE(std::memcmp(got.data(), want.data(), nt * sizeof(igs_result)) == 0);
        uint32_t used = 0;
        for (uint32_t g = 0; g < np; ++g) used |= 1u << sharded.ShardOf(g);
        REQUIRE(used == 7u);
    }
    std::printf("test_scoring_engine: %zu requests, flags seen 0x%x: ok\n", res.size(), seen);
    return 0;
}
