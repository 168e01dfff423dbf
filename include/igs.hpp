// igs.hpp — C++17 host side over the C ABI (igs.h), header only.
//
// The reference's host code is Go (services/risk); this image has no Go toolchain, so the compiled-language
// mirror of its operator interface is written in C++ (the cgo equivalent is sketched in INTEGRATION.md).
// Names, argument meaning and error behaviour follow services/risk/internal/scoring/engine.go:
//   ScoreRequest :40-53, ScoreResponse :56-64, NewScoringEngine :231-259, Score :262-323,
//   GetThresholds / SetThresholds :491-504, ScoreWithExplanation :507-543, DefaultConfig :215-228.
// Everything here is marshalling; every number in a response comes from the CUDA library.
#pragma once

#include <algorithm>
#include <array>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <map>
#include <stdexcept>
#include <string>
#include <thread>
#include <tuple>
#include <utility>
#include <vector>

#include "igs.h"

namespace igs {

struct Error : std::runtime_error {
    int status;
    Error(int st, const std::string& what) : std::runtime_error(what), status(st) {}
};

using AccountID = std::array<uint8_t, 16>;   // uuid.UUID

struct ScoreRequest {          // scoring.ScoreRequest
    AccountID account_id{};
    std::string player_id;
    int64_t amount = 0;
    std::string tx_type = "bet";   // "deposit" | "withdraw" | "bet"
    std::string currency, game_id, ip, device_id, fingerprint, user_agent, session_id;
    int64_t timestamp = 0;         // unix seconds; 0 = now
};

struct ScoreResponse {         // scoring.ScoreResponse
    int score = 0;
    std::string action;
    std::vector<std::string> reason_codes;
    int rule_score = 0;
    double ml_score = 0;
    int64_t response_time_ms = 0;
};

inline igs_config DefaultConfig() {
    igs_config c;
    igs_default_config(&c);
    return c;
}

// account uuid -> dense index of this shard (unknown accounts map beyond every table: the engine then
// scores them with zero batch / velocity features, as the reference does on a feature-store error)
class PlayerDirectory {
public:
    uint32_t Add(const AccountID& id) { return idx_.emplace(id, (uint32_t)idx_.size()).first->second; }
    uint32_t Get(const AccountID& id) const {
        auto it = idx_.find(id);
        return it == idx_.end() ? 0xFFFFFFFEu : it->second;
    }
    size_t size() const { return idx_.size(); }
    static uint32_t ShardOf(const AccountID& id, uint32_t n_shards) {
        uint64_t lo = 0, hi = 0;
        for (int i = 0; i < 8; ++i) { lo |= (uint64_t)id[15 - i] << (8 * i); hi |= (uint64_t)id[7 - i] << (8 * i); }
        return igs_shard_of(lo, hi, n_shards);
    }
private:
    std::map<AccountID, uint32_t> idx_;
};

using IPIntelligence = std::function<std::tuple<bool, bool, bool>(const std::string& ip)>;           // vpn, proxy, tor
using Blacklist = std::function<bool(const std::string& device, const std::string& fp, const std::string& ip)>;
using Uniques = std::function<std::pair<int, int>(const AccountID&)>;                                // devices, ips (24h)

template <typename T>
class Pinned {                  // host staging from igs_alloc_pinned
public:
    ~Pinned() { igs_free_pinned(p_); }
    T* reserve(size_t n) {
        if (n > cap_) {
            igs_free_pinned(p_);
            p_ = static_cast<T*>(igs_alloc_pinned((n ? n : 1) * sizeof(T)));
            if (!p_) throw Error(IGS_ERR_NOMEM, "igs_alloc_pinned failed");
            cap_ = n;
        }
        return p_;
    }
private:
    T* p_ = nullptr;
    size_t cap_ = 0;
};

class ScoringEngine {
public:
    // NewScoringEngine(features, model, ipIntel, blacklist, logger, config): the feature store and the
    // model live on the device (Load*), ipIntel / blacklist / uniques may be empty like the reference's nil.
    ScoringEngine(int device, const igs_config& config, PlayerDirectory* directory, IPIntelligence ip_intel = nullptr,
                  Blacklist blacklist = nullptr, Uniques uniques = nullptr)
        : cfg_(config), dir_(directory), ip_(std::move(ip_intel)), bl_(std::move(blacklist)), uq_(std::move(uniques)) {
        const int rc = igs_engine_create(device, &e_);
        if (rc != IGS_OK) throw Error(rc, igs_last_error(nullptr));
    }
    ~ScoringEngine() { igs_engine_destroy(e_); }
    ScoringEngine(const ScoringEngine&) = delete;
    ScoringEngine& operator=(const ScoringEngine&) = delete;

    igs_engine* handle() { return e_; }
    void LoadGBDT(const igs_gbdt_desc& d) { check(igs_load_gbdt(e_, &d)); }
    void LoadPlayers(const igs_player_row* rows, uint64_t n) { check(igs_load_players(e_, rows, n)); }
    void LoadEvents(const igs_event_cols& ev, uint64_t n, uint32_t n_accounts) { check(igs_load_events(e_, &ev, n, n_accounts)); }
    // History with device / IP identifier columns: from then on UniqueDevices24h / UniqueIPs24h come from the
    // event index, and ScoreBatch evaluates rule 8 against the device-side blacklist sets.
    void LoadEvents(const igs_event_cols& ev, const igs_event_ids& ids, uint64_t n, uint32_t n_accounts) {
        check(igs_load_events_ids(e_, &ev, &ids, n, n_accounts));
        device_ids_ = true;
    }
    static uint64_t HashId(const std::string& s) { return igs_hash_id(s.data(), s.size()); }
    // RedisFeatureStore.AddToBlacklist, features/redis_store.go:251-264
    void AddToBlacklist(const std::string& list_type, const std::string& value) {
        const int kind = list_type == "device" ? IGS_BL_DEVICE : list_type == "ip" ? IGS_BL_IP
                       : list_type == "fingerprint" ? IGS_BL_FINGERPRINT : -1;
        if (kind < 0) throw Error(IGS_ERR_INVALID, "unknown blacklist type: " + list_type);   // :261
        const uint64_t id = HashId(value);
        check(igs_blacklist_add(e_, kind, &id, 1));
        device_ids_ = true;
    }
    // Rebuilds the linked-accounts table of a single-shard deployment from the loaded history (a periodic job;
    // with several shards the three igs_links_* stages are driven with an all-to-all between them).
    void BuildLinks() {
        const igs_link_pair* pairs = nullptr; const igs_link* links = nullptr;
        uint64_t n_pairs = 0, n_links = 0;
        check(igs_links_pairs(e_, 0, 1, &pairs, &n_pairs));
        check(igs_links_group(e_, pairs, n_pairs, 1, &links, &n_links));
        check(igs_links_finish(e_, links, n_links, 0));
    }
    // RiskChecker.CheckBonusAbuse, services/bonus/internal/service/bonus_engine.go:139-141
    bool CheckBonusAbuse(const AccountID& account_id, igs_abuse* detail = nullptr) {
        const uint32_t idx = dir_ ? dir_->Get(account_id) : 0xFFFFFFFEu;
        igs_abuse a{};
        check(igs_check_bonus_abuse(e_, &idx, 1, &a));
        if (detail) *detail = a;
        return a.is_abuser != 0;
    }
    // ScoringEngine.UpdateFeatures (engine.go:486-488) for a micro-batch of completed transactions.
    struct TransactionEvent {      // features.TransactionEvent, features/redis_store.go:49-56
        std::array<uint8_t, 16> account_id{};
        int64_t amount = 0;
        std::string tx_type = "bet", ip, device_id;
        int64_t timestamp = 0;
    };
    void UpdateFeatures(const std::vector<TransactionEvent>& events) {
        std::vector<uint32_t> acct, ts; std::vector<int64_t> amt; std::vector<uint8_t> ty; std::vector<uint64_t> dev, ip;
        for (const TransactionEvent& ev : events) {
            const uint32_t a = dir_ ? dir_->Get(ev.account_id) : 0xFFFFFFFEu;
            if (a == 0xFFFFFFFEu) continue;              // unknown account: the device tables are dense
            acct.push_back(a); ts.push_back((uint32_t)ev.timestamp); amt.push_back(ev.amount);
            ty.push_back(TxType(ev.tx_type)); dev.push_back(HashId(ev.device_id)); ip.push_back(HashId(ev.ip));
        }
        const igs_event_cols cols{acct.data(), ts.data(), amt.data(), ty.data()};
        const igs_event_ids ids{dev.data(), ip.data()};
        check(device_ids_ ? igs_append_events_ids(e_, &cols, &ids, acct.size()) : igs_append_events(e_, &cols, acct.size()));
    }

    std::vector<igs_tx> Marshal(const std::vector<ScoreRequest>& reqs) const {
        const int64_t now = std::chrono::duration_cast<std::chrono::seconds>(
                                std::chrono::system_clock::now().time_since_epoch()).count();
        std::vector<igs_tx> tx(reqs.size());
        for (size_t i = 0; i < reqs.size(); ++i) {
            const ScoreRequest& r = reqs[i];
            igs_tx t{};
            t.amount = r.amount;
            t.account = dir_ ? dir_->Get(r.account_id) : 0xFFFFFFFEu;
            t.ts = (uint32_t)(r.timestamp ? r.timestamp : now);
            t.seq_limit = 0xFFFFFFFFu;
            t.tx_type = TxType(r.tx_type);
            uint8_t bits = IGS_BIT_VALID;
            if (ip_ && !r.ip.empty()) {                                   // engine.go:394-405
                auto [vpn, proxy, tor] = ip_(r.ip);
                bits |= (vpn ? IGS_BIT_VPN : 0) | (proxy ? IGS_BIT_PROXY : 0) | (tor ? IGS_BIT_TOR : 0);
            }
            if (bl_ && bl_(r.device_id, r.fingerprint, r.ip)) bits |= IGS_BIT_BLACKLISTED;   // :469-475
            t.bits = bits;
            if (uq_) {
                auto [d, p] = uq_(r.account_id);
                t.unique_devices_24h = (uint16_t)(d > 0xFFFF ? 0xFFFF : d);
                t.unique_ips_24h = (uint16_t)(p > 0xFFFF ? 0xFFFF : p);
            }
            tx[i] = t;
        }
        return tx;
    }

    // rpc ScoreBatch (risk.proto:13): one C-ABI call per batch.
    std::vector<ScoreResponse> ScoreBatch(const std::vector<ScoreRequest>& reqs) {
        const auto t0 = std::chrono::steady_clock::now();
        const std::vector<igs_tx> tx = Marshal(reqs);
        const size_t n = tx.size();
        igs_tx* h_tx = tx_.reserve(n);
        igs_result* h_out = out_.reserve(n);
        std::copy(tx.begin(), tx.end(), h_tx);
        if (device_ids_) {                  // rule 8 on the device: identifiers travel as 64-bit ids
            igs_tx_ids* h_ids = ids_.reserve(n);
            for (size_t i = 0; i < n; ++i)
                h_ids[i] = igs_tx_ids{HashId(reqs[i].device_id), HashId(reqs[i].fingerprint), HashId(reqs[i].ip)};
            check(igs_score_batch_ids(e_, &cfg_, h_tx, h_ids, n, h_out, nullptr, nullptr));
        } else {
            check(igs_score_batch(e_, &cfg_, h_tx, n, h_out, nullptr, nullptr));
        }
        const int64_t ms = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
        std::vector<ScoreResponse> res(n);
        const char* names[IGS_NUM_REASONS];
        for (size_t i = 0; i < n; ++i) {
            ScoreResponse& r = res[i];
            r.score = h_out[i].score;
            r.action = igs_action_name(h_out[i].action);
            const int k = igs_reason_codes(h_out[i].flags, names);
            r.reason_codes.assign(names, names + k);
            r.rule_score = igs_rule_score_from_flags(&cfg_, h_out[i].flags);
            r.ml_score = h_out[i].ml_score;
            r.response_time_ms = ms;
        }
        return res;
    }
    ScoreResponse Score(const ScoreRequest& req) { return ScoreBatch({req})[0]; }

    std::pair<int, int> GetThresholds() const { return {cfg_.block_threshold, cfg_.review_threshold}; }
    void SetThresholds(int block, int review) { cfg_.block_threshold = block; cfg_.review_threshold = review; }

    std::string ScoreWithExplanation(const ScoreRequest& req) {
        const ScoreResponse r = Score(req);
        char buf[256];
        std::snprintf(buf, sizeof(buf), "\nFraud Score Analysis\n====================\nFinal Score: %d/100\nAction: %s\n"
                      "Response Time: %lldms\n\nRule Contribution: %d\nML Contribution: %.2f (%.2f * 100)\n\nTriggered Rules:\n",
                      r.score, r.action.c_str(), (long long)r.response_time_ms, r.rule_score, r.ml_score, r.ml_score);
        std::string s = buf;
        static const char* order[IGS_NUM_REASONS] = {"HIGH_VELOCITY", "NEW_ACCOUNT_LARGE_TX", "IP_COUNTRY_MISMATCH",
            "MULTIPLE_DEVICES", "SUSPICIOUS_PATTERN", "VPN_DETECTED", "KNOWN_FRAUDSTER", "RAPID_DEPOSIT_WITHDRAW",
            "BONUS_ABUSE", "ML_HIGH_RISK"};
        for (const std::string& code : r.reason_codes)
            for (int b = 0; b < IGS_NUM_REASONS; ++b)
                if (code == order[b]) s += "  - " + code + " (+" + std::to_string(cfg_.reason_weight[b]) + ")\n";
        return s;
    }

private:
    static uint8_t TxType(const std::string& t) {
        return t == "bet" ? IGS_TX_BET : t == "deposit" ? IGS_TX_DEPOSIT : t == "withdraw" ? IGS_TX_WITHDRAW : IGS_TX_OTHER;
    }
    void check(int rc) const {
        if (rc != IGS_OK) throw Error(rc, igs_last_error(e_));
    }
    igs_engine* e_ = nullptr;
    igs_config cfg_;
    PlayerDirectory* dir_;
    IPIntelligence ip_;
    Blacklist bl_;
    Uniques uq_;
    Pinned<igs_tx> tx_;
    Pinned<igs_tx_ids> ids_;
    Pinned<igs_result> out_;
    bool device_ids_ = false;
};

// ShardedEngine — the in-process shape of the risk service on a multi-GPU box (services/risk/cmd/main.go:72-164 is
// ONE process): one igs_engine, one host thread and one set of pinned staging buffers per shard, shards placed
// round-robin on the visible devices, records routed by igs_shard_of(account).  Every feature, rule and score
// depends on one account only, so the shards never talk to each other; a batch is split, the shards
// score their parts concurrently, the results land back in the caller's order.  Accounts are GLOBAL dense indices
// here; the engine of a shard sees the local index (position among the shard's accounts in increasing global order).
class ShardedEngine {
public:
    explicit ShardedEngine(int n_shards, int n_devices = 1) : shards_(n_shards) {
        if (n_shards < 1 || n_shards > 256 || n_devices < 1) throw Error(IGS_ERR_INVALID, "1 <= n_shards <= 256 and n_devices >= 1");
        for (int s = 0; s < n_shards; ++s)
            if (int rc = igs_engine_create(s % n_devices, &shards_[s].e); rc != IGS_OK) {
                const std::string msg = igs_last_error(nullptr);
                for (int k = 0; k < s; ++k) igs_engine_destroy(shards_[k].e);
                throw Error(rc, msg);
            }
    }
    ~ShardedEngine() { for (Shard& sh : shards_) igs_engine_destroy(sh.e); }
    ShardedEngine(const ShardedEngine&) = delete;
    ShardedEngine& operator=(const ShardedEngine&) = delete;

    int n_shards() const { return (int)shards_.size(); }
    igs_engine* shard(int s) { return shards_[s].e; }
    uint32_t ShardOf(uint32_t account) const { return igs_shard_of(account, 0, (uint32_t)shards_.size()); }

    void LoadGbdt(const igs_gbdt_desc& d) { for (Shard& sh : shards_) check(sh, igs_load_gbdt(sh.e, &d)); }

    // FeatureStore.GetBatchFeatures for every shard: rows[g] = BatchFeatures of global account g
    void LoadPlayers(const igs_player_row* rows, uint32_t n) {
        local_.assign(n, 0);
        std::vector<std::vector<igs_player_row>> part(shards_.size());
        for (uint32_t g = 0; g < n; ++g) {
            auto& p = part[ShardOf(g)];
            local_[g] = (uint32_t)p.size();
            p.push_back(rows[g]);
        }
        for (size_t s = 0; s < shards_.size(); ++s) {
            shards_[s].n_accounts = (uint32_t)part[s].size();
            check(shards_[s], igs_load_players(shards_[s].e, part[s].data(), part[s].size()));
        }
    }
    // the event history, split by account shard (arrival order is kept inside a shard); needs LoadPlayers first
    void LoadEvents(const uint32_t* account, const uint32_t* ts, const int64_t* amount, const uint8_t* type, uint64_t n) {
        struct Cols { std::vector<uint32_t> a, t; std::vector<int64_t> m; std::vector<uint8_t> y; };
        std::vector<Cols> part(shards_.size());
        for (uint64_t i = 0; i < n; ++i) {
            if (account[i] >= local_.size()) throw Error(IGS_ERR_INVALID, "event of an account beyond the player table");
            Cols& c = part[ShardOf(account[i])];
            c.a.push_back(local_[account[i]]); c.t.push_back(ts[i]); c.m.push_back(amount[i]); c.y.push_back(type[i]);
        }
        for (size_t s = 0; s < shards_.size(); ++s) {
            const igs_event_cols cols{part[s].a.data(), part[s].t.data(), part[s].m.data(), part[s].y.data()};
            check(shards_[s], igs_load_events(shards_[s].e, &cols, part[s].a.size(), shards_[s].n_accounts));
        }
    }

    // rpc ScoreBatch over all shards: tx[i].account is a GLOBAL index; out[i] belongs to tx[i].
    // Host side = a parallel stable partition: `route_threads` threads count their slice of the batch per shard, the
    // counts become offsets, the same threads copy their records (local account index filled in) straight into the
    // shards' pinned staging; then one host thread per shard (Go: LockOSThread per device) calls igs_score_batch, and
    // the routing threads scatter the results back.  Records keep their arrival order inside a shard.
    void ScoreBatch(const igs_config& cfg, const igs_tx* tx, uint64_t n, igs_result* out) {
        const size_t S = shards_.size();
        const size_t P = std::max<size_t>(1, std::min<size_t>(route_threads_, n / 65536 + 1));
        shard_id_.resize(n);
        std::vector<std::vector<uint64_t>> off(P, std::vector<uint64_t>(S, 0));
        auto slice = [&](size_t p) { return std::pair<uint64_t, uint64_t>(n * p / P, n * (p + 1) / P); };
        parallel(P, [&](size_t p) {
            auto [lo, hi] = slice(p);
            std::vector<uint64_t>& c = off[p];
            for (uint64_t i = lo; i < hi; ++i) {
                const uint8_t sh = (uint8_t)ShardOf(tx[i].account);
                shard_id_[i] = sh;
                ++c[sh];
            }
        });
        std::vector<uint64_t> m(S, 0);
        for (size_t sh = 0; sh < S; ++sh)
            for (size_t p = 0; p < P; ++p) { const uint64_t c = off[p][sh]; off[p][sh] = m[sh]; m[sh] += c; }
        std::vector<igs_tx*> in(S);
        std::vector<igs_result*> res(S);
        for (size_t sh = 0; sh < S; ++sh) {
            shards_[sh].rows.resize(m[sh]);
            in[sh] = shards_[sh].tx.reserve(std::max<uint64_t>(m[sh], 1));
            res[sh] = shards_[sh].out.reserve(std::max<uint64_t>(m[sh], 1));
        }
        parallel(P, [&](size_t p) {
            auto [lo, hi] = slice(p);
            std::vector<uint64_t> k = off[p];
            for (uint64_t i = lo; i < hi; ++i) {
                const size_t sh = shard_id_[i];
                const uint64_t at = k[sh]++;
                igs_tx x = tx[i];
                x.account = x.account < local_.size() ? local_[x.account] : 0xFFFFFFFEu;   // unknown account: zero features
                in[sh][at] = x;
                shards_[sh].rows[at] = i;
            }
        });
        parallel(S, [&](size_t sh) {
            Shard& s = shards_[sh];
            s.rc = m[sh] ? igs_score_batch(s.e, &cfg, in[sh], m[sh], res[sh], nullptr, nullptr) : IGS_OK;
        });
        for (Shard& sh : shards_) check(sh, sh.rc);
        parallel(P, [&](size_t p) {
            for (size_t sh = 0; sh < S; ++sh) {
                const uint64_t lo = m[sh] * p / P, hi = m[sh] * (p + 1) / P;
                const uint64_t* rows = shards_[sh].rows.data();
                for (uint64_t k = lo; k < hi; ++k) out[rows[k]] = res[sh][k];
            }
        });
    }
    // host threads that route a batch (default: the hardware threads, at most 16)
    void set_route_threads(unsigned t) { route_threads_ = t ? t : 1; }

private:
    struct Shard {
        igs_engine* e = nullptr;
        uint32_t n_accounts = 0;
        std::vector<uint64_t> rows;
        Pinned<igs_tx> tx;
        Pinned<igs_result> out;
        int rc = IGS_OK;
    };
    static void check(const Shard& sh, int rc) {
        if (rc != IGS_OK) throw Error(rc, igs_last_error(sh.e));
    }
    template <typename F>
    static void parallel(size_t n_threads, F&& body) {               // body(0 .. n_threads - 1), one std::thread each
        if (n_threads <= 1) { body(0); return; }
        std::vector<std::thread> w;
        w.reserve(n_threads);
        for (size_t t = 0; t < n_threads; ++t) w.emplace_back([&body, t] { body(t); });
        for (std::thread& x : w) x.join();
    }
    std::vector<Shard> shards_;
    std::vector<uint32_t> local_;      // global account -> local index on its shard
    std::vector<uint8_t> shard_id_;    // per record of the batch in flight
    unsigned route_threads_ = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
};

}  // namespace igs
