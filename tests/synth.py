"""Seeded synthetic inputs: one splitmix64 stream per column, master seed 20251106.

Shared by the tests, __graft_entry__.smoke() and bench.py so that the GPU path, the oracle and the
CPU baseline all see identical bytes.  Pure numpy; no reference code is involved.
"""
from __future__ import annotations

import os
import sys

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
import __graft_entry__ as _ge  # noqa: E402

pkg = _ge.load_package()
abi = pkg.abi

MASTER_SEED = 20251106
T0 = 1_700_000_000          # base unix time of the synthetic clock
_GOLDEN = np.uint64(0x9E3779B97F4A7C15)


def splitmix64(seed: int, stream: int, n: int) -> np.ndarray:
    """n outputs of splitmix64 started at hash(seed, stream) — vectorised, uint64."""
    with np.errstate(over="ignore"):
        s0 = np.uint64((seed * 0x9E3779B97F4A7C15 + stream * 0xBF58476D1CE4E5B9 + 0x94D049BB133111EB)
                       & 0xFFFFFFFFFFFFFFFF)
        z = s0 + _GOLDEN * (np.arange(1, n + 1, dtype=np.uint64))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def uniform(seed: int, stream: int, n: int) -> np.ndarray:
    """U[0,1) float64."""
    return (splitmix64(seed, stream, n) >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))


def normal(seed: int, stream: int, n: int) -> np.ndarray:
    u1 = uniform(seed, stream, n)
    u2 = uniform(seed, stream + 1000003, n)
    return np.sqrt(-2.0 * np.log(1.0 - u1)) * np.cos(2.0 * np.pi * u2)


def geometric(u: np.ndarray, mean: float) -> np.ndarray:
    p = 1.0 / (mean + 1.0)
    return np.floor(np.log(1.0 - u) / np.log(1.0 - p)).astype(np.int64)


def poisson(u: np.ndarray, lam: float, kmax: int = 24) -> np.ndarray:
    pmf = [np.exp(-lam)]
    for k in range(1, kmax):
        pmf.append(pmf[-1] * lam / k)
    return np.searchsorted(np.cumsum(pmf), u).astype(np.int64)


def zipf_accounts(u: np.ndarray, n_accounts: int, s: float) -> np.ndarray:
    w = 1.0 / np.power(np.arange(1, n_accounts + 1, dtype=np.float64), s)
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    return np.minimum(np.searchsorted(cdf, u), n_accounts - 1).astype(np.uint32)


def tx_types(u: np.ndarray) -> np.ndarray:
    """70 % bet / 20 % deposit / 10 % withdraw."""
    t = np.full(u.shape, abi.TX_BET, dtype=np.uint8)
    t[u >= 0.7] = abi.TX_DEPOSIT
    t[u >= 0.9] = abi.TX_WITHDRAW
    return t


def amounts(u: np.ndarray) -> np.ndarray:
    """log-uniform 100 .. 5 000 000 cents."""
    return np.floor(100.0 * np.power(50000.0, u)).astype(np.int64)


# ------------------------------------------------------------------------------------------------
# models
# ------------------------------------------------------------------------------------------------
def make_gbdt(n_trees: int = 500, seed: int = MASTER_SEED, n_features: int = abi.NUM_FEATURES,
              post_transform: int = abi.POST_LOGISTIC, base_score: float = 0.0):
    """C2 model: complete depth-8 trees, feature U{0..n_features-1}, thr U(0,1), leaves N(0,0.05)."""
    nn = n_trees * abi.TREE_NODES
    feat = (splitmix64(seed, 1, nn) % np.uint64(n_features)).astype(np.uint8).reshape(n_trees, abi.TREE_NODES)
    thr = uniform(seed, 2, nn).astype(np.float32).reshape(n_trees, abi.TREE_NODES)
    leaf = (0.05 * normal(seed, 3, n_trees * abi.TREE_LEAVES)).astype(np.float32).reshape(
        n_trees, abi.TREE_LEAVES)
    return pkg.GbdtModel(feat, thr, leaf, base_score=base_score, post_transform=post_transform,
                         n_features=n_features)


def make_mlp(seed: int = MASTER_SEED):
    """C4 weights: N(0, 2/fan_in) rounded to bf16, bias 0.01."""
    def w(stream, n_out, n_in):
        a = (normal(seed, stream, n_out * n_in) * np.sqrt(2.0 / n_in)).astype(np.float32)
        return f32_to_bf16(a).reshape(n_out, n_in)

    def b(n):
        return np.full(n, 0.01, dtype=np.float32)

    H, I = abi.MLP_HIDDEN, abi.MLP_IN
    return pkg.MlpModel(w(11, H, I), b(H), w(12, H, H), b(H), w(13, H, H), b(H), w(14, 1, H), b(1))


def f32_to_bf16(a: np.ndarray) -> np.ndarray:
    """round-to-nearest-even bf16 bit patterns of a float32 array (finite inputs)."""
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32).astype(np.uint64)
    r = np.uint64(0x7FFF) + ((u >> np.uint64(16)) & np.uint64(1))
    return ((u + r) >> np.uint64(16)).astype(np.uint16)


def bf16_to_f32(b: np.ndarray) -> np.ndarray:
    return (b.astype(np.uint32) << np.uint32(16)).view(np.float32)


# ------------------------------------------------------------------------------------------------
# path A: pre-assembled feature rows (C2)
# ------------------------------------------------------------------------------------------------
def _min_max(x: np.ndarray, hi: float) -> np.ndarray:
    x = x.astype(np.float32)
    out = x / np.float32(hi)
    out[x < 0] = 0
    out[x > np.float32(hi)] = 1
    return out.astype(np.float32)


def make_feature_batch(n: int, seed: int = MASTER_SEED, invalid_rate: float = 0.0):
    """n x 64 normalised fp32 features + n rule-sideband rows.

    idx 0..29 follow ml.FeatureVector order with the reference's normalisation applied to
    C1-style raw fields (so rules and mock model see consistent values); idx 30..63 are U(0,1).
    """
    s = seed
    U = lambda k: uniform(s, k, n)  # noqa: E731
    c1m = geometric(U(10), 3.0)
    c5m = c1m + geometric(U(11), 4.0)
    c1h = c5m + geometric(U(12), 12.0)
    amt = amounts(U(13))
    sum1h = (c1h * (amounts(U(14)) // 8)).astype(np.int64)
    ttype = tx_types(U(15))
    age = geometric(U(16), 200.0)
    dep = np.floor(np.exp(8.0 + 2.0 * normal(s, 17, n))).astype(np.int64)
    dep[U(18) < 0.15] = 0
    wd = np.floor(dep * np.minimum(U(19) * 1.3, 1.25)).astype(np.int64)
    dep_cnt = np.where(dep > 0, 1 + geometric(U(20), 6.0), 0)
    wd_cnt = np.where(wd > 0, 1 + geometric(U(21), 2.0), 0)
    tslt = np.floor(np.exp(U(22) * 12.0)).astype(np.int64)          # 1 s .. ~45 h
    tslt[U(23) < 0.1] = 0                                             # no history
    sess = np.minimum(tslt, 1800) * (U(24) < 0.6)
    avg_bet = np.exp(5.0 + 1.5 * normal(s, 25, n))
    win_rate = np.clip(0.45 + 0.1 * normal(s, 26, n), 0, 1)
    claims = poisson(U(27), 1.5)
    wager = U(28)
    bonus_only = (claims > 3) & (dep < 5000)
    udev = 1 + geometric(U(29), 0.9)
    uip = 1 + geometric(U(30), 1.6)
    vpn = U(31) < 0.02
    proxy = U(32) < 0.01
    tor = U(33) < 0.005
    black = U(34) < 0.001
    valid = U(35) >= invalid_rate

    rows = np.zeros(n, dtype=abi.rule_row_dtype)
    rows["amount"] = amt
    rows["total_deposits"] = dep
    rows["total_withdrawals"] = wd
    rows["tx_count_1m"] = c1m
    rows["account_age_days"] = age
    rows["time_since_last_tx"] = tslt
    rows["deposit_count"] = dep_cnt
    rows["unique_devices_24h"] = udev
    rows["unique_ips_24h"] = uip
    rows["tx_type"] = ttype
    rows["bits"] = (vpn * abi.BIT_VPN + proxy * abi.BIT_PROXY + tor * abi.BIT_TOR
                    + bonus_only * abi.BIT_BONUS_ONLY + black * abi.BIT_BLACKLISTED
                    + valid * abi.BIT_VALID).astype(np.uint8)

    x = uniform(s, 40, n * abi.NUM_FEATURES).astype(np.float32).reshape(n, abi.NUM_FEATURES)
    f32 = lambda a: np.asarray(a).astype(np.float32)  # noqa: E731
    pos = lambda a: np.where(f32(a) > 0, f32(a), np.float32(0))  # noqa: E731  logTransform (identity)
    x[:, 0] = _min_max(c1m, 20)
    x[:, 1] = _min_max(c5m, 50)
    x[:, 2] = _min_max(c1h, 200)
    x[:, 3] = pos(sum1h)
    x[:, 4] = f32(np.where(c1h > 0, sum1h / np.maximum(c1h, 1), 0.0))
    x[:, 5] = _min_max(udev, 10)
    x[:, 6] = _min_max(uip, 20)
    x[:, 7] = 0
    x[:, 8] = 0
    x[:, 9] = _min_max(age, 365)
    x[:, 10] = pos(dep)
    x[:, 11] = pos(wd)
    x[:, 12] = f32(dep - wd)
    x[:, 13] = f32(dep_cnt)
    x[:, 14] = f32(wd_cnt)
    x[:, 15] = _min_max(tslt, 86400)
    x[:, 16] = f32(sess)
    x[:, 17] = f32(avg_bet)
    x[:, 18] = f32(win_rate)
    x[:, 19] = f32(vpn)
    x[:, 20] = f32(proxy)
    x[:, 21] = f32(tor)
    x[:, 22] = 0
    x[:, 23] = f32(claims)
    x[:, 24] = f32(wager)
    x[:, 25] = f32(bonus_only)
    x[:, 26] = pos(amt)
    x[:, 27] = f32(ttype == abi.TX_DEPOSIT)
    x[:, 28] = f32(ttype == abi.TX_WITHDRAW)
    x[:, 29] = f32(ttype == abi.TX_BET)
    return np.ascontiguousarray(x), rows


def make_unit_features(n: int, seed: int = MASTER_SEED) -> np.ndarray:
    """n x 64 U(0,1) fp32 — the plain C2 matrix for tree-walk-only checks."""
    return uniform(seed, 41, n * abi.NUM_FEATURES).astype(np.float32).reshape(n, abi.NUM_FEATURES)


# ------------------------------------------------------------------------------------------------
# path B: players, event history, compact tx (C1 / C3 / C5)
# ------------------------------------------------------------------------------------------------
def make_players(n_players: int, seed: int = MASTER_SEED, now: int = T0 + 48 * 3600) -> np.ndarray:
    s = seed + 101
    U = lambda k: uniform(s, k, n_players)  # noqa: E731
    p = np.zeros(n_players, dtype=abi.player_row_dtype)
    age_days = geometric(U(1), 200.0)
    p["account_created_at"] = now - age_days * 86400 - (U(2) * 86400).astype(np.int64)
    dep = np.floor(np.exp(8.0 + 2.0 * normal(s, 3, n_players))).astype(np.int64)
    dep[U(4) < 0.15] = 0
    p["total_deposits"] = dep
    p["total_withdrawals"] = np.floor(dep * np.minimum(U(5) * 1.3, 1.25)).astype(np.int64)
    p["deposit_count"] = np.where(dep > 0, 1 + geometric(U(6), 6.0), 0)
    p["withdraw_count"] = np.where(p["total_withdrawals"] > 0, 1 + geometric(U(7), 2.0), 0)
    p["bet_count"] = geometric(U(8), 300.0)
    p["win_count"] = np.floor(p["bet_count"] * np.clip(0.45 + 0.1 * normal(s, 9, n_players), 0, 1))
    p["total_bets"] = p["bet_count"] * np.floor(np.exp(5.0 + 1.5 * normal(s, 10, n_players))).astype(np.int64)
    p["total_wins"] = np.floor(p["total_bets"] * 0.95 * U(11) * 2).astype(np.int64)
    p["avg_bet_size"] = np.where(p["bet_count"] > 0, p["total_bets"] / np.maximum(p["bet_count"], 1), 0.0)
    p["bonus_claim_count"] = poisson(U(12), 1.5)
    p["bonus_wager_complete"] = U(13)
    return p


def make_events(n_events: int, n_accounts: int, seed: int = MASTER_SEED, zipf_s: float = 1.05,
                span: int = 48 * 3600, collision_rate: float = 0.01):
    """C3 history: Zipf accounts, ts over `span` seconds with per-account bursts, forced
    (ts, amount) collisions to exercise the ZSET member dedupe."""
    s = seed + 202
    U = lambda k: uniform(s, k, n_events)  # noqa: E731
    acct = zipf_accounts(U(1), n_accounts, zipf_s)
    ts = (U(2) * span).astype(np.int64)
    # bursts: half of the events land within 90 s of one of the account's 4 burst centres
    burst = U(3) < 0.5
    k = (splitmix64(s, 4, n_events) % np.uint64(4)).astype(np.uint64)
    with np.errstate(over="ignore"):
        h = (acct.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15) + k * np.uint64(0xD1B54A32D192ED03))
        h ^= h >> np.uint64(29)
        centre = (h % np.uint64(max(span - 120, 1))).astype(np.int64)
    ts = np.where(burst, centre + (U(5) * 90).astype(np.int64), ts)
    amt = amounts(U(6))
    ty = tx_types(U(7))
    # forced collisions: copy (account, ts, amount) of the previous event
    col = np.nonzero(U(8) < collision_rate)[0]
    col = col[col > 0]
    acct[col] = acct[col - 1]
    ts[col] = ts[col - 1]
    amt[col] = amt[col - 1]
    return pkg.Events(acct.astype(np.uint32), (T0 + ts).astype(np.uint32), amt.astype(np.int64), ty)


def mix64(x: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser, vectorised (the mix igs_hash_id ends with)."""
    x = x.astype(np.uint64).copy()
    with np.errstate(over="ignore"):
        x ^= x >> np.uint64(30); x *= np.uint64(0xbf58476d1ce4e5b9)
        x ^= x >> np.uint64(27); x *= np.uint64(0x94d049bb133111eb)
        x ^= x >> np.uint64(31)
    return x


def colliding_ids(count: int = 4, seed: int = 1) -> np.ndarray:
    """`count` distinct non-zero 64-bit ids whose mix64 values share their top 32 bits pairwise with a
    partner (birthday search) — they land in the same radix-sort key of the index build and must still
    be counted as different identifiers."""
    cand = splitmix64(seed, 99, 400_000) | np.uint64(1)
    key = mix64(cand) >> np.uint64(32)
    order = np.argsort(key, kind="stable")
    k = key[order]
    dup = np.nonzero(k[1:] == k[:-1])[0]
    out = []
    for d in dup:
        a, b = cand[order[d]], cand[order[d + 1]]
        if a != b:
            out += [a, b]
        if len(out) >= count:
            break
    assert len(out) >= count, "no hash collisions found"
    return np.array(out[:count], dtype=np.uint64)


def make_event_ids(ev, seed: int = MASTER_SEED, empty_rate: float = 0.1, shared_rate: float = 0.05,
                   devices_per_account: float = 1.2, ips_per_account: float = 2.5):
    """Device / IP ids for an event history: each account draws from its own small pools (so R3 / R4
    thresholds are straddled), some events carry no identifier, some devices are shared across accounts."""
    n = len(ev)
    s = seed + 404
    U = lambda k: uniform(s, k, n)  # noqa: E731
    acct = ev.account.astype(np.uint64)
    kd = geometric(U(1), devices_per_account).astype(np.uint64)
    ki = geometric(U(2), ips_per_account).astype(np.uint64)
    dev = mix64(acct * np.uint64(1_000_003) + kd + np.uint64(17)) | np.uint64(1)
    ip = mix64(acct * np.uint64(7_000_003) + ki + np.uint64(0xABCDEF)) | np.uint64(1)
    shared = U(3) < shared_rate
    dev = np.where(shared, mix64((splitmix64(s, 4, n) % np.uint64(50)) + np.uint64(5)) | np.uint64(1), dev)
    dev = np.where(U(5) < empty_rate, np.uint64(0), dev)
    ip = np.where(U(6) < empty_rate, np.uint64(0), ip)
    return pkg.Events(ev.account, ev.ts, ev.amount, ev.tx_type, device=dev.astype(np.uint64), ip=ip.astype(np.uint64))


def make_tx(n: int, n_accounts: int, seed: int = MASTER_SEED, zipf_s: float = 1.1,
            t_lo: int = T0 + 24 * 3600, t_hi: int = T0 + 48 * 3600, invalid_rate: float = 0.0) -> np.ndarray:
    """Compact tx records (queries) with C1 marginals; ts uniform in [t_lo, t_hi)."""
    s = seed + 303
    U = lambda k: uniform(s, k, n)  # noqa: E731
    tx = np.zeros(n, dtype=abi.tx_dtype)
    tx["account"] = zipf_accounts(U(1), n_accounts, zipf_s)
    tx["ts"] = (t_lo + U(2) * (t_hi - t_lo)).astype(np.uint32)
    tx["seq_limit"] = abi.SEQ_ALL
    tx["amount"] = amounts(U(3))
    tx["tx_type"] = tx_types(U(4))
    tx["unique_devices_24h"] = 1 + geometric(U(5), 0.9)
    tx["unique_ips_24h"] = 1 + geometric(U(6), 1.6)
    bits = ((U(7) < 0.02) * abi.BIT_VPN + (U(8) < 0.01) * abi.BIT_PROXY + (U(9) < 0.005) * abi.BIT_TOR
            + (U(10) < 0.001) * abi.BIT_BLACKLISTED + (U(11) >= invalid_rate) * abi.BIT_VALID)
    tx["bits"] = bits.astype(np.uint8)
    return tx


def queries_of(tx: np.ndarray) -> np.ndarray:
    q = np.zeros(tx.shape[0], dtype=abi.query_dtype)
    q["account"] = tx["account"]
    q["ts"] = tx["ts"]
    q["seq_limit"] = tx["seq_limit"]
    return q


def make_ltv_inputs(n: int, seed: int = MASTER_SEED):
    """C4: n x 256 bf16 N(0,1) model input + epilogue sideband rows."""
    s = seed + 404
    x = f32_to_bf16(normal(s, 1, n * abi.MLP_IN).astype(np.float32)).reshape(n, abi.MLP_IN)
    U = lambda k: uniform(s, k, n)  # noqa: E731
    r = np.zeros(n, dtype=abi.ltv_row_dtype)
    r["days_since_registration"] = geometric(U(10), 200.0)
    r["days_since_last_deposit"] = geometric(U(11), 12.0)
    r["days_since_last_bet"] = geometric(U(12), 8.0)
    r["sessions_per_week"] = np.round(U(13) * 8.0, 1)
    r["deposit_frequency"] = np.round(U(14) * 6.0, 1)
    dep = np.floor(np.exp(5.0 + 2.0 * normal(s, 15, n)))
    r["total_deposits"] = dep
    r["total_withdrawals"] = np.floor(dep * U(16) * 1.3)
    r["net_revenue"] = r["total_deposits"] - r["total_withdrawals"]
    r["bonus_conversion_rate"] = U(17)
    r["bet_count"] = geometric(U(18), 80.0)
    r["support_tickets"] = poisson(U(19), 1.2)
    r["bonuses_claimed"] = poisson(U(20), 2.0)
    r["games_played"] = geometric(U(21), 6.0)
    r["push_enabled"] = U(22) < 0.5
    r["email_opt_in"] = U(23) < 0.6
    r["has_vip_manager"] = U(24) < 0.05
    return x, r
