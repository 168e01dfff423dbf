"""ctypes binding of oracle/liboracle_igs.so — the CPU restatement of the reference path.

TEST INFRASTRUCTURE ONLY (see oracle/igs_oracle.h).  Imported by tests/, by
__graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs; never by the
product package.
"""
from __future__ import annotations

import ctypes as C
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
import __graft_entry__ as _ge  # noqa: E402

pkg = _ge.load_package()
abi = pkg.abi

LIB_PATH = os.path.join(_HERE, "liboracle_igs.so")


class Features(C.Structure):
    """orc_features = scoring.FeatureVector (scoring/engine.go:67-105)."""
    _fields_ = [
        ("tx_count_1m", C.c_int64), ("tx_count_5m", C.c_int64), ("tx_count_1h", C.c_int64),
        ("tx_sum_1h", C.c_int64), ("tx_avg_1h", C.c_double),
        ("unique_devices_24h", C.c_int64), ("unique_ips_24h", C.c_int64),
        ("ip_country_changes", C.c_int64), ("device_age", C.c_int64),
        ("account_age", C.c_int64), ("total_deposits", C.c_int64),
        ("total_withdrawals", C.c_int64), ("net_deposit", C.c_int64),
        ("deposit_count", C.c_int64), ("withdraw_count", C.c_int64),
        ("time_since_last_tx", C.c_int64), ("session_duration", C.c_int64),
        ("avg_bet_size", C.c_double), ("win_rate", C.c_double),
        ("is_vpn", C.c_int32), ("is_proxy", C.c_int32), ("is_tor", C.c_int32),
        ("disposable_email", C.c_int32), ("bonus_claim_count", C.c_int64),
        ("bonus_wager_rate", C.c_double), ("bonus_only_player", C.c_int32), ("pad_", C.c_int32),
    ]


class Realtime(C.Structure):
    _fields_ = [(k, C.c_int64) for k in (
        "tx_count_1m", "tx_count_5m", "tx_count_1h", "tx_sum_1h", "unique_devices_24h",
        "unique_ips_24h", "last_tx_timestamp", "session_start")]


class Blacklist(C.Structure):
    _fields_ = [("device", C.c_void_p), ("n_device", C.c_uint64),
                ("fingerprint", C.c_void_p), ("n_fingerprint", C.c_uint64),
                ("ip", C.c_void_p), ("n_ip", C.c_uint64)]


class Score(C.Structure):
    _fields_ = [("score", C.c_int32), ("rule_score", C.c_int32), ("ml_score", C.c_double),
                ("flags", C.c_uint32), ("action", C.c_int32)]


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        _ge.build_oracle()
    L = C.CDLL(LIB_PATH)
    vp, u64 = C.c_void_p, C.c_uint64
    L.orc_default_config.argtypes = [C.POINTER(abi.Config)]
    L.orc_default_config.restype = None
    L.orc_extract_features.argtypes = [C.POINTER(Realtime), vp, C.c_int64, C.c_uint32, C.POINTER(Features)]
    L.orc_extract_features.restype = None
    L.orc_apply_rules.argtypes = [C.POINTER(abi.Config), C.c_int64, C.c_int, C.POINTER(Features),
                                  C.c_int, C.POINTER(C.c_uint32)]
    L.orc_apply_rules.restype = C.c_int32
    L.orc_blend.argtypes = [C.POINTER(abi.Config), C.c_int32, C.c_uint32, C.c_double, C.c_int,
                            C.POINTER(Score)]
    L.orc_blend.restype = None
    L.orc_to_ml.argtypes = [C.POINTER(Features), C.c_int64, C.c_int, vp]
    L.orc_to_ml.restype = None
    L.orc_normalize.argtypes = [vp]
    L.orc_normalize.restype = None
    L.orc_mock_predict.argtypes = [vp]
    L.orc_mock_predict.restype = C.c_double
    L.orc_sigmoid.argtypes = [C.c_float]
    L.orc_sigmoid.restype = C.c_float
    L.orc_gbdt_margin.argtypes = [C.POINTER(abi.GbdtDesc), vp, vp]
    L.orc_gbdt_margin.restype = C.c_float
    L.orc_gbdt_predict.argtypes = [C.POINTER(abi.GbdtDesc), vp, vp]
    L.orc_gbdt_predict.restype = C.c_float
    L.orc_score_features_row.argtypes = [C.POINTER(abi.Config), C.POINTER(abi.GbdtDesc), vp, vp,
                                         C.POINTER(Score), vp]
    L.orc_score_features_row.restype = None
    L.orc_score_features_batch.argtypes = [C.POINTER(abi.Config), C.POINTER(abi.GbdtDesc), vp, vp,
                                           u64, vp, vp, C.c_int]
    L.orc_score_features_batch.restype = None
    L.orc_redis_new.argtypes = []
    L.orc_redis_new.restype = vp
    L.orc_redis_free.argtypes = [vp]
    L.orc_redis_free.restype = None
    L.orc_redis_get.argtypes = [vp, C.c_int64, C.POINTER(Realtime)]
    L.orc_redis_get.restype = None
    L.orc_redis_update.argtypes = [vp, C.c_int64, C.c_int64]
    L.orc_redis_update.restype = None
    L.orc_velocity_closed_form.argtypes = [vp, vp, vp, C.c_uint32, C.c_uint32, vp]
    L.orc_velocity_closed_form.restype = None
    L.orc_velocity_batch.argtypes = [C.POINTER(abi.EventCols), u64, C.c_uint32, vp, u64, vp]
    L.orc_velocity_batch.restype = None
    L.orc_assemble.argtypes = [vp, vp, vp, C.POINTER(Features), vp, vp]
    L.orc_assemble.restype = None
    L.orc_score_batch.argtypes = [C.POINTER(abi.Config), C.POINTER(abi.GbdtDesc), vp, u64,
                                  C.POINTER(abi.EventCols), u64, vp, u64, vp, vp, vp]
    L.orc_score_batch.restype = None
    L.orc_redis_update_ids.argtypes = [vp, C.c_int64, C.c_int64, u64, u64]
    L.orc_redis_update_ids.restype = None
    L.orc_uniques_closed_form.argtypes = [vp, vp, C.c_uint32, C.c_uint32]
    L.orc_uniques_closed_form.restype = C.c_uint32
    L.orc_uniques_batch.argtypes = [C.POINTER(abi.EventCols), C.POINTER(abi.EventIds), u64, C.c_uint32,
                                    vp, u64, vp]
    L.orc_uniques_batch.restype = None
    L.orc_hash_id.argtypes = [C.c_char_p, C.c_size_t]
    L.orc_hash_id.restype = u64
    L.orc_blacklist_check.argtypes = [C.POINTER(Blacklist), vp]
    L.orc_blacklist_check.restype = C.c_int
    L.orc_score_batch_ids.argtypes = [C.POINTER(abi.Config), C.POINTER(abi.GbdtDesc), vp, u64,
                                      C.POINTER(abi.EventCols), C.POINTER(abi.EventIds), u64,
                                      C.POINTER(Blacklist), vp, vp, u64, vp, vp]
    L.orc_score_batch_ids.restype = None
    L.orc_fast_index_build.argtypes = [C.POINTER(abi.EventCols), u64, C.c_uint32]
    L.orc_fast_index_build.restype = vp
    L.orc_fast_index_free.argtypes = [vp]
    L.orc_fast_index_free.restype = None
    L.orc_fast_velocity_batch.argtypes = [vp, vp, u64, vp]
    L.orc_fast_velocity_batch.restype = None
    L.orc_score_batch_indexed.argtypes = [C.POINTER(abi.Config), C.POINTER(abi.GbdtDesc), vp, u64, vp, vp, u64, vp,
                                          C.c_int]
    L.orc_score_batch_indexed.restype = None
    L.orc_links.argtypes = [vp, vp, u64, vp, u64, vp]
    L.orc_links.restype = None
    L.orc_check_bonus_abuse.argtypes = [vp, vp, vp]
    L.orc_check_bonus_abuse.restype = None
    for f in ("orc_ltv_engagement", "orc_ltv_heuristic", "orc_ltv_churn"):
        getattr(L, f).argtypes = [vp]
        getattr(L, f).restype = C.c_double
    L.orc_ltv_epilogue.argtypes = [vp, C.c_double, vp]
    L.orc_ltv_epilogue.restype = None
    for f in ("orc_ltv_model_engagement", "orc_ltv_model_churn", "orc_ltv_model_predict"):
        getattr(L, f).argtypes = [vp]
        getattr(L, f).restype = C.c_double
    L.orc_bf16_to_f32.argtypes = [C.c_uint16]
    L.orc_bf16_to_f32.restype = C.c_float
    L.orc_f32_to_bf16.argtypes = [C.c_float]
    L.orc_f32_to_bf16.restype = C.c_uint16
    L.orc_mlp_forward.argtypes = [C.POINTER(abi.MlpDesc), vp, vp]
    L.orc_mlp_forward.restype = C.c_float
    L.orc_mlp_layer.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp]
    L.orc_mlp_layer.restype = None
    _lib = L
    return L


# ------------------------------------------------------------------------------------------------
# convenience wrappers
# ------------------------------------------------------------------------------------------------
def default_config(**overrides) -> abi.Config:
    cfg = abi.Config()
    lib().orc_default_config(C.byref(cfg))
    for k, v in overrides.items():
        if k == "reason_weight":
            for i, w in enumerate(v):
                cfg.reason_weight[i] = w
        else:
            setattr(cfg, k, v)
    return cfg


def sigmoid(m: float) -> np.float32:
    return np.float32(lib().orc_sigmoid(C.c_float(float(m))))


def score_features(cfg, model, features, rows, want_leaves=False, threads=1):
    """Path A over a batch.  model may be None unless cfg.model_kind == MODEL_GBDT."""
    features = np.ascontiguousarray(features, dtype=np.float32)
    rows = np.ascontiguousarray(rows, dtype=abi.rule_row_dtype)
    n = rows.shape[0]
    out = np.zeros(n, dtype=abi.result_dtype)
    d = model.desc() if model is not None else None
    leaves = np.zeros((n, model.n_trees), dtype=np.uint8) if (want_leaves and model is not None) else None
    lib().orc_score_features_batch(C.byref(cfg), C.byref(d) if d is not None else None,
                                   abi.ptr(features), abi.ptr(rows), n, abi.ptr(out),
                                   abi.ptr(leaves), threads)
    return (out, leaves) if want_leaves else out


def score_row(cfg, model, x64, row) -> Score:
    x64 = np.ascontiguousarray(x64, dtype=np.float32)
    row = np.ascontiguousarray(row, dtype=abi.rule_row_dtype).reshape(1)
    s = Score()
    d = model.desc() if model is not None else None
    lib().orc_score_features_row(C.byref(cfg), C.byref(d) if d is not None else None,
                                 abi.ptr(x64), abi.ptr(row), C.byref(s), None)
    return s


def velocity_batch(ev, n_accounts: int, queries: np.ndarray) -> np.ndarray:
    queries = np.ascontiguousarray(queries, dtype=abi.query_dtype)
    out = np.zeros(queries.shape[0], dtype=abi.velocity_dtype)
    c = ev.cols()
    lib().orc_velocity_batch(C.byref(c), len(ev), n_accounts, abi.ptr(queries), queries.shape[0],
                             abi.ptr(out))
    return out


def score_batch(cfg, model, players, ev, tx, want_features=False, want_velocity=False):
    players = np.ascontiguousarray(players, dtype=abi.player_row_dtype)
    tx = np.ascontiguousarray(tx, dtype=abi.tx_dtype)
    n = tx.shape[0]
    out = np.zeros(n, dtype=abi.result_dtype)
    feats = np.zeros((n, abi.NUM_FEATURES), dtype=np.float32) if want_features else None
    vel = np.zeros(n, dtype=abi.velocity_dtype) if want_velocity else None
    d = model.desc() if model is not None else None
    c = ev.cols()
    lib().orc_score_batch(C.byref(cfg), C.byref(d) if d is not None else None, abi.ptr(players),
                          players.shape[0], C.byref(c), len(ev), abi.ptr(tx), n, abi.ptr(out),
                          abi.ptr(feats), abi.ptr(vel))
    if want_features or want_velocity:
        return out, feats, vel
    return out


class FastIndex:
    """Prefix-column index over a history (orc_fast_index): O(log n) lookups, the timed CPU baseline of the fused
    workload.  Build once, query / score many times."""

    def __init__(self, ev, n_accounts: int):
        c = ev.cols()
        self._keep = (ev, c)
        self.n_events, self.n_accounts = len(ev), n_accounts
        self._h = lib().orc_fast_index_build(C.byref(c), len(ev), n_accounts)

    def close(self):
        if getattr(self, "_h", None):
            lib().orc_fast_index_free(self._h)
            self._h = None

    __del__ = close

    def velocity(self, queries: np.ndarray) -> np.ndarray:
        queries = np.ascontiguousarray(queries, dtype=abi.query_dtype)
        out = np.zeros(queries.shape[0], dtype=abi.velocity_dtype)
        lib().orc_fast_velocity_batch(self._h, abi.ptr(queries), queries.shape[0], abi.ptr(out))
        return out

    def score_batch(self, cfg, model, players, tx, threads: int = 1) -> np.ndarray:
        players = np.ascontiguousarray(players, dtype=abi.player_row_dtype)
        tx = np.ascontiguousarray(tx, dtype=abi.tx_dtype)
        out = np.zeros(tx.shape[0], dtype=abi.result_dtype)
        d = model.desc() if model is not None else None
        lib().orc_score_batch_indexed(C.byref(cfg), C.byref(d) if d is not None else None, abi.ptr(players),
                                      players.shape[0], self._h, abi.ptr(tx), tx.shape[0], abi.ptr(out), threads)
        return out


def uniques_batch(ev, n_accounts: int, queries: np.ndarray) -> np.ndarray:
    queries = np.ascontiguousarray(queries, dtype=abi.query_dtype)
    out = np.zeros(queries.shape[0], dtype=abi.uniques_dtype)
    c, i = ev.cols(), ev.ids()
    lib().orc_uniques_batch(C.byref(c), C.byref(i), len(ev), n_accounts, abi.ptr(queries),
                            queries.shape[0], abi.ptr(out))
    return out


def hash_id(s) -> int:
    b = s.encode() if isinstance(s, str) else bytes(s)
    return int(lib().orc_hash_id(b, len(b)))


def make_blacklist(device=None, fingerprint=None, ip=None):
    """-> (Blacklist struct, keep-alive arrays)"""
    arrs = [np.ascontiguousarray(a if a is not None else [], dtype=np.uint64) for a in (device, fingerprint, ip)]
    bl = Blacklist(abi.ptr(arrs[0]), arrs[0].shape[0], abi.ptr(arrs[1]), arrs[1].shape[0],
                   abi.ptr(arrs[2]), arrs[2].shape[0])
    return bl, arrs


def blacklist_check(bl: Blacklist, ids: np.ndarray) -> np.ndarray:
    ids = np.ascontiguousarray(ids, dtype=abi.tx_ids_dtype)
    L = lib()
    return np.array([L.orc_blacklist_check(C.byref(bl), ids[i:i + 1].ctypes.data) for i in range(ids.shape[0])],
                    dtype=np.uint8)


def score_batch_ids(cfg, model, players, ev, tx, bl=None, tx_ids=None, want_features=False):
    """Path B with uniques from the history's identifier columns (when ev has them) and R8 from bl."""
    players = np.ascontiguousarray(players, dtype=abi.player_row_dtype)
    tx = np.ascontiguousarray(tx, dtype=abi.tx_dtype)
    n = tx.shape[0]
    out = np.zeros(n, dtype=abi.result_dtype)
    feats = np.zeros((n, abi.NUM_FEATURES), dtype=np.float32) if want_features else None
    d = model.desc() if model is not None else None
    c = ev.cols()
    i = ev.ids() if ev.has_ids() else None
    if tx_ids is not None:
        tx_ids = np.ascontiguousarray(tx_ids, dtype=abi.tx_ids_dtype)
    lib().orc_score_batch_ids(C.byref(cfg), C.byref(d) if d is not None else None, abi.ptr(players),
                              players.shape[0], C.byref(c), C.byref(i) if i is not None else None, len(ev),
                              C.byref(bl) if bl is not None else None, abi.ptr(tx), abi.ptr(tx_ids), n,
                              abi.ptr(out), abi.ptr(feats))
    return (out, feats) if want_features else out


def links(device: np.ndarray, account: np.ndarray, query: np.ndarray) -> np.ndarray:
    """Linked accounts over all shards at once: device / global account id per event -> rows of `query`."""
    device = np.ascontiguousarray(device, dtype=np.uint64)
    account = np.ascontiguousarray(account, dtype=np.uint64)
    query = np.ascontiguousarray(query, dtype=np.uint64)
    out = np.zeros(query.shape[0], dtype=abi.links_row_dtype)
    lib().orc_links(abi.ptr(device), abi.ptr(account), device.shape[0], abi.ptr(query), query.shape[0], abi.ptr(out))
    return out


def check_bonus_abuse(players, rows: np.ndarray, accounts: np.ndarray) -> np.ndarray:
    """players: BatchFeatures table (or None); rows[i] = links row of accounts[i]."""
    rows = np.ascontiguousarray(rows, dtype=abi.links_row_dtype)
    out = np.zeros(rows.shape[0], dtype=abi.abuse_dtype)
    L = lib()
    for i, a in enumerate(accounts):
        p = players[int(a):int(a) + 1].ctypes.data if players is not None and int(a) < players.shape[0] else None
        L.orc_check_bonus_abuse(p, rows[i:i + 1].ctypes.data, out[i:i + 1].ctypes.data)
    return out


def ltv_epilogue(rows: np.ndarray, ltv: np.ndarray) -> np.ndarray:
    rows = np.ascontiguousarray(rows, dtype=abi.ltv_row_dtype)
    out = np.zeros(rows.shape[0], dtype=abi.ltv_result_dtype)
    L = lib()
    for i in range(rows.shape[0]):
        L.orc_ltv_epilogue(rows[i:i + 1].ctypes.data, float(ltv[i]), out[i:i + 1].ctypes.data)
    return out


def ltv_predict_batch(rows: np.ndarray):
    """LTVPredictor.BatchPredict as the reference runs it (prediction/ltv.go:385-398): calculateLTV, then the
    post-model steps.  Returns (results, calculateLTV values)."""
    rows = np.ascontiguousarray(rows, dtype=abi.ltv_row_dtype)
    L = lib()
    raw = np.array([L.orc_ltv_heuristic(rows[i:i + 1].ctypes.data) for i in range(rows.shape[0])], dtype=np.float64)
    return ltv_epilogue(rows, raw), raw


def ltv_model_predict(feats: np.ndarray) -> np.ndarray:
    """ml.LTVModel.Predict (ml/onnx_model.go:405-432) per row of ml.LTVFeatures."""
    feats = np.ascontiguousarray(feats, dtype=abi.ltv_features_dtype)
    L = lib()
    return np.array([L.orc_ltv_model_predict(feats[i:i + 1].ctypes.data) for i in range(feats.shape[0])],
                    dtype=np.float64)


def mlp_forward(model, x_bf16: np.ndarray, want_hidden=False):
    x_bf16 = np.ascontiguousarray(x_bf16, dtype=np.uint16)
    n = x_bf16.shape[0]
    y = np.zeros(n, dtype=np.float32)
    hid = np.zeros((n, 3, abi.MLP_HIDDEN), dtype=np.uint16) if want_hidden else None
    d = model.desc()
    L = lib()
    for i in range(n):
        y[i] = L.orc_mlp_forward(C.byref(d), x_bf16[i].ctypes.data,
                                 hid[i].ctypes.data if want_hidden else None)
    return (y, hid) if want_hidden else y
