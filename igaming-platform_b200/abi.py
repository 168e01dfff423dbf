"""ctypes / numpy view of include/igs.h.

Row layouts are numpy structured dtypes (so a batch is one contiguous array that can be handed
to the C ABI, to pinned memory or to the device unchanged); descriptors are ctypes Structures.
Nothing here computes anything: the only compute path is the CUDA library behind `load_library`.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

NUM_FEATURES = 64
REF_FEATURES = 30
TREE_DEPTH = 8
TREE_NODES = 255
TREE_LEAVES = 256
NUM_SLOTS = 4
NUM_REASONS = 10
MLP_IN = 256
MLP_HIDDEN = 1024

OK, ERR_INVALID, ERR_CUDA, ERR_STATE, ERR_NOMEM = 0, -1, -2, -3, -4

# Reason bits, order of the ReasonCode consts (scoring/engine.go:17-28).
R_HIGH_VELOCITY = 1 << 0
R_NEW_ACCOUNT_LARGE_TX = 1 << 1
R_IP_COUNTRY_MISMATCH = 1 << 2
R_MULTIPLE_DEVICES = 1 << 3
R_SUSPICIOUS_PATTERN = 1 << 4
R_VPN_DETECTED = 1 << 5
R_KNOWN_FRAUDSTER = 1 << 6
R_RAPID_DEPOSIT_WITHDRAW = 1 << 7
R_BONUS_ABUSE = 1 << 8
R_ML_HIGH_RISK = 1 << 9

ACTION_APPROVE, ACTION_REVIEW, ACTION_BLOCK = 1, 2, 3
TX_BET, TX_DEPOSIT, TX_WITHDRAW, TX_OTHER = 0, 1, 2, 3
BIT_VPN, BIT_PROXY, BIT_TOR, BIT_BONUS_ONLY, BIT_BLACKLISTED, BIT_VALID = 1, 2, 4, 8, 16, 32
MODEL_NONE, MODEL_MOCK, MODEL_GBDT = 0, 1, 2
POST_NONE, POST_LOGISTIC = 0, 1
SEG_VIP, SEG_HIGH, SEG_MEDIUM, SEG_LOW, SEG_CHURNING = 1, 2, 3, 4, 5
SEQ_ALL = 0xFFFFFFFF
PARTITION_STATIC_WORDS = 256 * 1024
BL_DEVICE, BL_FINGERPRINT, BL_IP = 0, 1, 2
LATE_REJECT, LATE_REORDER = 0, 1
LINK_MAX_ACCOUNTS = 64
IPC_HANDLE_BYTES = 64
ABUSE_BONUS_ONLY, ABUSE_MULTI_ACCOUNT, ABUSE_ACCOUNT_RING = 1, 2, 4

rule_row_dtype = np.dtype([
    ("amount", "<i8"), ("total_deposits", "<i8"), ("total_withdrawals", "<i8"),
    ("tx_count_1m", "<u4"), ("account_age_days", "<i4"), ("time_since_last_tx", "<u4"),
    ("deposit_count", "<u4"), ("unique_devices_24h", "<u2"), ("unique_ips_24h", "<u2"),
    ("tx_type", "u1"), ("bits", "u1"), ("pad", "u1", (2,)),
])
result_dtype = np.dtype([("ml_score", "<f4"), ("flags", "<u2"), ("score", "u1"), ("action", "u1")])
tx_dtype = np.dtype([
    ("amount", "<i8"), ("account", "<u4"), ("ts", "<u4"), ("seq_limit", "<u4"),
    ("unique_devices_24h", "<u2"), ("unique_ips_24h", "<u2"), ("tx_type", "u1"), ("bits", "u1"),
    ("pad", "u1", (6,)),
])
player_row_dtype = np.dtype([
    ("total_deposits", "<i8"), ("total_withdrawals", "<i8"), ("total_bets", "<i8"),
    ("total_wins", "<i8"), ("account_created_at", "<i8"), ("avg_bet_size", "<f8"),
    ("bonus_wager_complete", "<f8"), ("deposit_count", "<i4"), ("withdraw_count", "<i4"),
    ("bet_count", "<i4"), ("win_count", "<i4"), ("bonus_claim_count", "<i4"),
    ("reserved", "<i4", (5,)),
])
query_dtype = np.dtype([("account", "<u4"), ("ts", "<u4"), ("seq_limit", "<u4"), ("reserved", "<u4")])
window_dtype = np.dtype([("sum", "<i8"), ("count", "<u4"), ("pad", "<u4")])
velocity_dtype = np.dtype([
    ("tx_sum_1h", "<i8"), ("tx_count_1m", "<u4"), ("tx_count_5m", "<u4"), ("tx_count_1h", "<u4"),
    ("time_since_last_tx", "<u4"), ("session_duration", "<u4"), ("n_prior", "<u4"),
    ("win", window_dtype, (2, 3)),
])
tx_ids_dtype = np.dtype([("device", "<u8"), ("fingerprint", "<u8"), ("ip", "<u8")])
uniques_dtype = np.dtype([("devices_24h", "<u4"), ("ips_24h", "<u4")])
link_pair_dtype = np.dtype([("device", "<u8"), ("account", "<u8")])
link_dtype = np.dtype([("account", "<u8"), ("other", "<u8")])
links_row_dtype = np.dtype([("n_linked", "<u4"), ("reserved", "<u4"), ("linked", "<u8", (3,))])
abuse_dtype = np.dtype([("abuse_score", "<f4"), ("n_linked", "<u4"), ("is_abuser", "u1"), ("signals", "u1"),
                        ("pad", "u1", (6,))])
ltv_row_dtype = np.dtype([
    ("sessions_per_week", "<f8"), ("deposit_frequency", "<f8"), ("total_deposits", "<f8"),
    ("total_withdrawals", "<f8"), ("net_revenue", "<f8"), ("bonus_conversion_rate", "<f8"),
    ("days_since_registration", "<i4"), ("days_since_last_deposit", "<i4"),
    ("days_since_last_bet", "<i4"), ("bet_count", "<i4"), ("support_tickets", "<i4"),
    ("bonuses_claimed", "<i4"), ("games_played", "<i4"),
    ("push_enabled", "u1"), ("email_opt_in", "u1"), ("has_vip_manager", "u1"), ("pad", "u1"),
])
ltv_result_dtype = np.dtype([
    ("predicted_ltv", "<f8"), ("churn_risk", "<f8"), ("confidence", "<f4"), ("days_survival", "<i4"),
    ("segment", "u1"), ("next_best_action", "u1"), ("pad", "u1", (6,)),
])

# ml.LTVFeatures, ml/onnx_model.go:375-393 (declaration order)
ltv_features_dtype = np.dtype([(k, "<f4") for k in (
    "days_since_registration", "days_since_last_deposit", "days_since_last_bet", "total_active_days",
    "sessions_per_week", "total_deposits", "total_withdrawals", "net_revenue", "avg_deposit_amount",
    "deposit_frequency", "total_bets", "bet_count", "win_rate", "avg_bet_size", "games_played",
    "bonuses_claimed", "bonus_conversion_rate")])

assert ltv_features_dtype.itemsize == 68
assert rule_row_dtype.itemsize == 48
assert result_dtype.itemsize == 8
assert tx_dtype.itemsize == 32
assert player_row_dtype.itemsize == 96
assert query_dtype.itemsize == 16
assert velocity_dtype.itemsize == 128
assert tx_ids_dtype.itemsize == 24 and uniques_dtype.itemsize == 8
assert link_pair_dtype.itemsize == 16 and link_dtype.itemsize == 16 and links_row_dtype.itemsize == 32 and abuse_dtype.itemsize == 16
assert ltv_row_dtype.itemsize == 80
assert ltv_result_dtype.itemsize == 32


class Config(C.Structure):
    """igs_config — ScoringConfig (scoring/engine.go:196-228) + ruleWeights (:246-257)."""
    _fields_ = [
        ("block_threshold", C.c_int32), ("review_threshold", C.c_int32),
        ("max_tx_per_minute", C.c_int32), ("max_tx_per_hour", C.c_int32),
        ("new_account_days", C.c_int32), ("max_devices_per_day", C.c_int32),
        ("max_ips_per_day", C.c_int32), ("model_kind", C.c_int32),
        ("large_deposit_amount", C.c_int64), ("ml_weight", C.c_double), ("rule_weight", C.c_double),
        ("reason_weight", C.c_int32 * NUM_REASONS), ("reserved", C.c_int32 * 2),
    ]


class GbdtDesc(C.Structure):
    _fields_ = [
        ("n_trees", C.c_uint32), ("n_features", C.c_uint32), ("post_transform", C.c_uint32),
        ("base_score", C.c_float),
        ("feat", C.c_void_p), ("thr", C.c_void_p), ("leaf", C.c_void_p),
    ]


class MlpDesc(C.Structure):
    _fields_ = [
        ("w1", C.c_void_p), ("b1", C.c_void_p), ("w2", C.c_void_p), ("b2", C.c_void_p),
        ("w3", C.c_void_p), ("b3", C.c_void_p), ("w4", C.c_void_p), ("b4", C.c_void_p),
    ]


class EventCols(C.Structure):
    _fields_ = [("account", C.c_void_p), ("ts", C.c_void_p), ("amount", C.c_void_p),
                ("tx_type", C.c_void_p)]


class EventIds(C.Structure):
    _fields_ = [("device", C.c_void_p), ("ip", C.c_void_p)]


class HistoryInfo(C.Structure):
    _fields_ = [("n_events", C.c_uint64), ("late_events", C.c_uint64), ("n_accounts", C.c_uint32),
                ("has_identifiers", C.c_uint32), ("delta_events", C.c_uint64), ("reserved", C.c_uint64 * 3)]


assert C.sizeof(Config) == 104 and C.sizeof(HistoryInfo) == 56

# Every symbol include/igs.h declares; tests check the built library exports each of them.
EXPORTS = [
    "igs_default_config", "igs_engine_create", "igs_engine_destroy", "igs_last_error",
    "igs_alloc_pinned", "igs_free_pinned", "igs_load_gbdt", "igs_load_mlp", "igs_load_players",
    "igs_load_events", "igs_append_events", "igs_score_features", "igs_score_features_async", "igs_wait",
    "igs_score_features_device", "igs_get_velocity", "igs_get_velocity_device", "igs_score_batch",
    "igs_score_batch_async", "igs_score_batch_device", "igs_partition_tx_device", "igs_partition_tx_static",
    "igs_ltv_batch", "igs_ltv_batch_device", "igs_ltv_hidden_device",
    "igs_rule_score_from_flags", "igs_reason_codes", "igs_action_name", "igs_ltv_action_name",
    "igs_segment_name", "igs_shard_of", "igs_version",
    "igs_load_events_ids", "igs_append_events_ids", "igs_get_uniques", "igs_get_uniques_device",
    "igs_blacklist_load", "igs_blacklist_add", "igs_blacklist_check", "igs_blacklist_check_device",
    "igs_score_batch_ids", "igs_score_batch_ids_device", "igs_hash_id",
    "igs_links_pairs", "igs_links_group", "igs_links_finish", "igs_get_links", "igs_check_bonus_abuse",
    "igs_abuse_signals", "igs_links_window_create", "igs_links_window_open", "igs_links_window_attach",
    "igs_links_window_reset", "igs_links_pairs_peer", "igs_links_group_peer", "igs_links_finish_peer",
    "igs_evict_events",
    "igs_set_compaction_threshold", "igs_prepare_ingest", "igs_kernel_launches", "igs_load_events_device", "igs_set_late_policy", "igs_get_history_info",
    "igs_ltv_predict_batch", "igs_ltv_predict_batch_device", "igs_ltv_mock_predict", "igs_ltv_mock_predict_device",
]

LIB_NAME = "libigs.so"


def library_path() -> str:
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), LIB_NAME)


_lib = None


def load_library() -> C.CDLL:
    """dlopen the CUDA library built by __graft_entry__.build(); fail loudly if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} is missing: the CUDA extension was not built "
            "(run `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback.")
    lib = C.CDLL(path)
    vp, u64, i32 = C.c_void_p, C.c_uint64, C.c_int
    lib.igs_default_config.argtypes = [C.POINTER(Config)]
    lib.igs_default_config.restype = None
    lib.igs_engine_create.argtypes = [i32, C.POINTER(vp)]
    lib.igs_engine_destroy.argtypes = [vp]
    lib.igs_engine_destroy.restype = None
    lib.igs_last_error.argtypes = [vp]
    lib.igs_last_error.restype = C.c_char_p
    lib.igs_alloc_pinned.argtypes = [C.c_size_t]
    lib.igs_alloc_pinned.restype = vp
    lib.igs_free_pinned.argtypes = [vp]
    lib.igs_free_pinned.restype = None
    lib.igs_load_gbdt.argtypes = [vp, C.POINTER(GbdtDesc)]
    lib.igs_load_mlp.argtypes = [vp, C.POINTER(MlpDesc)]
    lib.igs_load_players.argtypes = [vp, vp, u64]
    lib.igs_load_events.argtypes = [vp, C.POINTER(EventCols), u64, C.c_uint32]
    lib.igs_append_events.argtypes = [vp, C.POINTER(EventCols), u64]
    lib.igs_score_features.argtypes = [vp, C.POINTER(Config), vp, vp, u64, vp, vp]
    lib.igs_score_features_async.argtypes = [vp, i32, C.POINTER(Config), vp, vp, u64, vp]
    lib.igs_wait.argtypes = [vp, i32]
    lib.igs_score_features_device.argtypes = [vp, C.POINTER(Config), vp, vp, u64, vp, vp, vp]
    lib.igs_get_velocity.argtypes = [vp, vp, u64, vp]
    lib.igs_get_velocity_device.argtypes = [vp, vp, u64, vp, vp]
    lib.igs_score_batch.argtypes = [vp, C.POINTER(Config), vp, u64, vp, vp, vp]
    lib.igs_score_batch_async.argtypes = [vp, i32, C.POINTER(Config), vp, u64, vp]
    lib.igs_score_batch_device.argtypes = [vp, C.POINTER(Config), vp, u64, vp, vp, vp, vp]
    lib.igs_partition_tx_device.argtypes = [vp, vp, u64, C.c_uint32, vp, vp, vp, vp]
    lib.igs_partition_tx_static.argtypes = [vp, vp, u64, C.c_uint32, vp, vp, vp, vp]
    lib.igs_ltv_batch.argtypes = [vp, vp, vp, u64, vp, vp]
    lib.igs_ltv_batch_device.argtypes = [vp, vp, vp, u64, vp, vp, vp]
    lib.igs_ltv_hidden_device.argtypes = [vp, vp, u64, i32, vp, vp]
    lib.igs_rule_score_from_flags.argtypes = [C.POINTER(Config), C.c_uint32]
    lib.igs_rule_score_from_flags.restype = C.c_int32
    lib.igs_reason_codes.argtypes = [C.c_uint32, C.POINTER(C.c_char_p * NUM_REASONS)]
    lib.igs_action_name.argtypes = [i32]
    lib.igs_action_name.restype = C.c_char_p
    lib.igs_ltv_action_name.argtypes = [i32]
    lib.igs_ltv_action_name.restype = C.c_char_p
    lib.igs_segment_name.argtypes = [i32]
    lib.igs_segment_name.restype = C.c_char_p
    lib.igs_shard_of.argtypes = [u64, u64, C.c_uint32]
    lib.igs_shard_of.restype = C.c_uint32
    lib.igs_load_events_ids.argtypes = [vp, C.POINTER(EventCols), C.POINTER(EventIds), u64, C.c_uint32]
    lib.igs_append_events_ids.argtypes = [vp, C.POINTER(EventCols), C.POINTER(EventIds), u64]
    lib.igs_get_uniques.argtypes = [vp, vp, u64, vp]
    lib.igs_get_uniques_device.argtypes = [vp, vp, u64, vp, vp]
    lib.igs_blacklist_load.argtypes = [vp, i32, vp, u64]
    lib.igs_blacklist_add.argtypes = [vp, i32, vp, u64]
    lib.igs_blacklist_check.argtypes = [vp, vp, u64, vp]
    lib.igs_blacklist_check_device.argtypes = [vp, vp, u64, vp, vp]
    lib.igs_score_batch_ids.argtypes = [vp, C.POINTER(Config), vp, vp, u64, vp, vp, vp]
    lib.igs_score_batch_ids_device.argtypes = [vp, C.POINTER(Config), vp, vp, u64, vp, vp, vp, vp]
    lib.igs_links_pairs.argtypes = [vp, C.c_uint32, C.c_uint32, C.POINTER(vp), C.POINTER(u64)]
    lib.igs_links_group.argtypes = [vp, vp, u64, C.c_uint32, C.POINTER(vp), C.POINTER(u64)]
    lib.igs_links_finish.argtypes = [vp, vp, u64, C.c_uint32]
    lib.igs_get_links.argtypes = [vp, vp, u64, vp]
    lib.igs_check_bonus_abuse.argtypes = [vp, vp, u64, vp]
    lib.igs_abuse_signals.argtypes = [C.c_uint32, C.POINTER(C.c_char_p * 3)]
    lib.igs_links_window_create.argtypes = [vp, u64, u64, vp, C.POINTER(vp)]
    lib.igs_links_window_open.argtypes = [vp, C.c_uint32, vp]
    lib.igs_links_window_attach.argtypes = [vp, C.c_uint32, vp]
    lib.igs_links_window_reset.argtypes = [vp]
    lib.igs_links_pairs_peer.argtypes = [vp, C.c_uint32, C.c_uint32]
    lib.igs_links_group_peer.argtypes = [vp, C.c_uint32, C.c_uint32]
    lib.igs_links_finish_peer.argtypes = [vp, C.c_uint32]
    lib.igs_ltv_predict_batch.argtypes = [vp, vp, u64, vp, vp]
    lib.igs_ltv_predict_batch_device.argtypes = [vp, vp, u64, vp, vp, vp]
    lib.igs_ltv_mock_predict.argtypes = [vp, vp, u64, vp]
    lib.igs_ltv_mock_predict_device.argtypes = [vp, vp, u64, vp, vp]
    lib.igs_load_events_device.argtypes = [vp, C.POINTER(EventCols), C.POINTER(EventIds), u64, C.c_uint32]
    lib.igs_set_compaction_threshold.argtypes = [vp, u64]
    lib.igs_prepare_ingest.argtypes = [vp, u64]
    lib.igs_set_late_policy.argtypes = [vp, i32]
    lib.igs_get_history_info.argtypes = [vp, C.POINTER(HistoryInfo)]
    lib.igs_evict_events.argtypes = [vp, C.c_uint32, C.POINTER(u64)]
    lib.igs_hash_id.argtypes = [C.c_char_p, C.c_size_t]
    lib.igs_hash_id.restype = u64
    lib.igs_kernel_launches.argtypes = []
    lib.igs_kernel_launches.restype = C.c_ulonglong
    lib.igs_version.argtypes = []
    lib.igs_version.restype = C.c_char_p
    _lib = lib
    return lib


def ptr(a) -> int:
    """Host address of a numpy array (must be C-contiguous), or 0 for None."""
    if a is None:
        return 0
    if isinstance(a, int):
        return a
    assert a.flags["C_CONTIGUOUS"], "array handed to the C ABI must be contiguous"
    return a.ctypes.data
