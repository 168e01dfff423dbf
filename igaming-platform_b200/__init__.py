"""B200-native batched fraud-scoring engine — Python host side over the C ABI (include/igs.h).

`Engine` is a thin handle over `libigs.so`; `scoring.ScoringEngine` / `scoring.BatchScorer`
mirror the reference's operator interface (services/risk/internal/scoring/engine.go).  The only
compute path is the CUDA library: importing works anywhere, creating an Engine needs a B200.
"""
from __future__ import annotations

import ctypes as C
import functools
import threading
from dataclasses import dataclass

import numpy as np

from . import abi


class IgsError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"igs status {status}: {message}")
        self.status = status


def default_config(**overrides) -> abi.Config:
    """DefaultConfig() of the reference (scoring/engine.go:215-228) through the C ABI."""
    cfg = abi.Config()
    abi.load_library().igs_default_config(C.byref(cfg))
    for k, v in overrides.items():
        if k == "reason_weight":
            for i, w in enumerate(v):
                cfg.reason_weight[i] = w
        else:
            setattr(cfg, k, v)
    return cfg


@dataclass
class GbdtModel:
    """Host copy of a tree ensemble in the igs_gbdt_desc layout."""
    feat: np.ndarray   # [n_trees, 255] uint8
    thr: np.ndarray    # [n_trees, 255] float32
    leaf: np.ndarray   # [n_trees, 256] float32
    base_score: float = 0.0
    post_transform: int = abi.POST_LOGISTIC
    n_features: int = abi.NUM_FEATURES

    @property
    def n_trees(self) -> int:
        return int(self.feat.shape[0])

    def desc(self) -> abi.GbdtDesc:
        self.feat = np.ascontiguousarray(self.feat, dtype=np.uint8)
        self.thr = np.ascontiguousarray(self.thr, dtype=np.float32)
        self.leaf = np.ascontiguousarray(self.leaf, dtype=np.float32)
        assert self.feat.shape == (self.n_trees, abi.TREE_NODES)
        assert self.thr.shape == (self.n_trees, abi.TREE_NODES)
        assert self.leaf.shape == (self.n_trees, abi.TREE_LEAVES)
        return abi.GbdtDesc(self.n_trees, self.n_features, self.post_transform,
                            float(np.float32(self.base_score)),
                            abi.ptr(self.feat), abi.ptr(self.thr), abi.ptr(self.leaf))


@dataclass
class MlpModel:
    """Host copy of the LTV MLP in the igs_mlp_desc layout (bf16 bit patterns + fp32 biases)."""
    w1: np.ndarray; b1: np.ndarray
    w2: np.ndarray; b2: np.ndarray
    w3: np.ndarray; b3: np.ndarray
    w4: np.ndarray; b4: np.ndarray

    def desc(self) -> abi.MlpDesc:
        for k in ("w1", "w2", "w3", "w4"):
            setattr(self, k, np.ascontiguousarray(getattr(self, k), dtype=np.uint16))
        for k in ("b1", "b2", "b3", "b4"):
            setattr(self, k, np.ascontiguousarray(getattr(self, k), dtype=np.float32))
        assert self.w1.shape == (abi.MLP_HIDDEN, abi.MLP_IN)
        assert self.w2.shape == self.w3.shape == (abi.MLP_HIDDEN, abi.MLP_HIDDEN)
        assert self.w4.shape == (1, abi.MLP_HIDDEN)
        return abi.MlpDesc(*(abi.ptr(getattr(self, k)) for k in
                             ("w1", "b1", "w2", "b2", "w3", "b3", "w4", "b4")))


@dataclass
class Events:
    """Event history columns (igs_event_cols)."""
    account: np.ndarray  # uint32
    ts: np.ndarray       # uint32
    amount: np.ndarray   # int64
    tx_type: np.ndarray  # uint8
    device: np.ndarray | None = None   # uint64 ids (igs_hash_id), 0 = empty; igs_event_ids
    ip: np.ndarray | None = None       # uint64

    def has_ids(self) -> bool:
        return self.device is not None or self.ip is not None

    def ids(self) -> abi.EventIds:
        n = len(self)
        for k in ("device", "ip"):
            v = getattr(self, k)
            if v is not None:
                v = np.ascontiguousarray(v, dtype=np.uint64)
                assert v.shape == (n,)
                setattr(self, k, v)
        return abi.EventIds(abi.ptr(self.device), abi.ptr(self.ip))

    def __len__(self) -> int:
        return int(self.account.shape[0])

    def cols(self) -> abi.EventCols:
        self.account = np.ascontiguousarray(self.account, dtype=np.uint32)
        self.ts = np.ascontiguousarray(self.ts, dtype=np.uint32)
        self.amount = np.ascontiguousarray(self.amount, dtype=np.int64)
        self.tx_type = np.ascontiguousarray(self.tx_type, dtype=np.uint8)
        n = len(self)
        assert self.ts.shape == (n,) and self.amount.shape == (n,) and self.tx_type.shape == (n,)
        return abi.EventCols(abi.ptr(self.account), abi.ptr(self.ts), abi.ptr(self.amount),
                             abi.ptr(self.tx_type))


class Engine:
    """One engine per CUDA device (igs_engine).  Host-buffer calls are synchronous."""

    def __init__(self, device: int = 0):
        self._lib = abi.load_library()
        h = C.c_void_p()
        st = self._lib.igs_engine_create(device, C.byref(h))
        if st != abi.OK:
            raise IgsError(st, (self._lib.igs_last_error(None) or b"").decode())
        self._h = h
        self.device = device
        self._n_trees = 0
        # The synchronous host-buffer calls all stage through slot 0 and the loads replace device state
        # (include/igs.h, Threading): ctypes drops the GIL during a call, so a multi-threaded host (the gRPC
        # shim's worker pool) must not let two of them overlap.  They are serialised here.
        self._lock = threading.RLock()

    # -- life cycle ---------------------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.igs_engine_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st: int) -> None:
        if st != abi.OK:
            raise IgsError(st, (self._lib.igs_last_error(self._h) or b"").decode())

    @property
    def handle(self) -> C.c_void_p:
        return self._h

    # -- loads --------------------------------------------------------------------------------
    def load_gbdt(self, model: GbdtModel) -> None:
        d = model.desc()
        self._check(self._lib.igs_load_gbdt(self._h, C.byref(d)))
        self._n_trees = model.n_trees

    def load_mlp(self, model: MlpModel) -> None:
        d = model.desc()
        self._check(self._lib.igs_load_mlp(self._h, C.byref(d)))

    def load_players(self, rows: np.ndarray) -> None:
        rows = np.ascontiguousarray(rows, dtype=abi.player_row_dtype)
        self._check(self._lib.igs_load_players(self._h, abi.ptr(rows), rows.shape[0]))

    def load_events(self, ev: Events, n_accounts: int) -> None:
        """igs_load_events, or igs_load_events_ids when the events carry device / ip columns."""
        c = ev.cols()
        if ev.has_ids():
            i = ev.ids()
            self._check(self._lib.igs_load_events_ids(self._h, C.byref(c), C.byref(i), len(ev), n_accounts))
        else:
            self._check(self._lib.igs_load_events(self._h, C.byref(c), len(ev), n_accounts))

    def append_events(self, ev: Events) -> None:
        """UpdateFeatures for a batch of completed transactions (scoring/engine.go:486-488)."""
        c = ev.cols()
        if ev.has_ids():
            i = ev.ids()
            self._check(self._lib.igs_append_events_ids(self._h, C.byref(c), C.byref(i), len(ev)))
        else:
            self._check(self._lib.igs_append_events(self._h, C.byref(c), len(ev)))

    def set_late_policy(self, policy: int) -> None:
        """abi.LATE_REJECT (default): igs_append_events refuses batches with late events; abi.LATE_REORDER accepts
        them and orders by timestamp (see include/igs.h, LATE EVENTS)."""
        self._check(self._lib.igs_set_late_policy(self._h, policy))

    def set_compaction_threshold(self, min_delta_events: int) -> None:
        """Appended events live in a delta index folded into the base beyond max(min_delta_events, base / 8)."""
        self._check(self._lib.igs_set_compaction_threshold(self._h, min_delta_events))

    def prepare_ingest(self, max_batch_events: int) -> None:
        """Pre-allocate what appends of up to max_batch_events events will need (igs_prepare_ingest)."""
        self._check(self._lib.igs_prepare_ingest(self._h, max_batch_events))

    def history_info(self) -> dict:
        info = abi.HistoryInfo()
        self._check(self._lib.igs_get_history_info(self._h, C.byref(info)))
        return {"n_events": int(info.n_events), "late_events": int(info.late_events),
                "n_accounts": int(info.n_accounts), "has_identifiers": bool(info.has_identifiers),
                "delta_events": int(info.delta_events)}

    def evict_events(self, now: int) -> int:
        """igs_evict_events: drop what can no longer influence a query at t >= now; -> events dropped."""
        n = C.c_uint64(0)
        self._check(self._lib.igs_evict_events(self._h, now, C.byref(n)))
        return int(n.value)

    # -- identifiers: uniques + blacklist -----------------------------------------------------
    def get_uniques(self, queries: np.ndarray) -> np.ndarray:
        queries = np.ascontiguousarray(queries, dtype=abi.query_dtype)
        out = np.zeros(queries.shape[0], dtype=abi.uniques_dtype)
        self._check(self._lib.igs_get_uniques(self._h, abi.ptr(queries), queries.shape[0], abi.ptr(out)))
        return out

    def blacklist_load(self, kind: int, ids: np.ndarray) -> None:
        ids = np.ascontiguousarray(ids, dtype=np.uint64)
        self._check(self._lib.igs_blacklist_load(self._h, kind, abi.ptr(ids), ids.shape[0]))

    def blacklist_add(self, kind: int, ids: np.ndarray) -> None:
        """AddToBlacklist (features/redis_store.go:251-264) for a batch of ids."""
        ids = np.ascontiguousarray(ids, dtype=np.uint64)
        self._check(self._lib.igs_blacklist_add(self._h, kind, abi.ptr(ids), ids.shape[0]))

    def blacklist_check(self, ids: np.ndarray) -> np.ndarray:
        """CheckBlacklist (features/redis_store.go:267-293) for a batch of identifier triples."""
        ids = np.ascontiguousarray(ids, dtype=abi.tx_ids_dtype)
        out = np.zeros(ids.shape[0], dtype=np.uint8)
        self._check(self._lib.igs_blacklist_check(self._h, abi.ptr(ids), ids.shape[0], abi.ptr(out)))
        return out

    # -- linked accounts / bonus abuse (stages; links.build_links drives them) ------------------
    def links_pairs(self, shard: int, n_shards: int):
        """Stage 1 -> (device pointer to the grouped igs_link_pair records, per-shard counts)."""
        p = C.c_void_p()
        counts = (C.c_uint64 * n_shards)()
        self._check(self._lib.igs_links_pairs(self._h, shard, n_shards, C.byref(p), counts))
        return p.value or 0, np.array(list(counts), dtype=np.uint64)

    def links_group(self, d_pairs: int, n: int, n_shards: int):
        """Stage 2 -> (device pointer to the grouped igs_link records, per-shard counts)."""
        p = C.c_void_p()
        counts = (C.c_uint64 * n_shards)()
        self._check(self._lib.igs_links_group(self._h, d_pairs, n, n_shards, C.byref(p), counts))
        return p.value or 0, np.array(list(counts), dtype=np.uint64)

    def links_finish(self, d_links: int, n: int, shard: int) -> None:
        self._check(self._lib.igs_links_finish(self._h, d_links, n, shard))

    # peer-memory exchange (links.build_links_peer drives these)
    def links_window_create(self, cap_pairs: int, cap_links: int):
        """-> (64-byte IPC handle, device address of this shard's receive window)."""
        h = (C.c_uint8 * abi.IPC_HANDLE_BYTES)()
        p = C.c_void_p()
        self._check(self._lib.igs_links_window_create(self._h, cap_pairs, cap_links, h, C.byref(p)))
        return bytes(h), p.value

    def links_window_open(self, peer_shard: int, handle: bytes) -> None:
        buf = (C.c_uint8 * abi.IPC_HANDLE_BYTES).from_buffer_copy(handle)
        self._check(self._lib.igs_links_window_open(self._h, peer_shard, buf))

    def links_window_attach(self, peer_shard: int, d_window: int) -> None:
        self._check(self._lib.igs_links_window_attach(self._h, peer_shard, d_window))

    def links_window_reset(self) -> None:
        self._check(self._lib.igs_links_window_reset(self._h))

    def links_pairs_peer(self, shard: int, n_shards: int) -> None:
        self._check(self._lib.igs_links_pairs_peer(self._h, shard, n_shards))

    def links_group_peer(self, shard: int, n_shards: int) -> None:
        self._check(self._lib.igs_links_group_peer(self._h, shard, n_shards))

    def links_finish_peer(self, shard: int) -> None:
        self._check(self._lib.igs_links_finish_peer(self._h, shard))

    def get_links(self, accounts: np.ndarray) -> np.ndarray:
        accounts = np.ascontiguousarray(accounts, dtype=np.uint32)
        out = np.zeros(accounts.shape[0], dtype=abi.links_row_dtype)
        self._check(self._lib.igs_get_links(self._h, abi.ptr(accounts), accounts.shape[0], abi.ptr(out)))
        return out

    def check_bonus_abuse(self, accounts: np.ndarray) -> np.ndarray:
        """RiskChecker.CheckBonusAbuse (bonus_engine.go:139-141) for a batch of dense account indices."""
        accounts = np.ascontiguousarray(accounts, dtype=np.uint32)
        out = np.zeros(accounts.shape[0], dtype=abi.abuse_dtype)
        self._check(self._lib.igs_check_bonus_abuse(self._h, abi.ptr(accounts), accounts.shape[0], abi.ptr(out)))
        return out

    # -- path A -------------------------------------------------------------------------------
    def score_features(self, cfg: abi.Config, features: np.ndarray, rows: np.ndarray,
                       want_leaves: bool = False, n_trees: int | None = None):
        features = np.ascontiguousarray(features, dtype=np.float32)
        rows = np.ascontiguousarray(rows, dtype=abi.rule_row_dtype)
        n = rows.shape[0]
        assert features.shape == (n, abi.NUM_FEATURES)
        out = np.zeros(n, dtype=abi.result_dtype)
        leaves = None
        if want_leaves:
            nt = n_trees if n_trees is not None else self._n_trees
            leaves = np.zeros((n, nt), dtype=np.uint8)
        self._check(self._lib.igs_score_features(self._h, C.byref(cfg), abi.ptr(features),
                                                 abi.ptr(rows), n, abi.ptr(out), abi.ptr(leaves)))
        return (out, leaves) if want_leaves else out

    def score_features_device(self, cfg: abi.Config, d_features: int, d_rows: int, n: int,
                              d_out: int, d_leaves: int = 0, stream: int = 0) -> None:
        self._check(self._lib.igs_score_features_device(self._h, C.byref(cfg), d_features, d_rows,
                                                        n, d_out, d_leaves, stream))

    # -- velocity -----------------------------------------------------------------------------
    def get_velocity(self, queries: np.ndarray) -> np.ndarray:
        queries = np.ascontiguousarray(queries, dtype=abi.query_dtype)
        out = np.zeros(queries.shape[0], dtype=abi.velocity_dtype)
        self._check(self._lib.igs_get_velocity(self._h, abi.ptr(queries), queries.shape[0],
                                               abi.ptr(out)))
        return out

    # -- path B -------------------------------------------------------------------------------
    def score_batch(self, cfg: abi.Config, tx: np.ndarray, want_features: bool = False,
                    want_velocity: bool = False, ids: np.ndarray | None = None):
        """igs_score_batch; with `ids` (tx_ids_dtype, parallel to tx) igs_score_batch_ids."""
        tx = np.ascontiguousarray(tx, dtype=abi.tx_dtype)
        n = tx.shape[0]
        out = np.zeros(n, dtype=abi.result_dtype)
        feats = np.zeros((n, abi.NUM_FEATURES), dtype=np.float32) if want_features else None
        vel = np.zeros(n, dtype=abi.velocity_dtype) if want_velocity else None
        if ids is not None:
            ids = np.ascontiguousarray(ids, dtype=abi.tx_ids_dtype)
            assert ids.shape == (n,)
            self._check(self._lib.igs_score_batch_ids(self._h, C.byref(cfg), abi.ptr(tx), abi.ptr(ids), n,
                                                      abi.ptr(out), abi.ptr(feats), abi.ptr(vel)))
        else:
            self._check(self._lib.igs_score_batch(self._h, C.byref(cfg), abi.ptr(tx), n, abi.ptr(out),
                                                  abi.ptr(feats), abi.ptr(vel)))
        if want_features or want_velocity:
            return out, feats, vel
        return out

    # -- shard partition ----------------------------------------------------------------------
    def partition_tx_device(self, d_tx: int, n: int, n_shards: int, d_tx_out: int, d_perm: int,
                            d_offsets: int, stream: int = 0) -> None:
        self._check(self._lib.igs_partition_tx_device(self._h, d_tx, n, n_shards, d_tx_out, d_perm,
                                                      d_offsets, stream))

    def partition_tx_static(self, d_tx: int, n: int, n_shards: int, d_tx_out: int, d_perm: int,
                            d_offsets: int, stream: int = 0) -> None:
        """igs_partition_tx_static: the partition for callers that must not allocate (an ingest
        thread sharing the GPU with in-flight scoring, or a fixed GPU memory budget).  Counters come
        from a 1 MiB array inside the library, so n_shards * ceil(n / 2048) must stay within
        abi.PARTITION_STATIC_WORDS; calls that overlap on one GPU race on that array."""
        self._check(self._lib.igs_partition_tx_static(self._h, d_tx, n, n_shards, d_tx_out, d_perm,
                                                      d_offsets, stream))

    # -- LTV ----------------------------------------------------------------------------------
    def ltv_batch(self, x_bf16: np.ndarray, rows: np.ndarray, want_raw: bool = False):
        x_bf16 = np.ascontiguousarray(x_bf16, dtype=np.uint16)
        rows = np.ascontiguousarray(rows, dtype=abi.ltv_row_dtype)
        n = rows.shape[0]
        assert x_bf16.shape == (n, abi.MLP_IN)
        out = np.zeros(n, dtype=abi.ltv_result_dtype)
        raw = np.zeros(n, dtype=np.float32) if want_raw else None
        self._check(self._lib.igs_ltv_batch(self._h, abi.ptr(x_bf16), abi.ptr(rows), n,
                                            abi.ptr(out), abi.ptr(raw)))
        return (out, raw) if want_raw else out

    def ltv_predict_batch(self, rows: np.ndarray, want_raw: bool = False):
        """LTVPredictor.BatchPredict (prediction/ltv.go:385-398) with the reference's own calculateLTV."""
        rows = np.ascontiguousarray(rows, dtype=abi.ltv_row_dtype)
        n = rows.shape[0]
        out = np.zeros(n, dtype=abi.ltv_result_dtype)
        raw = np.zeros(n, dtype=np.float64) if want_raw else None
        self._check(self._lib.igs_ltv_predict_batch(self._h, abi.ptr(rows), n, abi.ptr(out), abi.ptr(raw)))
        return (out, raw) if want_raw else out

    def ltv_mock_predict(self, feats: np.ndarray) -> np.ndarray:
        """ml.LTVModel.Predict (ml/onnx_model.go:405-432) per ml.LTVFeatures row."""
        feats = np.ascontiguousarray(feats, dtype=abi.ltv_features_dtype)
        out = np.zeros(feats.shape[0], dtype=np.float64)
        self._check(self._lib.igs_ltv_mock_predict(self._h, abi.ptr(feats), feats.shape[0], abi.ptr(out)))
        return out

    def ltv_hidden_device(self, d_x: int, n: int, layer: int, d_hidden: int, stream: int = 0) -> None:
        self._check(self._lib.igs_ltv_hidden_device(self._h, d_x, n, layer, d_hidden, stream))


def _serialised(fn):
    @functools.wraps(fn)
    def call(self, *a, **k):
        with self._lock:
            return fn(self, *a, **k)
    return call


# every method that uses slot 0's staging or replaces device state; the *_device methods are ordered by the
# caller's streams and stay lock-free
for _name in ("load_gbdt", "load_mlp", "load_players", "load_events", "append_events", "set_late_policy", "set_compaction_threshold", "prepare_ingest",
              "history_info", "evict_events", "get_uniques", "blacklist_load", "blacklist_add", "blacklist_check",
              "links_pairs", "links_group", "links_finish", "get_links", "check_bonus_abuse", "score_features",
              "get_velocity", "score_batch", "ltv_batch", "ltv_predict_batch", "ltv_mock_predict"):
    setattr(Engine, _name, _serialised(getattr(Engine, _name)))
del _name


def hash_id(s: str | bytes) -> int:
    """igs_hash_id: the 64-bit id of an identifier string ("" -> 0)."""
    b = s.encode() if isinstance(s, str) else bytes(s)
    return int(abi.load_library().igs_hash_id(b, len(b)))


from . import links, scoring, sharding  # noqa: E402  (reference-facing mirror + shard bookkeeping)


def __getattr__(name):
    # grpc / protobuf are only needed by the service shim: import it on first use
    if name == "grpc_service":
        import importlib
        return importlib.import_module(__name__ + ".grpc_service")
    raise AttributeError(name)
