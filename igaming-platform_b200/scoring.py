"""Host-side mirror of the reference's operator interface for the scoring path.

Same names, argument meaning and error behaviour as `services/risk/internal/scoring/engine.go`
(`ScoreRequest` :40-53, `ScoreResponse` :56-64, `ScoringEngine.Score` :262-323, `UpdateFeatures`
:486-488, `SetThresholds` / `GetThresholds` :491-504, `ScoreWithExplanation` :507-543), but a batch
goes to the GPU in one C-ABI call (`igs_score_batch` / `igs_score_batch_ids`).  This is marshalling only — requests to `igs_tx` records, result records
to responses; every number in a response was computed by the CUDA library.
"""
from __future__ import annotations

import ctypes as C
import time
import uuid as _uuid
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np

from . import abi

ActionApprove, ActionReview, ActionBlock = "approve", "review", "block"      # engine.go:33-37
_TX_TYPES = {"bet": abi.TX_BET, "deposit": abi.TX_DEPOSIT, "withdraw": abi.TX_WITHDRAW}


@dataclass
class ScoreRequest:
    """scoring.ScoreRequest, engine.go:40-53.  `timestamp` (unix seconds) plays the role of
    time.Now() in extractFeatures (:350,:374); 0 = wall clock at call time."""
    account_id: _uuid.UUID
    amount: int = 0
    tx_type: str = "bet"
    player_id: str = ""
    currency: str = ""
    game_id: str = ""
    ip: str = ""
    device_id: str = ""
    fingerprint: str = ""
    user_agent: str = ""
    session_id: str = ""
    timestamp: int = 0


@dataclass
class TransactionEvent:
    """scoring.TransactionEvent (= features.TransactionEvent, features/redis_store.go:49-56)."""
    account_id: _uuid.UUID
    amount: int = 0
    tx_type: str = "bet"
    ip: str = ""
    device_id: str = ""
    timestamp: int = 0


@dataclass
class ScoreResponse:
    """scoring.ScoreResponse, engine.go:56-64."""
    score: int
    action: str
    reason_codes: List[str]
    rule_score: int
    ml_score: float
    response_time_ms: int
    features: Optional[np.ndarray] = None       # the 64-float model input, when requested


class PlayerDirectory:
    """account uuid -> dense index of this shard (host-side; the device tables are dense)."""

    def __init__(self, n_shards: int = 1, shard: int = 0):
        self._idx: Dict[_uuid.UUID, int] = {}
        self.n_shards, self.shard = n_shards, shard

    def shard_of(self, account_id: _uuid.UUID) -> int:
        v = account_id.int
        return abi.load_library().igs_shard_of(v & 0xFFFFFFFFFFFFFFFF, v >> 64, self.n_shards)

    def add(self, account_id: _uuid.UUID) -> int:
        return self._idx.setdefault(account_id, len(self._idx))

    def get(self, account_id: _uuid.UUID) -> int:
        # unknown account: an index beyond every table -> the engine scores it with zero batch /
        # velocity features, as the reference does when the feature store errors (engine.go:338-366)
        return self._idx.get(account_id, 0xFFFFFFFE)

    def uuid_of(self, index: int) -> Optional[_uuid.UUID]:
        """Reverse lookup (dense index -> account uuid) for reporting linked accounts."""
        if not hasattr(self, "_rev") or len(self._rev) != len(self._idx):
            self._rev = {i: u for u, i in self._idx.items()}
        return self._rev.get(index)

    def __len__(self) -> int:
        return len(self._idx)


def queries_of(tx: np.ndarray) -> np.ndarray:
    """The igs_query records of a batch of igs_tx records."""
    q = np.zeros(tx.shape[0], dtype=abi.query_dtype)
    q["account"], q["ts"], q["seq_limit"] = tx["account"], tx["ts"], tx["seq_limit"]
    return q


IPIntel = Callable[[str], Tuple[bool, bool, bool]]          # ip -> (vpn, proxy, tor)   engine.go:158-171
Blacklist = Callable[[str, str, str], bool]                  # (device, fingerprint, ip) engine.go:174-176
Uniques = Callable[[_uuid.UUID], Tuple[int, int]]            # account -> (devices24h, ips24h) redis_store.go:80-81


class ScoringEngine:
    """scoring.ScoringEngine over one `Engine` (one GPU).  `ip_intel`, `blacklist` and `uniques`
    may be None, exactly as the reference tolerates nil (engine.go:394, :469).

    `device_identifiers=True` moves the identifier work to the GPU: requests and events carry
    `igs_hash_id` of their device / fingerprint / ip strings, rule 8 is evaluated against the
    engine's blacklist sets (`AddToBlacklist`) and — when the history was loaded with identifiers —
    the unique device / IP counts come from the event index.  The host-side `blacklist` / `uniques`
    callables are then optional extras (a host verdict still ORs into rule 8)."""

    def __init__(self, engine, directory: PlayerDirectory, config: Optional[abi.Config] = None,
                 ip_intel: Optional[IPIntel] = None, blacklist: Optional[Blacklist] = None,
                 uniques: Optional[Uniques] = None, device_identifiers: bool = False):
        from . import default_config
        self.engine = engine
        self.directory = directory
        self.config = config if config is not None else default_config()
        self.ip_intel, self.blacklist, self.uniques = ip_intel, blacklist, uniques
        self.device_identifiers = device_identifiers
        self._lib = abi.load_library()

    def _hash(self, s: str) -> int:
        b = s.encode()
        return int(self._lib.igs_hash_id(b, len(b)))

    def marshal_ids(self, reqs: Sequence[ScoreRequest]) -> np.ndarray:
        """The identifier triples of a batch (igs_tx_ids); "" -> 0 = not checked (redis_store.go:271-279)."""
        ids = np.zeros(len(reqs), dtype=abi.tx_ids_dtype)
        for i, r in enumerate(reqs):
            ids[i] = (self._hash(r.device_id), self._hash(r.fingerprint), self._hash(r.ip))
        return ids

    def AddToBlacklist(self, list_type: str, value: str) -> None:
        """RedisFeatureStore.AddToBlacklist (features/redis_store.go:251-264) on the device sets."""
        kinds = {"device": abi.BL_DEVICE, "ip": abi.BL_IP, "fingerprint": abi.BL_FINGERPRINT}
        if list_type not in kinds:
            raise ValueError(f"unknown blacklist type: {list_type}")           # :261
        self.engine.blacklist_add(kinds[list_type], np.array([self._hash(value)], np.uint64))

    def UpdateFeatures(self, events: Sequence[TransactionEvent]) -> None:
        """ScoringEngine.UpdateFeatures (engine.go:486-488) for a micro-batch of completed transactions:
        one igs_append_events(_ids) call.  Events of accounts the directory does not know are dropped
        (the device tables are dense)."""
        from . import Events
        known = [e for e in events if self.directory.get(e.account_id) != 0xFFFFFFFE]
        now = int(time.time())
        ev = Events(np.array([self.directory.get(e.account_id) for e in known], np.uint32),
                    np.array([e.timestamp or now for e in known], np.uint32),
                    np.array([e.amount for e in known], np.int64),
                    np.array([_TX_TYPES.get(e.tx_type, abi.TX_OTHER) for e in known], np.uint8))
        if self.device_identifiers:
            ev.device = np.array([self._hash(e.device_id) for e in known], np.uint64)
            ev.ip = np.array([self._hash(e.ip) for e in known], np.uint64)
        self.engine.append_events(ev)

    def CheckBonusAbuse(self, account_id: _uuid.UUID) -> Tuple[bool, float, List[str], List[str]]:
        """RiskChecker.CheckBonusAbuse (services/bonus/internal/service/bonus_engine.go:139-141) with the
        detail of rpc CheckBonusAbuse (risk.proto:140-145): (is_abuser, abuse_score, signals, linked accounts).
        Linked accounts of this shard are reported by uuid, those of other shards as "shard:index"."""
        idx = np.array([self.directory.get(account_id)], np.uint32)
        a = self.engine.check_bonus_abuse(idx)[0]
        row = self.engine.get_links(idx)[0]
        names = (C.c_char_p * 3)()
        k = self._lib.igs_abuse_signals(int(a["signals"]), C.byref(names))
        linked = []
        for g in row["linked"][:min(int(row["n_linked"]), 3)]:
            shard, i = int(g) >> 32, int(g) & 0xFFFFFFFF
            u = self.directory.uuid_of(i) if shard == self.directory.shard else None
            linked.append(str(u) if u is not None else f"{shard}:{i}")
        return bool(a["is_abuser"]), float(a["abuse_score"]), [names[j].decode() for j in range(k)], linked

    # -- marshalling (pure host logic; unit-tested without a GPU) ---------------------------------
    def marshal(self, reqs: Sequence[ScoreRequest], now: Optional[int] = None) -> np.ndarray:
        now = int(time.time()) if now is None else now
        tx = np.zeros(len(reqs), dtype=abi.tx_dtype)
        for i, r in enumerate(reqs):
            bits = abi.BIT_VALID
            if self.ip_intel is not None and r.ip:                          # engine.go:394-405
                vpn, proxy, tor = self.ip_intel(r.ip)
                bits |= (abi.BIT_VPN if vpn else 0) | (abi.BIT_PROXY if proxy else 0) | (abi.BIT_TOR if tor else 0)
            if self.blacklist is not None and self.blacklist(r.device_id, r.fingerprint, r.ip):  # :469-475
                bits |= abi.BIT_BLACKLISTED
            dev, ips = self.uniques(r.account_id) if self.uniques is not None else (0, 0)
            tx[i] = (r.amount, self.directory.get(r.account_id), r.timestamp or now, abi.SEQ_ALL,
                     min(dev, 0xFFFF), min(ips, 0xFFFF), _TX_TYPES.get(r.tx_type, abi.TX_OTHER), bits,
                     np.zeros(6, np.uint8))
        return tx

    def unmarshal(self, out: np.ndarray, elapsed_ms: int, feats: Optional[np.ndarray] = None) -> List[ScoreResponse]:
        names = (C.c_char_p * abi.NUM_REASONS)()
        res = []
        for i in range(out.shape[0]):
            flags = int(out["flags"][i])
            k = self._lib.igs_reason_codes(flags, C.byref(names))
            res.append(ScoreResponse(
                score=int(out["score"][i]),
                action=self._lib.igs_action_name(int(out["action"][i])).decode(),
                reason_codes=[names[j].decode() for j in range(k)],
                rule_score=int(self._lib.igs_rule_score_from_flags(C.byref(self.config), flags)),
                ml_score=float(out["ml_score"][i]),
                response_time_ms=elapsed_ms,
                features=None if feats is None else feats[i]))
        return res

    # -- the reference's methods ------------------------------------------------------------------
    def ScoreBatch(self, reqs: Sequence[ScoreRequest], with_features: bool = False) -> List[ScoreResponse]:
        """rpc ScoreBatch (risk.proto:13): never raises for data reasons (engine.go:314-322)."""
        t0 = time.perf_counter()
        tx = self.marshal(reqs)
        ids = self.marshal_ids(reqs) if self.device_identifiers else None
        if with_features:
            out, feats, _ = self.engine.score_batch(self.config, tx, want_features=True, ids=ids)
        else:
            out, feats = self.engine.score_batch(self.config, tx, ids=ids), None
        return self.unmarshal(out, int((time.perf_counter() - t0) * 1000), feats)

    def Score(self, req: ScoreRequest) -> ScoreResponse:
        return self.ScoreBatch([req], with_features=True)[0]

    def GetThresholds(self) -> Tuple[int, int]:
        return self.config.block_threshold, self.config.review_threshold

    def SetThresholds(self, block: int, review: int) -> None:
        self.config.block_threshold, self.config.review_threshold = block, review

    def ScoreWithExplanation(self, req: ScoreRequest) -> str:
        r = self.Score(req)
        lines = ["", "Fraud Score Analysis", "====================",
                 f"Final Score: {r.score}/100", f"Action: {r.action}",
                 f"Response Time: {r.response_time_ms}ms", "",
                 f"Rule Contribution: {r.rule_score}",
                 f"ML Contribution: {r.ml_score:.2f} ({r.ml_score:.2f} * 100)", "", "Triggered Rules:"]
        order = ["HIGH_VELOCITY", "NEW_ACCOUNT_LARGE_TX", "IP_COUNTRY_MISMATCH", "MULTIPLE_DEVICES",
                 "SUSPICIOUS_PATTERN", "VPN_DETECTED", "KNOWN_FRAUDSTER", "RAPID_DEPOSIT_WITHDRAW",
                 "BONUS_ABUSE", "ML_HIGH_RISK"]
        for code in r.reason_codes:
            lines.append(f"  - {code} (+{self.config.reason_weight[order.index(code)]})")
        return "\n".join(lines) + "\n"
