"""gRPC shim for the scoring RPCs of `risk.v1.RiskService`.

The reference registers its handler with `riskv1.RegisterRiskServiceServer(grpcServer, NewRiskGRPCHandler(...))`
(services/risk/cmd/main.go:141-142) but ships neither the handler nor generated code.  This module is that
handler's body for the RPCs the engine can answer — `ScoreTransaction`, `ScoreBatch`, `GetThresholds`,
`UpdateThresholds` (proto/risk/v1/risk.proto:12-13,30-31), `AddToBlacklist`, `CheckBlacklist` (:23-24, on the
device-side sets), `CheckBonusAbuse` (:20, on the links table), `GetFeatures` (:27) and `PredictLTV`, `GetPlayerSegment` (:16-17,
on the LTV model) — speaking the same wire format.  There is no `protoc` in this image, so the message descriptors are built by hand from the .proto (same
package, message names, field names, numbers and types; risk.proto:38-257).  `GetFeatures` (:27) echoes
what extractFeatures would see for the account right now.  All ten RPCs of the service are answered.

Marshalling only: every score in a response was computed by the CUDA library through `scoring.ScoringEngine`.
"""
from __future__ import annotations

import time
import uuid as _uuid
from concurrent import futures
from typing import Optional

import numpy as np
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory, timestamp_pb2

from . import abi, scoring

PACKAGE = "risk.v1"
SERVICE = "risk.v1.RiskService"
_T = descriptor_pb2.FieldDescriptorProto

# (name, number, type[, type_name]) exactly as in proto/risk/v1/risk.proto
_SCORE_REQ = [("account_id", 1, _T.TYPE_STRING), ("player_id", 2, _T.TYPE_STRING), ("amount", 3, _T.TYPE_INT64),
              ("transaction_type", 4, _T.TYPE_STRING), ("currency", 5, _T.TYPE_STRING), ("game_id", 6, _T.TYPE_STRING),
              ("round_id", 7, _T.TYPE_STRING), ("ip_address", 8, _T.TYPE_STRING), ("device_id", 9, _T.TYPE_STRING),
              ("fingerprint", 10, _T.TYPE_STRING), ("user_agent", 11, _T.TYPE_STRING), ("session_id", 12, _T.TYPE_STRING)]
_FEATURES = [("tx_count_1m", 1, _T.TYPE_INT32), ("tx_count_5m", 2, _T.TYPE_INT32), ("tx_count_1h", 3, _T.TYPE_INT32),
             ("tx_sum_1h", 4, _T.TYPE_INT64), ("tx_avg_1h", 5, _T.TYPE_FLOAT), ("unique_devices_24h", 6, _T.TYPE_INT32),
             ("unique_ips_24h", 7, _T.TYPE_INT32), ("ip_country_changes_7d", 8, _T.TYPE_INT32),
             ("device_age_days", 9, _T.TYPE_INT32), ("account_age_days", 10, _T.TYPE_INT32),
             ("total_deposits", 11, _T.TYPE_INT64), ("total_withdrawals", 12, _T.TYPE_INT64),
             ("net_deposit", 13, _T.TYPE_INT64), ("deposit_count", 14, _T.TYPE_INT32), ("withdraw_count", 15, _T.TYPE_INT32),
             ("time_since_last_tx_sec", 16, _T.TYPE_INT32), ("session_duration_sec", 17, _T.TYPE_INT32),
             ("avg_bet_size", 18, _T.TYPE_FLOAT), ("win_rate", 19, _T.TYPE_FLOAT), ("is_vpn", 20, _T.TYPE_BOOL),
             ("is_proxy", 21, _T.TYPE_BOOL), ("is_tor", 22, _T.TYPE_BOOL), ("disposable_email", 23, _T.TYPE_BOOL),
             ("bonus_claim_count", 24, _T.TYPE_INT32), ("bonus_wager_completion_rate", 25, _T.TYPE_FLOAT),
             ("bonus_only_player", 26, _T.TYPE_BOOL)]


def _msg(fd, name, fields):
    m = fd.message_type.add()
    m.name = name
    for f in fields:
        x = m.field.add()
        x.name, x.number, x.type = f[0], f[1], f[2]
        x.label = f[4] if len(f) > 4 else _T.LABEL_OPTIONAL
        if len(f) > 3 and f[3]:
            x.type_name = f[3]
    return m


def build_file_descriptor() -> descriptor_pb2.FileDescriptorProto:
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "risk/v1/risk_scoring_subset.proto"
    fd.package = PACKAGE
    fd.syntax = "proto3"
    fd.dependency.append("google/protobuf/timestamp.proto")           # risk.proto:7
    seg = fd.enum_type.add()
    seg.name = "Segment"                                               # risk.proto:122-129
    for nm, num in (("SEGMENT_UNSPECIFIED", 0), ("SEGMENT_VIP", 1), ("SEGMENT_HIGH", 2), ("SEGMENT_MEDIUM", 3),
                    ("SEGMENT_LOW", 4), ("SEGMENT_CHURNING", 5)):
        v = seg.value.add()
        v.name, v.number = nm, num
    en = fd.enum_type.add()
    en.name = "Action"
    for nm, num in (("ACTION_UNSPECIFIED", 0), ("ACTION_APPROVE", 1), ("ACTION_REVIEW", 2), ("ACTION_BLOCK", 3)):
        v = en.value.add()
        v.name, v.number = nm, num
    req = _msg(fd, "ScoreTransactionRequest", _SCORE_REQ)
    entry = req.nested_type.add()                      # map<string, string> metadata = 13
    entry.name = "MetadataEntry"
    entry.options.map_entry = True
    for nm, num in (("key", 1), ("value", 2)):
        f = entry.field.add()
        f.name, f.number, f.type, f.label = nm, num, _T.TYPE_STRING, _T.LABEL_OPTIONAL
    md = req.field.add()
    md.name, md.number, md.type, md.label = "metadata", 13, _T.TYPE_MESSAGE, _T.LABEL_REPEATED
    md.type_name = f".{PACKAGE}.ScoreTransactionRequest.MetadataEntry"
    _msg(fd, "FeatureVector", _FEATURES)
    _msg(fd, "ScoreTransactionResponse", [
        ("score", 1, _T.TYPE_INT32), ("action", 2, _T.TYPE_ENUM, f".{PACKAGE}.Action"),
        ("reason_codes", 3, _T.TYPE_STRING, None, _T.LABEL_REPEATED), ("rule_score", 4, _T.TYPE_INT32),
        ("ml_score", 5, _T.TYPE_FLOAT), ("response_time_ms", 6, _T.TYPE_INT64),
        ("features", 7, _T.TYPE_MESSAGE, f".{PACKAGE}.FeatureVector")])
    _msg(fd, "ScoreBatchRequest", [("transactions", 1, _T.TYPE_MESSAGE, f".{PACKAGE}.ScoreTransactionRequest", _T.LABEL_REPEATED)])
    _msg(fd, "ScoreBatchResponse", [("results", 1, _T.TYPE_MESSAGE, f".{PACKAGE}.ScoreTransactionResponse", _T.LABEL_REPEATED)])
    _msg(fd, "UpdateThresholdsRequest", [("block_threshold", 1, _T.TYPE_INT32), ("review_threshold", 2, _T.TYPE_INT32)])
    _msg(fd, "UpdateThresholdsResponse", [("success", 1, _T.TYPE_BOOL), ("block_threshold", 2, _T.TYPE_INT32),
                                          ("review_threshold", 3, _T.TYPE_INT32)])
    _msg(fd, "GetThresholdsRequest", [])
    _msg(fd, "GetThresholdsResponse", [("block_threshold", 1, _T.TYPE_INT32), ("review_threshold", 2, _T.TYPE_INT32)])
    ts = ".google.protobuf.Timestamp"
    # blacklist management, risk.proto:148-182
    _msg(fd, "AddToBlacklistRequest", [("type", 1, _T.TYPE_STRING), ("value", 2, _T.TYPE_STRING), ("reason", 3, _T.TYPE_STRING),
                                       ("created_by", 4, _T.TYPE_STRING), ("expires_at", 5, _T.TYPE_MESSAGE, ts)])
    _msg(fd, "AddToBlacklistResponse", [("success", 1, _T.TYPE_BOOL), ("id", 2, _T.TYPE_STRING)])
    _msg(fd, "CheckBlacklistRequest", [("device_id", 1, _T.TYPE_STRING), ("fingerprint", 2, _T.TYPE_STRING),
                                       ("ip_address", 3, _T.TYPE_STRING), ("email", 4, _T.TYPE_STRING)])
    _msg(fd, "BlacklistMatch", [("type", 1, _T.TYPE_STRING), ("value", 2, _T.TYPE_STRING), ("reason", 3, _T.TYPE_STRING),
                                ("created_at", 4, _T.TYPE_MESSAGE, ts)])
    _msg(fd, "CheckBlacklistResponse", [("is_blacklisted", 1, _T.TYPE_BOOL),
                                        ("matches", 2, _T.TYPE_MESSAGE, f".{PACKAGE}.BlacklistMatch", _T.LABEL_REPEATED)])
    # feature inspection, risk.proto:188-195
    _msg(fd, "GetFeaturesRequest", [("account_id", 1, _T.TYPE_STRING)])
    _msg(fd, "GetFeaturesResponse", [("account_id", 1, _T.TYPE_STRING), ("features", 2, _T.TYPE_MESSAGE, f".{PACKAGE}.FeatureVector"),
                                     ("computed_at", 3, _T.TYPE_MESSAGE, ts)])
    # bonus abuse, risk.proto:135-145
    _msg(fd, "CheckBonusAbuseRequest", [("account_id", 1, _T.TYPE_STRING), ("bonus_id", 2, _T.TYPE_STRING)])
    _msg(fd, "CheckBonusAbuseResponse", [("is_abuser", 1, _T.TYPE_BOOL), ("abuse_score", 2, _T.TYPE_FLOAT),
                                         ("signals", 3, _T.TYPE_STRING, None, _T.LABEL_REPEATED),
                                         ("linked_accounts", 4, _T.TYPE_STRING, None, _T.LABEL_REPEATED)])
    # LTV & segmentation, risk.proto:94-120
    _msg(fd, "PredictLTVRequest", [("account_id", 1, _T.TYPE_STRING)])
    _msg(fd, "PredictLTVResponse", [("account_id", 1, _T.TYPE_STRING), ("predicted_ltv", 2, _T.TYPE_FLOAT),
                                    ("segment", 3, _T.TYPE_ENUM, f".{PACKAGE}.Segment"), ("churn_risk", 4, _T.TYPE_FLOAT),
                                    ("predicted_active_days", 5, _T.TYPE_INT32), ("confidence", 6, _T.TYPE_FLOAT),
                                    ("next_best_action", 7, _T.TYPE_STRING), ("predicted_at", 8, _T.TYPE_MESSAGE, ts)])
    _msg(fd, "GetPlayerSegmentRequest", [("account_id", 1, _T.TYPE_STRING)])
    _msg(fd, "GetPlayerSegmentResponse", [("account_id", 1, _T.TYPE_STRING), ("segment", 2, _T.TYPE_ENUM, f".{PACKAGE}.Segment"),
                                          ("ltv", 3, _T.TYPE_FLOAT), ("churn_risk", 4, _T.TYPE_FLOAT),
                                          ("recommended_actions", 5, _T.TYPE_STRING, None, _T.LABEL_REPEATED)])
    return fd


_pool = descriptor_pool.DescriptorPool()
_pool.AddSerializedFile(timestamp_pb2.DESCRIPTOR.serialized_pb)
_pool.Add(build_file_descriptor())


def message_class(name: str):
    return message_factory.GetMessageClass(_pool.FindMessageTypeByName(f"{PACKAGE}.{name}"))


ScoreTransactionRequest = message_class("ScoreTransactionRequest")
ScoreTransactionResponse = message_class("ScoreTransactionResponse")
ScoreBatchRequest = message_class("ScoreBatchRequest")
ScoreBatchResponse = message_class("ScoreBatchResponse")
FeatureVector = message_class("FeatureVector")
UpdateThresholdsRequest = message_class("UpdateThresholdsRequest")
UpdateThresholdsResponse = message_class("UpdateThresholdsResponse")
GetThresholdsRequest = message_class("GetThresholdsRequest")
GetThresholdsResponse = message_class("GetThresholdsResponse")
AddToBlacklistRequest = message_class("AddToBlacklistRequest")
AddToBlacklistResponse = message_class("AddToBlacklistResponse")
CheckBlacklistRequest = message_class("CheckBlacklistRequest")
CheckBlacklistResponse = message_class("CheckBlacklistResponse")
BlacklistMatch = message_class("BlacklistMatch")
GetFeaturesRequest = message_class("GetFeaturesRequest")
GetFeaturesResponse = message_class("GetFeaturesResponse")
CheckBonusAbuseRequest = message_class("CheckBonusAbuseRequest")
CheckBonusAbuseResponse = message_class("CheckBonusAbuseResponse")
PredictLTVRequest = message_class("PredictLTVRequest")
PredictLTVResponse = message_class("PredictLTVResponse")
GetPlayerSegmentRequest = message_class("GetPlayerSegmentRequest")
GetPlayerSegmentResponse = message_class("GetPlayerSegmentResponse")

_METHODS = {
    "ScoreTransaction": (ScoreTransactionRequest, ScoreTransactionResponse),
    "ScoreBatch": (ScoreBatchRequest, ScoreBatchResponse),
    "UpdateThresholds": (UpdateThresholdsRequest, UpdateThresholdsResponse),
    "GetThresholds": (GetThresholdsRequest, GetThresholdsResponse),
    "AddToBlacklist": (AddToBlacklistRequest, AddToBlacklistResponse),
    "CheckBlacklist": (CheckBlacklistRequest, CheckBlacklistResponse),
    "CheckBonusAbuse": (CheckBonusAbuseRequest, CheckBonusAbuseResponse),
    "GetFeatures": (GetFeaturesRequest, GetFeaturesResponse),
    "PredictLTV": (PredictLTVRequest, PredictLTVResponse),
    "GetPlayerSegment": (GetPlayerSegmentRequest, GetPlayerSegmentResponse),
}
_UNIMPLEMENTED = []
_ACTION_ENUM = {"approve": 1, "review": 2, "block": 3}


def to_score_request(m) -> scoring.ScoreRequest:
    """risk.v1.ScoreTransactionRequest -> scoring.ScoreRequest (the wire message carries no timestamp:
    like the reference, scoring happens at wall-clock now, redis_store.go:62)."""
    try:
        acc = _uuid.UUID(m.account_id)
    except (ValueError, AttributeError, TypeError):
        acc = _uuid.UUID(int=0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFF)     # unknown account -> zero features
    return scoring.ScoreRequest(account_id=acc, amount=m.amount, tx_type=m.transaction_type, player_id=m.player_id,
                                currency=m.currency, game_id=m.game_id, ip=m.ip_address, device_id=m.device_id,
                                fingerprint=m.fingerprint, user_agent=m.user_agent, session_id=m.session_id)


def fill_features(fv, tx_row, vel, player) -> None:
    """The 26-field echo (risk.proto:197-235 = scoring.FeatureVector, engine.go:67-105) from what the
    engine looked up: its velocity block, the host copy of the player row and the request's own bits."""
    fv.tx_count_1m, fv.tx_count_5m, fv.tx_count_1h = int(vel["tx_count_1m"]), int(vel["tx_count_5m"]), int(vel["tx_count_1h"])
    fv.tx_sum_1h = int(vel["tx_sum_1h"])
    if vel["tx_count_1h"] > 0:
        fv.tx_avg_1h = float(np.float32(float(vel["tx_sum_1h"]) / float(vel["tx_count_1h"])))
    fv.unique_devices_24h, fv.unique_ips_24h = int(tx_row["unique_devices_24h"]), int(tx_row["unique_ips_24h"])
    fv.time_since_last_tx_sec, fv.session_duration_sec = int(vel["time_since_last_tx"]), int(vel["session_duration"])
    bits = int(tx_row["bits"])
    fv.is_vpn, fv.is_proxy, fv.is_tor = bool(bits & abi.BIT_VPN), bool(bits & abi.BIT_PROXY), bool(bits & abi.BIT_TOR)
    if player is not None:
        fv.account_age_days = int((int(tx_row["ts"]) - int(player["account_created_at"])) // 86400)
        fv.total_deposits, fv.total_withdrawals = int(player["total_deposits"]), int(player["total_withdrawals"])
        fv.net_deposit = int(player["total_deposits"]) - int(player["total_withdrawals"])
        fv.deposit_count, fv.withdraw_count = int(player["deposit_count"]), int(player["withdraw_count"])
        fv.avg_bet_size = float(np.float32(player["avg_bet_size"]))
        if player["bet_count"] > 0:
            fv.win_rate = float(np.float32(float(player["win_count"]) / float(player["bet_count"])))
        fv.bonus_claim_count = int(player["bonus_claim_count"])
        fv.bonus_wager_completion_rate = float(np.float32(player["bonus_wager_complete"]))
        fv.bonus_only_player = bool(player["bonus_claim_count"] > 3 and player["total_deposits"] < 5000)


class RiskServicer:
    """Handler body for the scoring RPCs.  `players` (optional host copy of the BatchFeatures table)
    enables the 26-field `features` echo of ScoreTransactionResponse; `ltv_inputs` (optional host copy of
    the per-player LTV model input: n x 256 bf16 bit patterns + igs_ltv_row sideband, indexed like the
    directory) enables PredictLTV / GetPlayerSegment."""

    def __init__(self, engine: scoring.ScoringEngine, players: Optional[np.ndarray] = None, ltv_inputs=None):
        self.engine = engine
        self.players = players
        self.ltv_inputs = ltv_inputs

    def _score(self, reqs, now: Optional[int] = None):
        t0 = time.perf_counter()
        sreqs = [to_score_request(m) for m in reqs]
        tx = self.engine.marshal(sreqs, now=now)
        if self.engine.device_identifiers:      # rule 8 on the device; echo the uniques the engine used
            out, _, vel = self.engine.engine.score_batch(self.engine.config, tx, want_velocity=True,
                                                         ids=self.engine.marshal_ids(sreqs))
            try:
                uq = self.engine.engine.get_uniques(scoring.queries_of(tx))
                tx["unique_devices_24h"], tx["unique_ips_24h"] = uq["devices_24h"], uq["ips_24h"]
            except Exception:                   # history without identifiers: the tx's own counts were used
                pass
        else:
            out, _, vel = self.engine.engine.score_batch(self.engine.config, tx, want_velocity=True)
        res = self.engine.unmarshal(out, int((time.perf_counter() - t0) * 1000))
        msgs = []
        for i, r in enumerate(res):
            m = ScoreTransactionResponse(score=r.score, action=_ACTION_ENUM[r.action], reason_codes=r.reason_codes,
                                         rule_score=r.rule_score, ml_score=r.ml_score,
                                         response_time_ms=r.response_time_ms)
            acc = int(tx["account"][i])
            player = self.players[acc] if self.players is not None and acc < self.players.shape[0] else None
            fill_features(m.features, tx[i], vel[i], player)
            msgs.append(m)
        return msgs

    def ScoreTransaction(self, request, context):
        return self._score([request])[0]

    def ScoreBatch(self, request, context):
        return ScoreBatchResponse(results=self._score(list(request.transactions)))

    def GetThresholds(self, request, context):
        b, r = self.engine.GetThresholds()
        return GetThresholdsResponse(block_threshold=b, review_threshold=r)

    def UpdateThresholds(self, request, context):
        self.engine.SetThresholds(request.block_threshold, request.review_threshold)
        b, r = self.engine.GetThresholds()
        return UpdateThresholdsResponse(success=True, block_threshold=b, review_threshold=r)


    # -- blacklist management (risk.proto:23-24) on the device-side sets -----------------------------
    def AddToBlacklist(self, request, context):
        try:
            self.engine.AddToBlacklist(request.type, request.value)
        except ValueError:                       # "unknown blacklist type" (redis_store.go:261), e.g. email
            return AddToBlacklistResponse(success=False, id="")
        return AddToBlacklistResponse(success=True, id=f"{request.type}:{self.engine._hash(request.value):016x}")

    def CheckBlacklist(self, request, context):
        h = self.engine._hash
        ids = np.zeros(3, dtype=abi.tx_ids_dtype)            # one single-identifier probe per list type
        ids["device"][0], ids["fingerprint"][1], ids["ip"][2] = h(request.device_id), h(request.fingerprint), h(request.ip_address)
        hit = self.engine.engine.blacklist_check(ids)
        resp = CheckBlacklistResponse(is_blacklisted=bool(hit.any()))
        for k, (typ, val) in enumerate((("device", request.device_id), ("fingerprint", request.fingerprint),
                                        ("ip", request.ip_address))):
            if hit[k]:
                resp.matches.add(type=typ, value=val)
        return resp

    # -- feature inspection (risk.proto:27): extractFeatures for the account at wall-clock now, no tx ---
    def GetFeatures(self, request, context):
        probe = ScoreTransactionRequest(account_id=request.account_id)          # amount 0, no identifiers
        resp = GetFeaturesResponse(account_id=request.account_id)
        resp.features.CopyFrom(self._score([probe])[0].features)
        resp.computed_at.GetCurrentTime()
        return resp

    # -- bonus abuse (risk.proto:20) on the links table + player rows -------------------------------
    def CheckBonusAbuse(self, request, context):
        try:
            acc = _uuid.UUID(request.account_id)
        except ValueError:
            acc = _uuid.UUID(int=0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFF)
        is_abuser, score, signals, linked = self.engine.CheckBonusAbuse(acc)
        return CheckBonusAbuseResponse(is_abuser=is_abuser, abuse_score=score, signals=signals, linked_accounts=linked)

    # -- LTV & segmentation (risk.proto:16-17) on the LTV model --------------------------------------
    def _ltv(self, account_id: str, context):
        import grpc
        if self.ltv_inputs is None:
            context.abort(grpc.StatusCode.FAILED_PRECONDITION, "no LTV inputs loaded")
        try:
            idx = self.engine.directory.get(_uuid.UUID(account_id))
        except ValueError:
            idx = 0xFFFFFFFE
        x, rows = self.ltv_inputs
        if idx >= rows.shape[0]:
            context.abort(grpc.StatusCode.NOT_FOUND, "unknown account")       # GetPlayerFeatures error, ltv.go:115-118
        return self.engine.engine.ltv_batch(x[idx:idx + 1], rows[idx:idx + 1])[0]

    def PredictLTV(self, request, context):
        r = self._ltv(request.account_id, context)
        lib = abi.load_library()
        resp = PredictLTVResponse(account_id=request.account_id, predicted_ltv=float(r["predicted_ltv"]),
                                  segment=int(r["segment"]), churn_risk=float(r["churn_risk"]),
                                  predicted_active_days=int(r["days_survival"]), confidence=float(r["confidence"]),
                                  next_best_action=lib.igs_ltv_action_name(int(r["next_best_action"])).decode())
        resp.predicted_at.GetCurrentTime()
        return resp

    def GetPlayerSegment(self, request, context):
        r = self._ltv(request.account_id, context)
        lib = abi.load_library()
        return GetPlayerSegmentResponse(account_id=request.account_id, segment=int(r["segment"]),
                                        ltv=float(r["predicted_ltv"]), churn_risk=float(r["churn_risk"]),
                                        recommended_actions=[lib.igs_ltv_action_name(int(r["next_best_action"])).decode()])


def add_to_server(servicer: RiskServicer, server) -> None:
    """RegisterRiskServiceServer: generic handlers, no generated code needed."""
    import grpc

    handlers = {}
    for name, (req_cls, resp_cls) in _METHODS.items():
        handlers[name] = grpc.unary_unary_rpc_method_handler(
            getattr(servicer, name), request_deserializer=req_cls.FromString,
            response_serializer=lambda m: m.SerializeToString())

    def _unimplemented(request, context):
        context.abort(grpc.StatusCode.UNIMPLEMENTED, "not on the scoring path of this shim")

    for name in _UNIMPLEMENTED:
        handlers[name] = grpc.unary_unary_rpc_method_handler(_unimplemented, request_deserializer=lambda b: b,
                                                             response_serializer=lambda m: b"")
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(SERVICE, handlers),))


def serve(servicer: RiskServicer, address: str = "127.0.0.1:9082", max_workers: int = 4):
    """grpc.Server on the reference's risk port (main.go:57, docker-compose.yml:163-186); returns it started."""
    import grpc

    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers),
                         options=[("grpc.max_receive_message_length", 256 << 20),
                                  ("grpc.max_send_message_length", 256 << 20)])
    add_to_server(servicer, server)
    port = server.add_insecure_port(address)
    server.start()
    server.bound_port = port
    return server


class RiskClient:
    """Minimal client stub (what wallet's RiskService interface needs, wallet_service.go:40-42)."""

    def __init__(self, address: str):
        import grpc

        self.channel = grpc.insecure_channel(address, options=[("grpc.max_receive_message_length", 256 << 20),
                                                              ("grpc.max_send_message_length", 256 << 20)])
        self._calls = {}
        for name, (req_cls, resp_cls) in _METHODS.items():
            self._calls[name] = self.channel.unary_unary(f"/{SERVICE}/{name}",
                                                         request_serializer=lambda m: m.SerializeToString(),
                                                         response_deserializer=resp_cls.FromString)

    def __getattr__(self, name):
        try:
            return self.__dict__["_calls"][name]
        except KeyError:
            raise AttributeError(name)

    def close(self):
        self.channel.close()
