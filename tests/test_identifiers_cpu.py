"""Oracle-side checks of the identifier features: exact unique devices / IPs per
account (features/redis_store.go:80-81,141-152) and the blacklist sets (:245-293).  CPU only."""
import ctypes as C
import json
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def load(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def test_kat_uniques_redis_emulation(orc):
    """Command-level emulation (PFADD + EXPIRE 24h with exact sets) reproduces KAT set D."""
    kat = load("kat_uniques.json")
    L = orc.lib()
    acc = L.orc_redis_new()
    try:
        for r in kat["rows"]:
            now = kat["t0"] + r["dt"]
            rt = orc.Realtime()
            L.orc_redis_get(acc, now, C.byref(rt))
            assert (rt.unique_devices_24h, rt.unique_ips_24h) == (r["devices"], r["ips"]), f"dt={r['dt']}: {r['why']}"
            L.orc_redis_update_ids(acc, now, 100, orc.hash_id(r["device"]), orc.hash_id(r["ip"]))
    finally:
        L.orc_redis_free(acc)


def test_kat_uniques_closed_form(orc, pkg):
    """The closed form the device index implements reproduces KAT set D through the batch API."""
    abi = pkg.abi
    kat = load("kat_uniques.json")
    rows = kat["rows"]
    n = len(rows)
    ev = pkg.Events(np.zeros(n, np.uint32), np.array([kat["t0"] + r["dt"] for r in rows], np.uint32),
                    np.full(n, 100, np.int64), np.zeros(n, np.uint8),
                    device=np.array([orc.hash_id(r["device"]) for r in rows], np.uint64),
                    ip=np.array([orc.hash_id(r["ip"]) for r in rows], np.uint64))
    q = np.zeros(n, dtype=abi.query_dtype)
    q["ts"] = ev.ts
    q["seq_limit"] = np.arange(n)
    u = orc.uniques_batch(ev, 1, q)
    for i, r in enumerate(rows):
        assert (int(u[i]["devices_24h"]), int(u[i]["ips_24h"])) == (r["devices"], r["ips"]), f"dt={r['dt']}"


def test_uniques_closed_form_equals_emulation_on_random_streams(orc, pkg):
    """Property: random single-account streams with gaps around the 24h TTL, empty identifiers and
    repeated identifiers — closed form == command-level emulation, for queries at and between events."""
    abi = pkg.abi
    L = orc.lib()
    rng = np.random.default_rng(777)
    gaps_pool = np.array([0, 0, 1, 5, 60, 3600, 43200, 86399, 86400, 86401, 90000, 172800])
    for trial in range(60):
        n = int(rng.integers(1, 80))
        ts = 1_700_000_000 + np.cumsum(rng.choice(gaps_pool, size=n))
        dev = rng.integers(0, 5, size=n).astype(np.uint64)        # 0 = empty
        ip = rng.integers(0, 4, size=n).astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)
        ev = pkg.Events(np.zeros(n, np.uint32), ts.astype(np.uint32), np.full(n, 7, np.int64),
                        np.zeros(n, np.uint8), device=dev, ip=ip)
        later = rng.choice(np.array([0, 1, 86399, 86400, 86401]), size=n)
        q = np.zeros(2 * n, dtype=abi.query_dtype)
        q["ts"][:n] = ev.ts
        q["seq_limit"][:n] = np.arange(n)
        q["ts"][n:] = ts + later                                   # a query some time after event i
        q["seq_limit"][n:] = abi.SEQ_ALL
        u = orc.uniques_batch(ev, 1, q)
        # emulation: replay, answering the queries in time order
        acc = L.orc_redis_new()
        try:
            for i in range(n):
                rt = orc.Realtime()
                L.orc_redis_get(acc, int(ts[i]), C.byref(rt))
                assert (rt.unique_devices_24h, rt.unique_ips_24h) == (u[i]["devices_24h"], u[i]["ips_24h"])
                L.orc_redis_update_ids(acc, int(ts[i]), 7, int(dev[i]), int(ip[i]))
        finally:
            L.orc_redis_free(acc)
        for i in range(n):          # "later" queries see every event with ts <= query ts: replay that prefix
            tq = int(ts[i] + later[i])
            acc = L.orc_redis_new()
            try:
                for j in range(n):
                    if ts[j] <= tq:
                        L.orc_redis_update_ids(acc, int(ts[j]), 7, int(dev[j]), int(ip[j]))
                rt = orc.Realtime()
                L.orc_redis_get(acc, tq, C.byref(rt))
                assert (rt.unique_devices_24h, rt.unique_ips_24h) == (u[n + i]["devices_24h"], u[n + i]["ips_24h"]), \
                    (trial, i)
            finally:
                L.orc_redis_free(acc)


def test_hash_id_spec(orc, pkg):
    """igs_hash_id (host function of libigs.so, no GPU needed) equals the oracle's restatement of the
    spec; "" -> 0; known FNV-1a + finaliser value for one string computed independently here."""
    def spec(b: bytes) -> int:
        if not b:
            return 0
        M = (1 << 64) - 1
        h = 0xcbf29ce484222325
        for c in b:
            h = ((h ^ c) * 0x100000001b3) & M
        h ^= h >> 30; h = (h * 0xbf58476d1ce4e5b9) & M
        h ^= h >> 27; h = (h * 0x94d049bb133111eb) & M
        h ^= h >> 31
        return h or 1
    for s in ["", "a", "device-123", "203.0.113.7", "fp:9f2c" * 9, "\x00", "ü"]:
        b = s.encode()
        assert pkg.hash_id(s) == orc.hash_id(s) == spec(b)
    assert pkg.hash_id("") == 0


def test_blacklist_check_semantics(orc, pkg):
    """CheckBlacklist (:267-293): a hit in any of the three sets, each identifier only against its own
    set, empty identifiers never checked."""
    abi = pkg.abi
    bl, keep = orc.make_blacklist(device=[11, 12], fingerprint=[21], ip=[31, 0])
    ids = np.zeros(7, dtype=abi.tx_ids_dtype)
    ids["device"] = [11, 21, 0, 0, 5, 0, 12]
    ids["fingerprint"] = [0, 0, 21, 0, 6, 0, 21]
    ids["ip"] = [0, 0, 0, 31, 7, 0, 31]
    assert orc.blacklist_check(bl, ids).tolist() == [1, 0, 1, 1, 0, 0, 1]
    del keep
