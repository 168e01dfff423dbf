import sys, time, numpy as np
sys.path.insert(0, "/root/repo")
import __graft_entry__ as ge
pkg = ge.load_package()
eng = pkg.Engine(0)
n = 2_000_000
rng = np.random.default_rng(1)
for name, amt in (("2M events in one (account, second), 1000 distinct amounts", rng.integers(1, 1000, n)),
                  ("2M events in one (account, second), all distinct", np.arange(n))):
    ev = pkg.Events(np.zeros(n, np.uint32), np.full(n, 1_700_000_000, np.uint32), amt.astype(np.int64), np.zeros(n, np.uint8))
    eng.load_events(ev, 1)
    t0 = time.perf_counter(); eng.load_events(ev, 1); dt = time.perf_counter() - t0
    q = np.zeros(1, dtype=pkg.abi.query_dtype); q["ts"] = 1_700_000_001; q["seq_limit"] = 0xFFFFFFFF
    print(name, f"{dt*1e3:.1f} ms wall, count_1m = {int(eng.get_velocity(q)['tx_count_1m'][0])}")
