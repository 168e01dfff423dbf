"""CPU: the C-ABI library builds, loads, exports every symbol include/igs.h declares, its pure-host
helpers agree with the oracle, and the product fails loudly (no fallback) without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_header_symbols_match_export_list(pkg):
    hdr = open(os.path.join(ROOT, "include", "igs.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(igs_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(pkg.abi.EXPORTS)


def test_library_exports_every_declared_symbol(pkg, built_lib):
    lib = C.CDLL(built_lib)
    for name in pkg.abi.EXPORTS:
        assert hasattr(lib, name), f"{name} is declared in include/igs.h but not exported"
    assert b"sm_100a" in pkg.abi.load_library().igs_version()


def test_default_config_matches_oracle(pkg, orc, built_lib):
    a = pkg.default_config()
    b = orc.default_config()
    assert bytes(a) == bytes(b)
    assert (a.block_threshold, a.review_threshold, a.max_tx_per_minute, a.new_account_days) == (80, 50, 10, 7)
    assert (a.ml_weight, a.rule_weight, a.large_deposit_amount) == (0.6, 0.4, 100000)
    assert list(a.reason_weight) == [20, 30, 25, 15, 20, 15, 50, 25, 20, 30]


def test_struct_sizes_match_header(pkg):
    """numpy dtypes / ctypes structs == the C compiler's view of include/igs.h."""
    import subprocess, tempfile
    src = r'''
#include <stdio.h>
#include "igs.h"
int main(void){printf("%zu %zu %zu %zu %zu %zu %zu\n", sizeof(igs_event_ids), sizeof(igs_tx_ids), sizeof(igs_uniques),
 sizeof(igs_link_pair), sizeof(igs_link), sizeof(igs_links_row), sizeof(igs_abuse));
 printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\n", sizeof(igs_config),
 sizeof(igs_rule_row), sizeof(igs_result), sizeof(igs_tx), sizeof(igs_player_row), sizeof(igs_query),
 sizeof(igs_velocity), sizeof(igs_ltv_row), sizeof(igs_ltv_result), sizeof(igs_gbdt_desc),
 sizeof(igs_mlp_desc), sizeof(igs_event_cols));return 0;}'''
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "s.c"), "w").write(src)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(d, "s"),
                               os.path.join(d, "s.c")])
        got = [int(x) for x in subprocess.check_output([os.path.join(d, "s")]).split()]
    abi = pkg.abi
    want = [C.sizeof(abi.EventIds), abi.tx_ids_dtype.itemsize, abi.uniques_dtype.itemsize, abi.link_pair_dtype.itemsize,
            abi.link_dtype.itemsize, abi.links_row_dtype.itemsize, abi.abuse_dtype.itemsize,
            C.sizeof(abi.Config), abi.rule_row_dtype.itemsize, abi.result_dtype.itemsize,
            abi.tx_dtype.itemsize, abi.player_row_dtype.itemsize, abi.query_dtype.itemsize,
            abi.velocity_dtype.itemsize, abi.ltv_row_dtype.itemsize, abi.ltv_result_dtype.itemsize,
            C.sizeof(abi.GbdtDesc), C.sizeof(abi.MlpDesc), C.sizeof(abi.EventCols)]
    assert got == want


def test_host_helpers(pkg, built_lib):
    lib = pkg.abi.load_library()
    cfg = pkg.default_config()
    abi = pkg.abi
    assert lib.igs_rule_score_from_flags(C.byref(cfg), abi.R_HIGH_VELOCITY | abi.R_VPN_DETECTED) == 35
    assert lib.igs_rule_score_from_flags(C.byref(cfg), 0x3FF) == 100          # capped, ML excluded
    assert lib.igs_rule_score_from_flags(C.byref(cfg), abi.R_ML_HIGH_RISK) == 0
    arr = (C.c_char_p * abi.NUM_REASONS)()
    n = lib.igs_reason_codes(abi.R_ML_HIGH_RISK | abi.R_KNOWN_FRAUDSTER | abi.R_MULTIPLE_DEVICES
                             | abi.R_IP_COUNTRY_MISMATCH, C.byref(arr))
    # emission order of applyRules: devices (R3) before IPs (R4), blacklist (R8), then ML
    assert [arr[i].decode() for i in range(n)] == ["MULTIPLE_DEVICES", "IP_COUNTRY_MISMATCH",
                                                   "KNOWN_FRAUDSTER", "ML_HIGH_RISK"]
    assert [lib.igs_action_name(i).decode() for i in (1, 2, 3)] == ["approve", "review", "block"]
    assert lib.igs_segment_name(5) == b"churning"
    assert lib.igs_ltv_action_name(5) == b"ASSIGN_VIP_MANAGER"
    shards = [lib.igs_shard_of(i, i * 7919, 8) for i in range(4000)]
    assert set(shards) == set(range(8))
    counts = np.bincount(shards, minlength=8)
    assert counts.min() > 400 and counts.max() < 600


@pytest.mark.skipif(_has_gpu(), reason="only meaningful on a box without a GPU")
def test_engine_fails_loudly_without_gpu(pkg, built_lib):
    with pytest.raises(pkg.IgsError) as ei:
        pkg.Engine(device=0)
    assert "no CPU fallback" in str(ei.value) or "CUDA" in str(ei.value)


def test_product_never_touches_the_oracle():
    """The package must not import, link or call anything under oracle/."""
    pkg_dir = os.path.join(ROOT, "igaming-platform_b200")
    for dp, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "oracle" not in txt.replace("the oracle", "").replace("oracle/igs_oracle.c", "") \
                    .replace("with the oracle", "").replace("oracle's", "").replace("CPU oracle", ""), \
                    f"{f} mentions the oracle in code"
                assert "liboracle" not in txt and "orc_" not in txt.replace("orc_sigmoid", "")
