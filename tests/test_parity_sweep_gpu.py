"""A bounded slice of tools/parity_sweep.py (the config-size parity sweep against oracle/_ref) as a GPU test."""
import os, sys
import pytest
import refutil as R

pytestmark = pytest.mark.gpu
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))


@pytest.mark.skipif(not R.have_ref(), reason="oracle/_ref not shipped")
def test_quick_sweep_is_bit_exact():
    import parity_sweep as P
    res = P.Sweep(quick=True).run(["c1", "c2", "c3", "c5", "levels"])
    assert res["bit_exact_tier_mismatches"] == 0, [r for r in res["rows"] if r["mismatches"]]
    rdo = [r for r in res["rows"] if r["what"] == "encode+rdo"][0]
    assert rdo["psnr_within_0.01_db"] and rdo["mode_agreement"] >= 0.99, rdo
    c5 = [r for r in res["rows"] if r["config"].startswith("C5 base")][0]
    assert res["encode_blocks_checked"] > 100000 and c5["blocks"] > 0
