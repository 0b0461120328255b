"""Accuracy-vs-round parity (BASELINE.json: "matching accuracy-vs-round curves on synthetic data").

``profiles/curve_resnet9_sketch.json`` holds 300 rounds of ResNet-9 FetchSGD (5 x 500 000 sketch, k = 50 000) on
the shared synthetic task for three arms measured on one B200: ours at tf32 (matched precision), ours at bf16
and the UNMODIFIED reference (``bench/curves.py``).  Here the tf32 arm is re-run and must stay inside a band
around the committed reference curve; when the reference is installed (``baseline/_ref``) a short live run of
it must agree round by round."""
import importlib.util
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _curves():
    spec = importlib.util.spec_from_file_location("ceb_curves", os.path.join(ROOT, "bench", "curves.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_committed_curves_agree_across_arms():
    with open(os.path.join(ROOT, "profiles", "curve_resnet9_sketch.json")) as f:
        rec = json.load(f)
    ref = rec["reference"]
    assert len(ref["round"]) >= 300 and ref["val_acc"][-1] > 0.85            # the task is actually learned
    for arm in ("ours_tf32", "ours_bf16"):
        gap = _curves().compare(rec[arm], ref)
        assert gap["max_abs_train_loss_gap_smoothed"] < 0.02, (arm, gap)
        assert gap["max_abs_val_loss_gap"] < 0.02 and gap["max_abs_val_acc_gap"] < 0.08, (arm, gap)
        assert abs(rec[arm]["val_acc"][-1] - ref["val_acc"][-1]) < 0.03


def test_tf32_rerun_stays_in_the_band_of_the_committed_reference_curve():
    curves = _curves()
    with open(os.path.join(ROOT, "profiles", "curve_resnet9_sketch.json")) as f:
        rec = json.load(f)
    c = rec["config"]
    cfg = curves.parse(["--rounds", str(c["rounds"]), "--clients_per_round", str(c["clients_per_round"]),
                        "--per_client", str(c["per_client"]), "--noise", str(c["noise"]), "--lr_peak", str(c["lr_peak"]),
                        "--val_every", str(c["val_every"]), "--val_images", str(c["val_images"]), "--k", str(c["k"]),
                        "--num_cols", str(c["num_cols"])])
    ours = curves.run_ours(cfg, "tf32")
    gap = curves.compare(ours, rec["reference"])
    assert gap["max_abs_train_loss_gap_smoothed"] < 0.03, gap
    assert gap["max_abs_val_acc_gap"] < 0.12 and abs(ours["val_acc"][-1] - rec["reference"]["val_acc"][-1]) < 0.05, gap


def test_live_reference_short_run_matches_round_by_round(tmp_path):
    if not os.path.exists(os.path.join(ROOT, "baseline", "_ref", "CommEfficient", "fed_aggregator.py")):
        pytest.skip("reference not installed (baseline/_ref)")
    curves = _curves()
    out = str(tmp_path / "c.json")
    res = curves.main(["--rounds", "30", "--clients_per_round", "8", "--val_every", "15", "--arms", "tf32,reference",
                       "--out", out])
    if "reference" not in res:
        pytest.skip("the reference could not be run on this box")
    ours, ref = res["ours_tf32"], res["reference"]
    np.testing.assert_allclose(ours["train_loss"], ref["train_loss"], atol=5e-3)
    np.testing.assert_allclose(ours["val_loss"], ref["val_loss"], atol=5e-3)
