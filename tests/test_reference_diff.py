"""Function-by-function differential test against the UNMODIFIED reference (``baseline/_ref``) on the CPU:
the server rules of all five modes, the client step (momentum, error feedback, local top-k, clipping,
sketching, microbatches), every CV model class (same weights and input -> same output and gradients), the
round sampler (same numpy seed -> same rounds) and the federated datasets (same client files -> same items).  The comparison runs in a subprocess (``baseline/diff_rules.py``) because the
reference's modules are top-level names."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_rules_match_the_reference_function_by_function():
    if not os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "CommEfficient")):
        pytest.skip("baseline/_ref is not installed (python baseline/install_ref.py)")
    proc = subprocess.run([sys.executable, os.path.join(ROOT, "baseline", "diff_rules.py")], stdout=subprocess.PIPE,
                          stderr=subprocess.PIPE, text=True, timeout=600)
    assert proc.returncode == 0, proc.stderr[-2000:]
    res = json.loads(proc.stdout.strip().splitlines()[-1])
    assert "unavailable" not in res, res
    assert len(res) >= 46 and "loop/reference_gpt2_loop_drives_our_classes" in res and "loop/reference_train_loop_drives_our_classes" in res and "data/transforms" in res and "data/fed_emnist_prepared_from_leaf_json" in res and "rounds/fedavg_local_sgd" in res and "data/fed_imagenet_items" in res and "client/process_batch_topk_down" in res and "cli/flags_dests_defaults" in res and "data/fed_persona_items_and_collate" in res and any(c.startswith("models/") for c in res) and "data/fed_sampler_rounds" in res
    for case, err in res.items():
        if isinstance(err, str) and err.startswith("skipped"):
            continue
        assert isinstance(err, float) and err < 1e-4, (case, err)
