import sys, json
sys.path.insert(0, '.')
import importlib.util
spec = importlib.util.spec_from_file_location("curves", "bench/curves.py"); curves = importlib.util.module_from_spec(spec); spec.loader.exec_module(curves)
for noise, lr, rounds, W in ((0.5, 0.15, 300, 16), (0.5, 0.2, 400, 16), (1.0, 0.2, 400, 16), (0.5, 0.2, 300, 32), (1.0, 0.3, 300, 32)):
    r = curves.main(["--rounds", str(rounds), "--clients_per_round", str(W), "--noise", str(noise), "--lr_peak", str(lr),
                     "--arms", "tf32,bf16", "--val_every", "25", "--out", "/tmp/c.json"])
    for arm in ("ours_tf32", "ours_bf16"):
        o = r[arm]
        print("noise", noise, "lr", lr, "R", rounds, "W", W, arm, "train", [round(x, 2) for x in o["train_loss"][24::25]], "val_acc", [round(x,2) for x in o["val_acc"]], flush=True)
