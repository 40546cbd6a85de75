"""Prints the FP8 loss trajectories of tests/test_golden.py::test_loss_tracks_hf_bf16_run next to the HF bf16 fixture (GPU box)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _swap import build_swapped_causal_lm  # noqa: E402

import llm_fp8_b200  # noqa: E402

fx = json.load(open(os.path.join(ROOT, "tests", "golden", "hf_bf16_loss.json")))
tiny, run = fx["tiny_config"], fx["run"]
print("bf16   ", " ".join(f"{v:.3f}" for v in fx["losses"]))
for scenario in ("default", "hybrid", "mxfp8"):
    for calib in (0, 2):
        llm_fp8_b200.reset_fp8_state()
        model = build_swapped_causal_lm(tiny, run["seed"], scenario).cuda()
        g = torch.Generator().manual_seed(run["seed"] + 1)
        tokens = torch.randint(0, tiny["vocab_size"], (run["batch"], run["seq"]), generator=g).cuda()
        opt = torch.optim.AdamW(model.parameters(), lr=run["lr"])
        for _ in range(calib):
            model(input_ids=tokens, labels=tokens).loss.backward()
            opt.zero_grad(set_to_none=True)
        losses = []
        for _ in range(run["steps"]):
            out = model(input_ids=tokens, labels=tokens)
            out.loss.backward()
            opt.step()
            opt.zero_grad(set_to_none=True)
            losses.append(float(out.loss.detach()))
        rel = max(abs(a - b) / b for a, b in zip(losses, fx["losses"]))
        print(f"{scenario:8s} calib={calib}", " ".join(f"{v:.3f}" for v in losses), f"max rel dev {rel:.3f}")
