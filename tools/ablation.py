"""What each fusion / kernel buys in the benchmark step, measured inside one process on one GPU: the step is timed
with everything on, then with one feature at a time sent back to the library call it replaced (A/B/A/B interleaved).
usage: python tools/ablation.py"""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools import step_variants as sv
from comfy_gen_server_b200 import _native
from comfy_gen_server_b200.model_base import build_model
from comfy_gen_server_b200.model_management import ModelPatcher


def main():
    dev = torch.device("cuda")
    model = build_model("sdxl", dtype=torch.bfloat16, device=dev, seed=1234)
    patcher = ModelPatcher(model, dev)
    patcher.model_options["cuda_graph"] = False
    step = sv.make_stepper(model, patcher, dev)
    sv.timeit(step, warm=3, n=3)
    rows = {}
    for feat in list(_native.FEATURES):
        on, off = [], []
        for rep in range(3):
            _native.FEATURES[feat] = True
            on.append(sv.timeit(step, warm=1, n=4))
            _native.FEATURES[feat] = False
            off.append(sv.timeit(step, warm=1, n=4))
        _native.FEATURES[feat] = True
        rows[feat] = {"on_ms": round(min(on), 2), "off_ms": round(min(off), 2), "gain_ms": round(min(off) - min(on), 2)}
        print(feat, json.dumps(rows[feat]), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "ablation.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
