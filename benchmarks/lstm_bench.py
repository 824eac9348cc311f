"""PTB-shaped LSTM language-model layers (S/example/languagemodel/PTBModel.scala: 2 layers, hidden 650 / 1500, 35 steps,
batch 20): forward + backward of the recurrent stack, ours (one hoisted tcgen05 GEMM + one persistent kernel per layer
and direction, csrc/lstm_seq.cu) vs torch.nn.LSTM (cuDNN), words/s from CUDA events.
    python benchmarks/lstm_bench.py [--batch 20] [--steps 35]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch


def timeit(fn, iters=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=20)
    ap.add_argument("--steps", type=int, default=35)
    ap.add_argument("--layers", type=int, default=2)
    args = ap.parse_args()
    from bigdl_b200.nn.layer import LSTM, Recurrent, Sequential
    from bigdl_b200.utils import config

    B, T, L = args.batch, args.steps, args.layers
    for H in (650, 1500):
        x = torch.randn(B, T, H, device="cuda").bfloat16()
        g = (torch.randn(B, T, H, device="cuda") * 0.01).bfloat16()

        def build():
            m = Sequential()
            for _ in range(L):
                m.add(Recurrent().add(LSTM(H, H)))
            m.cuda().training()
            m.getParameters()
            return m

        res = {"bench": "ptb_lstm", "hidden": H, "batch": B, "steps": T, "layers": L}
        for name, persistent in (("ours_persistent", True), ("ours_step_loop", False)):
            config.set("bigdl.rnn.persistent", persistent)
            m = build()

            def step():
                xin = x.clone().requires_grad_(True)
                with torch.enable_grad():
                    y = m._call(xin)
                torch.autograd.backward(y, g)

            ms = timeit(step, iters=20 if persistent else 5, warm=3)
            res[name + "_ms"] = round(ms, 3)
            res[name + "_words_per_s"] = round(B * T / ms * 1e3)
            config.unset("bigdl.rnn.persistent")
        ref = torch.nn.LSTM(H, H, L, batch_first=True).cuda().bfloat16()

        def ref_step():
            xin = x.clone().requires_grad_(True)
            y, _ = ref(xin)
            y.backward(g)

        ms = timeit(ref_step)
        res["cudnn_ms"] = round(ms, 3)
        res["cudnn_words_per_s"] = round(B * T / ms * 1e3)
        res["persistent_vs_cudnn"] = round(res["cudnn_ms"] / res["ours_persistent_ms"], 3)
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
