"""Repeat-run bit equality of a full 20-step rollout under three launch regimes: one rollout then a synchronisation,
rollouts enqueued back to back (successive rollouts overlap through programmatic dependent launch), and replays of
a CUDA graph captured around the call.  Every regime must give the same bits."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import dlwp_benchmark_b200 as pkg  # noqa: E402
from bench import TFNO_C2, AFNO_C3  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    bad = 0
    for cls, kw, lead in (("TFNO2DModule", TFNO_C2, 20), ("FourCastNet", AFNO_C3, 5)):
        torch.manual_seed(1234)
        mod = getattr(pkg, cls)(**kw).to(dev).eval()
        Cg = kw["prognostic_channels"]
        for B in (8, 16, 32, 64, 256):
            T = 1 + lead
            const = torch.randn(B, 1, 4, 32, 64, device=dev)
            presc = torch.randn(B, T, 1, 32, 64, device=dev)
            prog = torch.randn(B, T, Cg, 32, 64, device=dev)
            plan = mod._plan_for(dev, B, 32, 64)
            outs = [torch.empty(B, lead, Cg, 32, 64, device=dev) for _ in range(6)]
            torch.cuda.synchronize()
            plan.rollout(const, presc, prog, outs[0])
            torch.cuda.synchronize()
            for o in outs[1:5]:      # back to back
                plan.rollout(const, presc, prog, o)
            torch.cuda.synchronize()
            s = torch.cuda.Stream()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.stream(s):
                with torch.cuda.graph(g, stream=s):
                    plan.rollout(const, presc, prog, outs[5])
            torch.cuda.synchronize()
            outs[5].zero_()
            torch.cuda.synchronize()
            for _ in range(4):
                g.replay()
            torch.cuda.synchronize()
            res = [bool(torch.equal(o, outs[0])) for o in outs[1:]]
            d = max(float((o - outs[0]).abs().max()) for o in outs[1:])
            bad += 0 if all(res) else 1
            print(f"{cls} B={B:3d}: back-to-back == synced: {res[:4]}, graph replay == synced: {res[4]}, max |diff| {d:.3e}", flush=True)
    print("RACE PROBE:", "all regimes bit-identical" if bad == 0 else f"{bad} configurations differ")


if __name__ == "__main__":
    main()
