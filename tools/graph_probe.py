"""Is the small-batch rollout bound by host launch overhead?  Times dlwp_rollout (20 lead steps, TFNO2D 32x64) launched
directly and replayed from a CUDA graph captured around the same call, for several member counts."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import dlwp_benchmark_b200 as pkg  # noqa: E402
from bench import TFNO_C2, AFNO_C3  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    for cls, kw, lead in (("TFNO2DModule", TFNO_C2, 20), ("FourCastNet", AFNO_C3, 5)):
        torch.manual_seed(1234)
        mod = getattr(pkg, cls)(**kw).to(dev).eval()
        Cg = kw["prognostic_channels"]
        for B in (8, 16, 32, 64, 128, 256):
            T = 1 + lead
            const = torch.randn(B, 1, 4, 32, 64, device=dev)
            presc = torch.randn(B, T, 1, 32, 64, device=dev)
            prog = torch.randn(B, T, Cg, 32, 64, device=dev)
            out = torch.empty(B, lead, Cg, 32, 64, device=dev)
            with torch.no_grad():
                ref = mod(constants=const, prescribed=presc, prognostic=prog)
            plan = next(iter(mod._plans.values())) if len(mod._plans) == 1 else mod._plan_for(dev, B, 32, 64)
            plan = mod._plan_for(dev, B, 32, 64)

            def direct():
                plan.rollout(const, presc, prog, out)

            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                for _ in range(3):
                    direct()
                s.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=s):
                    direct()
            torch.cuda.synchronize()

            def timeit(fn, n=20):
                fn()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0 = time.perf_counter()
                e0.record()
                for _ in range(n):
                    fn()
                e1.record()
                t_host = time.perf_counter() - t0
                torch.cuda.synchronize()
                return e0.elapsed_time(e1) / n, t_host / n * 1e3

            d_ms, d_host = timeit(direct)
            out.zero_()
            g_ms, g_host = timeit(g.replay)
            ok = torch.equal(out, ref)
            print(f"{cls} B={B:4d}: direct {d_ms * 1e3 / lead:7.1f} us/step (host enqueue {d_host * 1e3 / lead:6.1f} us/step), "
                  f"graph {g_ms * 1e3 / lead:7.1f} us/step (host {g_host * 1e3 / lead:5.1f}); graph output == direct: {ok}; "
                  f"{B * lead / d_ms / 1e3:8.1f} k -> {B * lead / g_ms / 1e3:8.1f} k samples*steps/s", flush=True)


if __name__ == "__main__":
    main()
