"""Probe: the c4 rollout as two half-ensembles on two streams (each plan's persistent grids capped with DLWP_SM_LIMIT) against one
256-member rollout on the whole device.  The issue-bound MLP kernels of one half can then overlap the HBM-bound layer kernels of
the other.   DLWP_SM_LIMIT=74 python tools/split_probe.py   /   python tools/split_probe.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import dlwp_benchmark_b200 as pkg  # noqa: E402


def main():
    cls, kw, H, W, T, members, _ = bench.WORKLOADS["c4"]
    ctx = kw["context_size"]
    torch.manual_seed(0)
    mods = [getattr(pkg, cls)(**kw).cuda().eval() for _ in range(2)]
    mods[1].load_state_dict(mods[0].state_dict())
    const, presc, prog = (t.cuda() if t is not None else None for t in bench.synthetic_inputs(kw, members, T + ctx, H, W, 1, False))
    reps = 6

    def run(parts, streams):
        outs = []
        for m, (lo, hi), st in zip(mods, parts, streams):
            with torch.cuda.stream(st), torch.no_grad():
                outs.append(m(constants=const[lo:hi], prescribed=presc[lo:hi], prognostic=prog[lo:hi]))
        return outs

    def timed(parts, streams):
        for _ in range(3):
            run(parts, streams)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for st in streams:
            st.wait_stream(torch.cuda.current_stream())
        for _ in range(reps):
            run(parts, streams)
        for st in streams:
            torch.cuda.current_stream().wait_stream(st)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        return ms, members * T / ms * 1e3

    s = [torch.cuda.Stream(), torch.cuda.Stream()]
    lim = os.environ.get("DLWP_SM_LIMIT", "none")
    one = timed([(0, members)], s[:1])
    two = timed([(0, members // 2), (members // 2, members)], s)
    print(f"DLWP_SM_LIMIT={lim}: one plan of {members}: {one[0]:.3f} ms ({one[1]:.0f} samples*steps/s); "
          f"two plans of {members // 2} on two streams: {two[0]:.3f} ms ({two[1]:.0f})")
    a = run([(0, members)], s[:1])[0]
    torch.cuda.synchronize()
    halves = run([(0, members // 2), (members // 2, members)], s)
    torch.cuda.synchronize()
    b = torch.cat(halves)
    print("max |one - two| =", float((a - b).abs().max()))


if __name__ == "__main__":
    main()
