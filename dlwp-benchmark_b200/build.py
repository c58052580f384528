"""Build ``libdlwp_b200.so`` in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libdlwp_b200.so")
SOURCES = ["kernels_fp32.cu", "plan.cu", "api.cu", "profile.cu", "tc_selftest.cu", "tc_fno.cu", "tc_mid.cu", "tc_mid2.cu", "tc_fno_lp.cu", "tc_afno_mid.cu", "tc_plan.cu", "tc_wide.cu", "tc_wide_mix.cu", "tc_wide_plan.cu", "tc_afno.cu", "tc_afno_plan.cu", "afno_fused.cu", "static_scratch.cu", "spectral_ops.cu", "eval_ops.cu", "train_ops.cu", "train_afno.cu"]
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC",
         "-diag-suppress", "550", "-shared"]


def _stale() -> bool:
    if not os.path.isfile(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "dlwp_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """One object per translation unit (compiled in parallel, only the stale ones), then one link."""
    if not force and not _stale():
        return OUT
    from concurrent.futures import ThreadPoolExecutor
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    # DLWP_TRACE=1: clock64 timestamps of the layer kernel's pipeline events (tools/trace_layer.py)
    trace = bool(os.environ.get("DLWP_TRACE"))
    extra = ["-DDLWP_TRACE"] if trace else []
    more = os.environ.get("DLWP_NVCC_EXTRA", "").split()      # experiment builds (-DNAME=value), in their own object directory
    extra += more
    objdir = os.path.join(HERE, "_build_x" if more else "_build_trace" if trace else "_build")
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if not f.endswith(".cu")]
    headers.append(os.path.join(HERE, "..", "include", "dlwp_b200.h"))
    hdr_t = max(os.path.getmtime(h) for h in headers)
    flags = [f for f in FLAGS if f != "-shared"]

    def compile_one(src: str) -> str:
        cu, obj = os.path.join(CSRC, src), os.path.join(objdir, src[:-3] + ".o")
        if not force and os.path.isfile(obj) and os.path.getmtime(obj) > max(hdr_t, os.path.getmtime(cu)):
            return obj
        cmd = [nvcc] + flags + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", cu, "-o", obj]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("nvcc failed on " + src + ":\n" + res.stdout + res.stderr)
        if verbose:
            print(res.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    res = subprocess.run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a"] + objs + ["-o", OUT],
                         capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("link failed:\n" + res.stdout + res.stderr)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
