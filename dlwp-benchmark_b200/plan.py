"""Host-side owner of a native ``dlwp_plan`` plus the bin bookkeeping both model families need."""
from __future__ import annotations

import ctypes as C
from typing import Dict, Iterable, List, Sequence, Tuple

import torch

from . import native


def neuralop_row_bins(H: int, m1: int) -> Tuple[List[int], List[int]]:
    """FFT row bins read / written by weight row ``j`` of a neuralop 0.3.x ``SpectralConv``.

    neuralop shifts the spectrum (``fftshift`` over rows), keeps the centred ``m1`` rows, and
    applies ``fftshift`` (not ``ifftshift``) again before the inverse transform.  For odd ``H``
    the second shift displaces every bin by one; that behaviour is kept, not fixed.
    """
    m = min(H, int(m1))
    st = H - m
    lo, hi = st // 2, H - (st - st // 2)
    pos = list(range(lo, hi))
    return [(i - H // 2) % H for i in pos], [(i + H // 2) % H for i in pos]


def centre_crop(max_m: int, m: int) -> slice:
    """Rows neuralop keeps of a weight stored with ``max_n_modes`` rows when ``n_modes`` is smaller."""
    st = max_m - min(max_m, m)
    return slice(st // 2, max_m - (st - st // 2))


def afno_kept_region(h: int, w_half: int, frac: float) -> Tuple[List[int], int]:
    """Rows and number of leading columns AFNO2D mixes (fourcastnet.py:93-96): both are
    derived from the ROW count ``h``; Python slicing clamps them to the spectrum."""
    total = h // 2 + 1
    kept = int(total * frac)
    rows = list(range(h))[max(total - kept, 0): total + kept]
    return rows, min(kept, w_half)


class NativePlan:
    """RAII wrapper: creates the plan, uploads weights, launches rollouts on torch's stream."""

    def __init__(self, desc_fields: Dict, rows_in: Sequence[int], rows_out: Sequence[int], device: torch.device):
        self._lib = native.lib()
        self._rows_in = native.int32_array(list(rows_in))
        self._rows_out = native.int32_array(list(rows_out))
        d = native.ModelDesc()
        for k, v in desc_fields.items():
            setattr(d, k, v)
        d.n_rows = len(rows_in)
        d.rows_in = C.cast(self._rows_in, C.POINTER(C.c_int32))
        d.rows_out = C.cast(self._rows_out, C.POINTER(C.c_int32))
        self.desc = d
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise native.NativeError("dlwp_benchmark_b200 runs on CUDA devices only (no CPU fallback)")
        self.max_batch = int(d.max_batch)
        handle = C.c_void_p()
        index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        native.check(self._lib.dlwp_plan_create(C.byref(d), index, C.byref(handle)), "dlwp_plan_create")
        self._h = handle
        self.weights_key = None

    def close(self):
        if getattr(self, "_h", None):
            self._lib.dlwp_plan_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def workspace_bytes(self) -> int:
        n = C.c_size_t()
        native.check(self._lib.dlwp_plan_workspace_bytes(self._h, C.byref(n)))
        return int(n.value)

    def uses_tensor_cores(self) -> bool:
        n = C.c_int()
        native.check(self._lib.dlwp_plan_uses_tensor_cores(self._h, C.byref(n)))
        return bool(n.value)

    def check_guards(self):
        """Raise when a kernel wrote outside a plan-owned buffer (guard zones around every allocation)."""
        native.check(self._lib.dlwp_plan_check_guards(self._h, self._stream()), "dlwp_plan_check_guards")

    def debug_read(self, name: str, nbytes: int) -> torch.Tensor:
        """First `nbytes` bytes of a named internal workspace buffer (diagnostics; see dlwp_plan_debug_read)."""
        out = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        native.check(self._lib.dlwp_plan_debug_read(self._h, name.encode(), C.c_void_p(out.data_ptr()), nbytes,
                                                    self._stream()), "debug_read")
        torch.cuda.current_stream(self.device).synchronize()
        return out

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def load(self, named: Iterable[Tuple[str, torch.Tensor]]):
        keep = []
        for name, t in named:
            t = t.detach().to(device=self.device, dtype=torch.float32).contiguous()
            keep.append(t)
            shape = (C.c_int64 * max(t.dim(), 1))(*t.shape)
            native.check(self._lib.dlwp_plan_set_tensor(self._h, name.encode(), C.c_void_p(t.data_ptr()), t.dim(),
                                                        shape, self._stream()), f"set_tensor({name})")
        native.check(self._lib.dlwp_plan_commit_weights(self._h, self._stream()), "commit_weights")
        torch.cuda.current_stream(self.device).synchronize()   # `keep` may be freed after this
        del keep

    @staticmethod
    def _ptr(t):
        return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p()

    def step(self, x: torch.Tensor, y: torch.Tensor):
        native.check(self._lib.dlwp_step(self._h, self._ptr(x), self._ptr(y), x.shape[0], self._stream()), "dlwp_step")

    def rollout(self, constants, prescribed, prognostic, outs):
        B, T = prognostic.shape[0], prognostic.shape[1]
        native.check(self._lib.dlwp_rollout(self._h, self._ptr(constants), self._ptr(prescribed),
                                            self._ptr(prognostic), B, T, self._ptr(outs), self._stream()),
                     "dlwp_rollout")

    def rollout_host(self, constants, prescribed, prognostic, outs):
        B, T = prognostic.shape[0], prognostic.shape[1]
        native.check(self._lib.dlwp_rollout_host(self._h, self._ptr(constants), self._ptr(prescribed),
                                                 self._ptr(prognostic), B, T, self._ptr(outs), self._stream()),
                     "dlwp_rollout_host")
