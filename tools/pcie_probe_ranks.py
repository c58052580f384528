"""Host-memory ceiling of the box for the end-to-end path: every rank (one per GPU, torchrun) copies the c4 forecast's bytes
device -> pinned host (row-pitched per lead time, as dlwp_rollout_host does, and as one contiguous copy) and the inputs'
bytes host -> device, all ranks at the same time; rank 0 prints the aggregate GB/s."""
import ctypes
import os
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
cudart = ctypes.CDLL("libcudart.so.12")
B, T, frame = int(os.environ.get("PROBE_MEMBERS", 256)), 20, 8 * 2048 * 4
d = torch.empty(B * T * frame // 4, device="cuda")
h = torch.empty(B * T * frame // 4).pin_memory()
st = torch.cuda.current_stream().cuda_stream


def copy2d(rows_bytes, nrows, pitch, off, kind):
    src, dst = (d, h) if kind == 2 else (h, d)
    rc = cudart.cudaMemcpy2DAsync(ctypes.c_void_p(dst.data_ptr() + off), ctypes.c_size_t(pitch), ctypes.c_void_p(src.data_ptr() + off),
                                  ctypes.c_size_t(pitch), ctypes.c_size_t(rows_bytes), ctypes.c_size_t(nrows), kind, ctypes.c_void_p(st))
    assert rc == 0, rc


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    dt = torch.tensor([(time.perf_counter() - t0) / reps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    return float(dt)


total = B * T * frame
cases = {
    "D2H per lead time (rows of 64 KB, pitched)": lambda: [copy2d(frame, B, T * frame, t * frame, 2) for t in range(T)],
    "D2H contiguous": lambda: copy2d(total, 1, total, 0, 2),
    "H2D contiguous": lambda: copy2d(total, 1, total, 0, 1),
}
for name, fn in cases.items():
    dt = timed(fn)
    if rank == 0:
        print(f"{world} ranks x {B} members: {name}: {dt * 1e3:.2f} ms per rank, aggregate {world * total / dt / 1e9:.1f} GB/s "
              f"({total / dt / 1e9:.1f} per GPU)", flush=True)
if world > 1:
    dist.destroy_process_group()
