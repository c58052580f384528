"""PCIe probe on the GPU box: contiguous and row-pitched device -> pinned-host copies at the c4 forecast's sizes."""
import ctypes, time, torch
cudart = ctypes.CDLL("libcudart.so.12")
B, T, frame = 256, 20, 8 * 2048 * 4
d = torch.empty(B * T * frame // 4, device="cuda")
h = torch.empty(B * T * frame // 4).pin_memory()
torch.cuda.synchronize()
st = torch.cuda.current_stream().cuda_stream


def copy2d(rows_bytes, nrows, pitch, off):
    rc = cudart.cudaMemcpy2DAsync(ctypes.c_void_p(h.data_ptr() + off), ctypes.c_size_t(pitch), ctypes.c_void_p(d.data_ptr() + off),
                                  ctypes.c_size_t(pitch), ctypes.c_size_t(rows_bytes), ctypes.c_size_t(nrows), 2, ctypes.c_void_p(st))
    assert rc == 0, rc


for g in (1, 2, 4, 5, 10, 20):      # lead times per copy: rows of g frames, pitch T frames
    for _ in range(2):
        for t in range(0, T, g):
            copy2d(g * frame, B, T * frame, t * frame)
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    for t in range(0, T, g):
        copy2d(g * frame, B, T * frame, t * frame)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"rows of {g * frame // 1024} KB x {B}, {T // g} copies: {dt * 1e3:.2f} ms = {B * T * frame / dt / 1e9:.1f} GB/s")
