"""host <-> device copy rates from pinned memory on this box: each direction alone, both at once
(two streams), in the chunk sizes the sliced batch path uses"""
import torch, time
dev = torch.device("cuda:0")
for mb in (4, 17, 34, 134):
    n = mb * (1 << 20)
    h_in, h_out = torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.uint8).pin_memory()
    d_in, d_out = torch.empty(n, dtype=torch.uint8, device=dev), torch.empty(n, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    def run(up, down, reps=10):
        torch.cuda.synchronize(); t = time.perf_counter()
        for _ in range(reps):
            if up:
                with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
            if down:
                with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize(); return (time.perf_counter() - t) / reps
    run(True, True, 2)
    a, b, c = run(True, False), run(False, True), run(True, True)
    print(f"{mb:4d} MB  H2D {n / a / 1e9:5.1f} GB/s  D2H {n / b / 1e9:5.1f} GB/s  both at once {c * 1e3:6.2f} ms ({n / c / 1e9:5.1f} GB/s each way)")
