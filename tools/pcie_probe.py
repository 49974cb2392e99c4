import torch, time
n = 256 << 20
h = torch.empty(n, dtype=torch.int8).pin_memory()
d = torch.empty(n, dtype=torch.int8, device="cuda")
h2 = torch.empty(32 << 20, dtype=torch.int8).pin_memory()
d2 = torch.empty(32 << 20, dtype=torch.int8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for sz in (1 << 20, 8 << 20, 11 << 20, 64 << 20, 256 << 20):
    for both in (False, True):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        reps = max(1, (1 << 30) // sz)
        for _ in range(reps):
            with torch.cuda.stream(s1):
                d[:sz].copy_(h[:sz], non_blocking=True)
            if both:
                with torch.cuda.stream(s2):
                    h2[:sz // 8].copy_(d2[:sz // 8], non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print(f"chunk {sz >> 20:4d} MiB  d2h_concurrent={both}  H2D {reps * sz / dt / 1e9:6.1f} GB/s")
