import torch, time
n = 979 * 1000 * 1000
a = torch.empty(n, dtype=torch.uint8).pin_memory()
b = torch.empty(n, dtype=torch.uint8, device='cuda')
for chunk in (n, 64 << 20, 16 << 20):
    for rep in range(3):
        torch.cuda.synchronize(); t = time.perf_counter()
        for o in range(0, n, chunk):
            b[o:o + chunk].copy_(a[o:o + chunk], non_blocking=True)
        torch.cuda.synchronize(); dt = time.perf_counter() - t
    print('chunk %d MiB: %.2f ms, %.1f GB/s' % (chunk >> 20, dt * 1e3, n / dt / 1e9))
