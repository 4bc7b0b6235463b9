import torch, time
n = 464486400
h = torch.empty(n, dtype=torch.uint8).pin_memory(); d = torch.empty(n, dtype=torch.uint8, device='cuda')
ho = torch.empty(132710400, dtype=torch.uint8).pin_memory(); do = torch.empty(132710400, dtype=torch.uint8, device='cuda')
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for mode in ('h2d', 'both'):
    for _ in range(2):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(5):
            with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
            if mode == 'both':
                with torch.cuda.stream(s2): ho.copy_(do, non_blocking=True)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(mode, 'H2D GB/s', 5 * n / dt / 1e9, ('D2H GB/s %.1f' % (5 * 132710400 / dt / 1e9)) if mode == 'both' else '')
# chunked like the library: 4 MB pieces
chunk = 4147200
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(3):
    for o in range(0, n - chunk + 1, chunk): d[o:o + chunk].copy_(h[o:o + chunk], non_blocking=True)
torch.cuda.synchronize(); dt = time.perf_counter() - t0
print('chunked 4MB H2D GB/s', 3 * (n // chunk) * chunk / dt / 1e9)
