import torch, time
x = torch.empty(608_000_000, dtype=torch.uint8).pin_memory()
y = torch.empty_like(x, device='cuda')
for n in (1, 2, 32):
    chunks = x.chunk(n); ychunks = y.chunk(n)
    for _ in range(3):
        torch.cuda.synchronize(); t=time.perf_counter()
        for a,b in zip(chunks, ychunks): b.copy_(a, non_blocking=True)
        torch.cuda.synchronize(); dt=time.perf_counter()-t
    print(n, "chunks: %.2f ms  %.1f GB/s" % (dt*1e3, x.numel()/dt/1e9))
