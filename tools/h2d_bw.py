"""Development aid: host->device copy rate of this box from page-locked and pageable memory (torch)."""
import torch, time
for pin in (True, False):
    h = torch.empty(256 << 20, dtype=torch.uint8, pin_memory=pin); h.fill_(1)
    d = torch.empty_like(h, device="cuda")
    for _ in range(2): d.copy_(h, non_blocking=True); torch.cuda.synchronize()
    t = time.time()
    for _ in range(5): d.copy_(h, non_blocking=True)
    torch.cuda.synchronize(); dt = (time.time() - t) / 5
    print("H2D %s: %.1f GB/s" % ("pinned" if pin else "pageable", h.numel() / dt / 1e9))
    t = time.time()
    for _ in range(5): h.copy_(d, non_blocking=True)
    torch.cuda.synchronize(); dt = (time.time() - t) / 5
    print("D2H %s: %.1f GB/s" % ("pinned" if pin else "pageable", h.numel() / dt / 1e9))
