"""Raw PCIe rates of the box with pinned memory: one direction alone, both directions at once, and the effect of
splitting the device->host side into several copies (what StepPipeline moves per KITTI step)."""
import torch
dev = torch.device("cuda:0")
MB = 1 << 20
def run(h2d_sizes, d2h_sizes, reps=50):
    hs = [torch.empty(s, dtype=torch.uint8).pin_memory() for s in h2d_sizes]
    ds = [torch.empty(s, dtype=torch.uint8, device=dev) for s in h2d_sizes]
    do = [torch.empty(s, dtype=torch.uint8, device=dev) for s in d2h_sizes]
    ho = [torch.empty(s, dtype=torch.uint8).pin_memory() for s in d2h_sizes]
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    def once():
        with torch.cuda.stream(s1):
            for h, d in zip(hs, ds): d.copy_(h, non_blocking=True)
        with torch.cuda.stream(s2):
            for d, h in zip(do, ho): h.copy_(d, non_blocking=True)
    for _ in range(5): once()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    s1.wait_event(a); s2.wait_event(a)
    for _ in range(reps): once()
    e1, e2 = torch.cuda.Event(), torch.cuda.Event()
    e1.record(s1); e2.record(s2)
    torch.cuda.current_stream().wait_event(e1); torch.cuda.current_stream().wait_event(e2)
    b.record(); b.synchronize()
    ms = a.elapsed_time(b) / reps
    return ms, sum(h2d_sizes) / ms / 1e6, sum(d2h_sizes) / ms / 1e6
k = 23040000
print("H2D alone 23 MB          : %.3f ms  %.1f GB/s" % run([k], [])[:2])
r = run([], [k]); print("D2H alone 23 MB          : %.3f ms  %.1f GB/s" % (r[0], r[2]))
print("both, one copy each      : %.3f ms  H2D %.1f  D2H %.1f GB/s" % run([k], [k]))
print("both, as StepPipeline    : %.3f ms  H2D %.1f  D2H %.1f GB/s" % run([k // 2, k // 2], [7680000, 1536000, 384000, 28, 1536000, 11520000]))
print("both, 4 x 5.76 MB each   : %.3f ms  H2D %.1f  D2H %.1f GB/s" % run([k // 4] * 4, [k // 4] * 4))
