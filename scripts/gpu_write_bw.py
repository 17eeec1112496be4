"""Development aid: pure-write and pure-read HBM bandwidth (cudaMemset of 16 GB; a sum over 16 GB) next to the copy figure."""
import torch
x = torch.empty(2 << 30, dtype=torch.float64, device="cuda")
for name, fn, bytes_ in (("memset (write only)", lambda: x.zero_(), x.numel() * 8), ("sum (read only)", lambda: x.sum(), x.numel() * 8)):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        fn()
    e1.record(); torch.cuda.synchronize()
    print(f"{name}: {3 * bytes_ / e0.elapsed_time(e1) / 1e6:.0f} GB/s")
