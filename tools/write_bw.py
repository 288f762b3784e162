"""What a pure WRITE stream reaches on this GPU (the 4-state traversal is one: 25.3 GB written, ~0.6 GB read per cfg2
evaluation), next to the copy figure MEASURED_PEAKS.json quotes.  CUDA events, 3 warm-ups, buffers far larger than L2."""
import json
import torch

GB = 1 << 30
n = 8 * GB
a = torch.empty(n, dtype=torch.uint8, device="cuda")
b = torch.empty(n, dtype=torch.uint8, device="cuda")


def timed(fn, reps=5):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e-3


af = a.view(torch.float64)
out = {}
t = timed(lambda: a.zero_())
out["memset_u8_GBps"] = n / t / 1e9
t = timed(lambda: af.fill_(1.5))
out["fill_f64_GBps"] = n / t / 1e9
t = timed(lambda: b.copy_(a))
out["copy_GBps_read_plus_write"] = 2 * n / t / 1e9
t = timed(lambda: torch.sum(af))
out["read_sum_f64_GBps"] = n / t / 1e9
print(json.dumps(out))
