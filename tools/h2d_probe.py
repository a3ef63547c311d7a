#!/usr/bin/env python
"""H2D copy-rate probe (pinned host memory -> device) at the copy sizes the front end uses:
one stereo image (466 616 B), one stereo pair, and one whole 64-sequence step in a single copy."""
import torch

W, H, S = 1241, 376, 64
img = W * H
host = torch.empty(S * 2 * img, dtype=torch.uint8).pin_memory()
dev = torch.empty_like(host, device="cuda")
st = torch.cuda.Stream()


def run(chunk):
    n = host.numel() // chunk
    with torch.cuda.stream(st):
        for _ in range(3):
            for i in range(n):
                dev[i * chunk:(i + 1) * chunk].copy_(host[i * chunk:(i + 1) * chunk], non_blocking=True)
        st.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        reps = 20
        for _ in range(reps):
            for i in range(n):
                dev[i * chunk:(i + 1) * chunk].copy_(host[i * chunk:(i + 1) * chunk], non_blocking=True)
        b.record(st)
        st.synchronize()
    ms = a.elapsed_time(b) / reps
    print(f"chunk {chunk:>9d} B x {n:4d}: {ms:7.3f} ms per 64-pair step, {host.numel() / ms / 1e6:6.1f} GB/s")


for c in (img, 2 * img, 8 * img, host.numel()):
    run(c)
