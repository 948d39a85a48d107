"""Raw PCIe bandwidth of this box (pinned host memory, CUDA events): one big copy per direction, both directions at once,
and the D2H direction cut into the copy sizes acb200_env_step_host issues.  Prints one JSON line."""
import json
import torch

dev = torch.device("cuda", 0)
n = 1 << 20
out = {}


def timed(fn, reps=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e-3


big_d = torch.empty(80 << 20, dtype=torch.uint8, device=dev); big_h = torch.empty(80 << 20, dtype=torch.uint8, pin_memory=True)
out["d2h_80MB_GBs"] = (80 << 20) / timed(lambda: big_h.copy_(big_d, non_blocking=True)) / 1e9
out["h2d_80MB_GBs"] = (80 << 20) / timed(lambda: big_d.copy_(big_h, non_blocking=True)) / 1e9
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
d2 = torch.empty(16 << 20, dtype=torch.uint8, device=dev); h2 = torch.empty(16 << 20, dtype=torch.uint8, pin_memory=True)


def both():
    s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s1):
        big_h.copy_(big_d, non_blocking=True)
    with torch.cuda.stream(s2):
        d2.copy_(h2, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)


out["d2h_80MB_with_h2d_16MB_GBs"] = (80 << 20) / timed(both) / 1e9
# the env's per-chunk copies: 16 chunks x (obs 4.7 MB, reward 256 KB, two flag arrays 64 KB)
per = n // 16
obs_d = torch.empty((n, 18), dtype=torch.float32, device=dev); obs_h = torch.empty((n, 18), dtype=torch.float32, pin_memory=True)
rew_d = torch.empty(n, dtype=torch.float32, device=dev); rew_h = torch.empty(n, dtype=torch.float32, pin_memory=True)
f1_d = torch.empty(n, dtype=torch.uint8, device=dev); f1_h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
f2_d = torch.empty(n, dtype=torch.uint8, device=dev); f2_h = torch.empty(n, dtype=torch.uint8, pin_memory=True)


def chunked():
    for c in range(16):
        sl = slice(c * per, (c + 1) * per)
        obs_h[sl].copy_(obs_d[sl], non_blocking=True); rew_h[sl].copy_(rew_d[sl], non_blocking=True)
        f1_h[sl].copy_(f1_d[sl], non_blocking=True); f2_h[sl].copy_(f2_d[sl], non_blocking=True)


out["d2h_env_chunks_GBs"] = n * 78 / timed(chunked) / 1e9


def obs_only():
    for c in range(16):
        sl = slice(c * per, (c + 1) * per)
        obs_h[sl].copy_(obs_d[sl], non_blocking=True)


out["d2h_obs_chunks_only_GBs"] = n * 72 / timed(obs_only) / 1e9
print(json.dumps(out))
