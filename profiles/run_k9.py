"""Minimal driver for profiling k9 under ncu: a few launches over device-resident
config-2 strings, nothing else (no CPU baseline, no multiprocessing)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import kubectl_b200  # noqa: E402,F401
from kubectl_b200 import strings, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
data, off = synth.command_csr(n, seed=2)
d_data, d_off = torch.from_numpy(data).cuda(), torch.from_numpy(off).cuda()
v = torch.empty(n, dtype=torch.uint8, device="cuda")
h = torch.empty(n, dtype=torch.int64, device="cuda")
for _ in range(reps):
    strings.validate_hash_device(d_data, d_off, 0, verdict=v, hash=h)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
strings.validate_hash_device(d_data, d_off, 0, verdict=v, hash=h)
b.record()
torch.cuda.synchronize()
print(f"k9 n={n} {a.elapsed_time(b):.3f} ms  safe={(v == 0).sum().item()}")
