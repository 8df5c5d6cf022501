#!/usr/bin/env python
"""Host-side limits of the checkpoint writer on this box: crc32 rate (the zip container needs one
per record), buffered file write + fsync rate, O_DIRECT write rate, for the checkpoint directory's filesystem."""
import json
import os
import sys
import tempfile
import time
import zlib

n = 1 << 30
buf = bytearray(os.urandom(1 << 20) * (n >> 20))
out = {"cpus": os.cpu_count()}
t = time.perf_counter(); zlib.crc32(buf); out["crc32_gbs"] = n / (time.perf_counter() - t) / 1e9
d = sys.argv[1] if len(sys.argv) > 1 else tempfile.gettempdir()
path = os.path.join(d, "pr_io_probe.bin")
mv = memoryview(buf)
t = time.perf_counter()
with open(path, "wb", buffering=0) as f:
    for rep in range(4):
        for o in range(0, n, 64 << 20):
            f.write(mv[o:o + (64 << 20)])
    t1 = time.perf_counter()
    os.fsync(f.fileno())
t2 = time.perf_counter()
out["write_gbs_before_fsync"] = 4 * n / (t1 - t) / 1e9
out["write_gbs_durable"] = 4 * n / (t2 - t) / 1e9
os.remove(path)
try:
    import mmap
    fd = os.open(path, os.O_WRONLY | os.O_CREAT | os.O_DIRECT, 0o644)
    m = mmap.mmap(-1, 64 << 20)
    m.write(bytes(buf[:64 << 20]))
    t = time.perf_counter()
    for _ in range(32):
        os.write(fd, m)
    os.fsync(fd)
    out["odirect_write_gbs"] = 32 * (64 << 20) / (time.perf_counter() - t) / 1e9
    os.close(fd)
    os.remove(path)
except Exception as e:  # noqa: BLE001
    out["odirect_write_gbs"] = f"unavailable: {e}"
st = os.statvfs(d)
out["dir"] = d
out["free_gb"] = st.f_bavail * st.f_frsize / 1e9
try:
    out["mount"] = [l.split()[:3] for l in open("/proc/mounts") if l.split()[1] in ("/", "/tmp", d)]
except Exception:
    pass
print(json.dumps(out))
