#!/usr/bin/env python3
"""One measured line of the UNMODIFIED Python reference (through oracle/ref_shim.py) next to the C port (oracle/) on the
same frames and machine (build container only: needs /root/reference).  VERDICT r1 weak #6: the two CPU figures must
never be confused -- bench.py's cpu_baseline / --impl reference legs time the C port (`kind: port`)."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as orc  # noqa: E402
from oracle import ref_shim  # noqa: E402
from tests import fullsize  # noqa: E402


def best(fn, reps=3):
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return min(ts)


out = {"cpu": open("/proc/cpuinfo").read().split("model name")[1].split("\n")[0].strip(": \t"), "cores": os.cpu_count()}
for name in ("fs_4k_U", "fs_4k_N", "fs_1080p_U"):
    c, s, m = fullsize.inputs(fullsize.CASES[name])
    mp = c.shape[0] * c.shape[1] / 1e6
    t_ref = best(lambda: ref_shim.ref_pipeline(c, s, m))
    t_port = best(lambda: orc.transfer(c, s, m))
    out[name] = {"python_reference_s": round(t_ref, 2), "python_reference_MPps": round(mp / t_ref, 2),
                 "c_port_1thread_s": round(t_port, 2), "c_port_1thread_MPps": round(mp / t_port, 2)}
print(json.dumps(out, indent=1))
