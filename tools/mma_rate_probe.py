#!/usr/bin/env python
"""tcgen05.mma issue rate by operand layout (no data movement): cycles per M=128, K=16 MMA for N in {64, 128, 256} and
every combination of K-major / MN-major A and B operands in shared memory.  One JSON line."""
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import layerwise_b200  # noqa: E402,F401
from layerwise_b200 import _native  # noqa: E402


def main():
    lib = _native.lib()
    torch.cuda.init()
    torch.zeros(1, device="cuda")
    lib.lwc_probe_mma_rate.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double)]
    lib.lwc_probe_mma_rate.restype = C.c_int
    out = {}
    for n in (64, 128, 256):
        for a_mn in (0, 1):
            for b_mn in (0, 1):
                v = C.c_double(0)
                for _ in range(2):          # second run is the measurement
                    rc = lib.lwc_probe_mma_rate(a_mn, b_mn, n, 2000, C.byref(v))
                    assert rc == 0, rc
                out[f"N{n}_A{'mn' if a_mn else 'k'}_B{'mn' if b_mn else 'k'}_cycles_per_mma"] = round(v.value, 1)
    out["ideal_cycles_per_mma"] = {"N64": 32, "N128": 64, "N256": 128}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
