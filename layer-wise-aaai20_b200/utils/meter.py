"""Running meters (reference IMAGENET/training/meter.py:4-86)."""
from __future__ import annotations

import time
from typing import Tuple


class AverageMeter:
    def __init__(self):
        self.reset()

    def reset(self):
        self.val = self.avg = self.sum = 0.0
        self.count = 0.0

    def update(self, val, n=1):
        self.val = float(val)
        self.sum += float(val) * n
        self.count += n
        self.avg = self.sum / self.count if self.count else 0.0


def network_bytes() -> Tuple[int, int]:
    """(received, transmitted) bytes summed over non-loopback NICs from /proc/net/dev.

    Measures Ethernet-class traffic only; NVLink traffic is invisible here -- use the
    exchange engine's byte counters for that.
    """
    recv = xmit = 0
    try:
        with open("/proc/net/dev") as f:
            lines = f.readlines()[2:]
    except OSError:
        return 0, 0
    for line in lines:
        if ":" not in line:
            continue
        name, rest = line.split(":", 1)
        if name.strip() == "lo":
            continue
        cols = rest.split()
        if len(cols) >= 9:
            recv += int(cols[0])
            xmit += int(cols[8])
    return recv, xmit


class NetworkMeter:
    def __init__(self):
        self.recv_meter = AverageMeter()
        self.transmit_meter = AverageMeter()
        self.last_recv, self.last_transmit = network_bytes()
        self.last_log = time.time()

    def update_bandwidth(self):
        now = time.time()
        recv, xmit = network_bytes()
        dt = max(now - self.last_log, 1e-9)
        recv_gbit = 8 * (recv - self.last_recv) / dt / 1e9
        xmit_gbit = 8 * (xmit - self.last_transmit) / dt / 1e9
        self.last_recv, self.last_transmit, self.last_log = recv, xmit, now
        self.recv_meter.update(recv_gbit)
        self.transmit_meter.update(xmit_gbit)
        return recv_gbit, xmit_gbit


class TimeMeter:
    """Data-loading time vs whole-batch time, wall clock.

    The reference notes that async CUDA work escapes this measurement
    (train_imagenet_nv.py:402); device-side numbers come from utils.timing.PhaseTimer.
    """

    def __init__(self):
        self.batch_time = AverageMeter()
        self.data_time = AverageMeter()
        self.start = time.time()

    def batch_start(self):
        self.data_time.update(time.time() - self.start)

    def batch_end(self):
        self.batch_time.update(time.time() - self.start)
        self.start = time.time()
