"""`@on_device`: run a method with its object's CUDA device current. The C ABI launches on the stream it is handed,
and kernel attributes / per-device caches inside the library belong to the CURRENT device, so a process that drives
several GPUs (device= arguments of the services) must switch before every entry point. A no-op when the device is
already current (the one-process-per-GPU layout of bench.py)."""
from __future__ import annotations

import functools

import torch


def on_device(fn):
    @functools.wraps(fn)
    def wrapper(self, *args, **kwargs):
        dev = getattr(self, "device", None)
        if dev is None:
            dev = self.eng.device
        if dev.index is None or torch.cuda.current_device() == dev.index:
            return fn(self, *args, **kwargs)
        with torch.cuda.device(dev):
            return fn(self, *args, **kwargs)
    return wrapper
