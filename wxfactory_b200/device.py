"""``B200Device`` -- the reference's backend seam (wx_factory/device/device.py:16-68, CudaDevice :125-220) for this path.

Same method surface and argument meaning as ``Device`` / ``CudaDevice``: ``synchronize``, ``array``, ``pinned``,
``to_host``, ``timestamp``, ``elapsed``, ``has_128_bits_float``, attributes ``comm``, ``xp``, ``main_stream``,
``copy_stream``.  Differences, all deliberate (BASELINE.json north_star excludes cupy):

* arrays are torch CUDA tensors (exchanged with anything else through DLPack); ``xp`` is the ``torch`` module, whose
  creation / elementwise / ``linalg`` functions cover what the integrators and Krylov solvers call on ``device.xp``;
* memory is plain device memory, not the reference's managed-memory pool (device.py:171);
* the device is chosen as ``rank % number of devices`` exactly as the reference does (device.py:166-167), the rank
  coming from ``torch.distributed`` (or the ``RANK`` environment variable) instead of an MPI communicator;
* no CUDA device -> ``ValueError`` with the reference's message (device.py:131-132); there is no host fallback.
"""

import os
from typing import Any, List, Optional

import numpy


class B200Device:
    _default = None

    def __init__(self, comm=None, device_list: Optional[List[int]] = None) -> None:
        import torch

        if not torch.cuda.is_available():
            raise ValueError("Unable to create a CudaDevice object, no GPU devices were detected")
        self.torch = torch
        self.xp = torch
        self.xalg = torch.linalg
        self.comm = comm  # a torch.distributed process group (None = the default group / single process)
        self.pde = None  # the compiled kernels live behind wxfactory_b200.capi, not in a pybind11 module
        ndev = torch.cuda.device_count()
        devs = [x for x in (device_list or []) if x < ndev] or list(range(ndev))
        rank = 0
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            rank = torch.distributed.get_rank(comm)
        else:
            rank = int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", "0")))
        self.index = devs[rank % len(devs)]
        torch.cuda.set_device(self.index)
        self.device = torch.device("cuda", self.index)
        self.main_stream = torch.cuda.current_stream(self.device)
        self.copy_stream = torch.cuda.Stream(device=self.device)
        if B200Device._default is None:
            B200Device._default = self

    # ---- Device interface -------------------------------------------------------------------------------
    def synchronize(self, **kwargs):
        """Host-device barrier on the main stream, or on the copy stream with ``copy_stream=True`` (device.py:181-187)."""
        (self.copy_stream if kwargs.get("copy_stream") else self.main_stream).synchronize()

    def array(self, a, *args, **kwargs):
        """Copy the given array to the device (device.py:189-191).  Accepts numpy arrays and DLPack exporters."""
        t = self.torch
        if isinstance(a, t.Tensor):
            return a.to(self.device)
        if isinstance(a, numpy.ndarray):
            return t.from_numpy(numpy.ascontiguousarray(a)).to(self.device)
        if hasattr(a, "__dlpack__"):
            return t.from_dlpack(a).to(self.device)
        return t.as_tensor(a, device=self.device)

    def pinned(self, *args, **kwargs):
        """Allocate a host array in page-locked memory (device.py:193-195); returned as a numpy array."""
        shape = args[0] if args else kwargs.pop("shape")
        dtype = kwargs.get("dtype", args[1] if len(args) > 1 else numpy.float64)
        tdt = self.torch.from_numpy(numpy.empty(0, dtype=dtype)).dtype
        return self.torch.empty(shape, dtype=tdt, pin_memory=True).numpy()

    def to_host(self, val: Any, **kwargs):
        """Copy the given array to the host (device.py:197-199)."""
        if isinstance(val, self.torch.Tensor):
            return val.detach().cpu().numpy()
        return val

    def timestamp(self, **kwargs):
        """Insert a timing event in the flow of execution (device.py:201-208)."""
        ev = self.torch.cuda.Event(enable_timing=True)
        ev.record(self.copy_stream if kwargs.get("copy_stream") else self.main_stream)
        return ev

    def elapsed(self, timestamps):
        """Seconds between consecutive timestamps, then the total (device.py:210-214)."""
        timestamps[-1].synchronize()
        out = [timestamps[i].elapsed_time(timestamps[i + 1]) / 1000.0 for i in range(len(timestamps) - 1)]
        out.append(timestamps[0].elapsed_time(timestamps[-1]) / 1000.0)
        return out

    def has_128_bits_float(self) -> bool:
        return False

    @staticmethod
    def get_default() -> "B200Device":
        if B200Device._default is None:
            B200Device._default = B200Device()
        return B200Device._default
