"""``TOD`` with maria's interface for the synthesis path (maria/tod/tod.py:34-154): ``data[field]`` (n_det, n_t) in
``dtype``, ``weight``, ``coords``, ``units``, ``dets``, ``metadata``, ``.signal``, ``.to(units)``.

The arrays live on the GPU as torch tensors (``device_data``); ``data[field]`` hands out a host copy on demand, and the
float64 all-ones ``weight`` of the reference (tod.py:68-73) is materialised only when read.  ``to("K_RJ")`` /
``to("pW")`` run the calibration kernel (calibration/functions.py:48-90 through band.py:235-255) on the device.
"""
from __future__ import annotations

import numpy as np

from . import _lib

_UNIT_CODES = {"pW": _lib.UNITS_PW, "K_RJ": _lib.UNITS_KRJ}


class _HostView(dict):
    """``tod.data``: field -> host ndarray, copied from the device when first read."""

    def __init__(self, tod):
        super().__init__()
        self._tod = tod

    def __getitem__(self, field):
        if not dict.__contains__(self, field):
            dict.__setitem__(self, field, self._tod.device_data[field].cpu().numpy())
        return dict.__getitem__(self, field)

    def keys(self):
        return self._tod.device_data.keys()

    def __iter__(self):
        return iter(self._tod.device_data)

    def __contains__(self, field):
        return field in self._tod.device_data

    def __len__(self):
        return len(self._tod.device_data)

    def items(self):
        return [(k, self[k]) for k in self.keys()]


class TOD:
    def __init__(self, data, weight=None, coords=None, units="K_RJ", dets=None, distributed=True, dtype=np.float32,
                 metadata=None, calibrator=None):
        import torch

        if units not in _UNIT_CODES:
            raise ValueError(f"Invalid units '{units}': the synthesis path supports {sorted(_UNIT_CODES)}.")
        self.coords, self.dets, self.units, self.dtype = coords, dets, units, np.dtype(dtype)
        self.metadata = dict(metadata or {})
        self.distributed = distributed
        self._calibrator = calibrator
        tdtype = torch.float64 if self.dtype == np.float64 else torch.float32
        self.device_data = {}
        for field in sorted(data):  # fields are kept sorted, tod.py:66
            v = data[field]
            if not isinstance(v, torch.Tensor):
                if calibrator is None:
                    raise RuntimeError("TOD needs device tensors or a calibrator bound to a device")
                v = torch.as_tensor(np.asarray(v)).to(f"cuda:{calibrator.ctx.device}")
            self.device_data[field] = v.to(tdtype)
        self.data = _HostView(self)
        self._weight = weight

    # ---- shape / content -------------------------------------------------------------------------------------
    @property
    def fields(self):
        return list(self.device_data.keys())

    @property
    def shape(self):
        return tuple(next(iter(self.device_data.values())).shape)

    @property
    def weight(self):
        if self._weight is None:
            # tod.py:68-73: float64 ones of the TOD's shape -- as a zero-stride, read-only view (144 GB at C4 otherwise)
            self._weight = np.broadcast_to(np.ones((), dtype=np.float64), self.shape)
        return self._weight

    def sample(self, field, dets=slice(None), samples=slice(None)):
        """``data[field][dets, samples]`` as a host array without copying the whole field off the device."""
        return self.device_data[field][dets, samples].cpu().numpy()

    @property
    def signal_device(self):
        return sum(self.device_data.values())

    @property
    def signal(self):
        """Sum of all fields (tod.py:152-154), as a host array."""
        return self.signal_device.cpu().numpy()

    @property
    def content(self):
        return dict(data=dict(self.device_data), weight=self._weight, coords=self.coords, units=self.units, dets=self.dets,
                    dtype=self.dtype, metadata=self.metadata, calibrator=self._calibrator)

    # ---- units -----------------------------------------------------------------------------------------------
    def to(self, units):
        """tod.py:106-142: returns a new TOD in ``units``; per band pW <-> K_RJ with the transmission integral looked up
        at (round(T0, 3), round(pwv, 3), full-rate elevation)."""
        if units not in _UNIT_CODES:
            raise ValueError(f"Invalid units '{units}': the synthesis path supports {sorted(_UNIT_CODES)}.")
        content = self.content
        if units == self.units:
            content["data"] = {k: v.clone() for k, v in self.device_data.items()}
            return TOD(**content)
        atm = self._calibrator
        if atm is None:
            raise RuntimeError("this TOD carries no calibrator (atmosphere with band tables): cannot convert units")
        ctx = atm.ctx
        if np.dtype(ctx.np_dtype) != self.dtype:
            raise RuntimeError(f"TOD dtype {self.dtype} does not match the device context precision {ctx.precision}")
        det0 = int(self.metadata.get("det0", 0))
        out = {}
        for field, v in self.device_data.items():
            o = ctx.torch.empty_like(v)
            ctx.check(ctx.lib.matm_calibrate(ctx.h, det0, v.shape[0], v.data_ptr(), o.data_ptr(), _UNIT_CODES[units],
                                             ctx.stream_ptr()))
            out[field] = o
        content["data"], content["units"] = out, units
        return TOD(**content)
