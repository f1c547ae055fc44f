"""Drop-in for the reference's pybind module `wlqppy`
(template/wlcontroller/pybind/wlqppy.cpp:15-19): WLController(popts, controlRate).update(u0, h0, pdotdes)
-> (u[4], w0[6]).  Like the reference C wrapper (wlcontroller/c/wlcontroller.c:35-45) the controller
state is process-global; `u0` is only used on the first call."""
import numpy as np

from . import _lib
from .uprightmpc2py import _fp, _v


class WLController:
    def __init__(self, popts, controlRate):
        self._lib = _lib.load()
        self._popts = _v(popts, 90)
        self._rate = float(controlRate)
        self._lib.wlControllerReset()

    def update(self, u0, h0, pdotdes):
        u, w0 = np.zeros(4, np.float32), np.zeros(6, np.float32)
        self._lib.wlControllerUpdateW(_fp(u), _fp(w0), _fp(_v(u0, 4)), _fp(_v(h0, 6)), _fp(_v(pdotdes, 6)),
                                      _fp(self._popts), self._rate)
        if np.isnan(u[0]):
            raise RuntimeError("wlControllerUpdate failed: %s" % self._lib.umpcLastError().decode())
        return u, w0
