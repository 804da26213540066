"""ctypes front-end of oracle/_ref/libref_uiclient.so: a remote-UI client made of the REFERENCE's packetcomms + cereal
(compiled unmodified by oracle/Makefile, with the packet table and wire structs cut out of src/InterfaceServer.hpp:21-83
at build time). Same surface as tests/ui_client.py's UiClient so that the protocol tests run against both.
TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "_ref", "libref_uiclient.so")


def available() -> bool:
    return os.path.exists(LIB)


class _Dummy:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class RefUiClient:
    def __init__(self, port, host="127.0.0.1", connect_timeout=20.0):
        self.lib = C.CDLL(LIB)
        self.lib.refui_connect.restype = C.c_void_p
        self.lib.refui_connect.argtypes = [C.c_char_p, C.c_int, C.c_double]
        for f in ("refui_ok", "refui_counts", "refui_progress", "refui_rates", "refui_image", "refui_send_float", "refui_send_bool",
                  "refui_send_u32", "refui_send_string", "refui_close"):
            getattr(self.lib, f).argtypes = None
        self.lib.refui_packet_type.restype = C.c_char_p
        self.h = self.lib.refui_connect(host.encode(), port, float(connect_timeout))
        if not self.h:
            raise ConnectionError(f"reference client could not connect to {host}:{port}")
        self.h = C.c_void_p(self.h)
        self.lock = _Dummy()              # the library locks internally
        self.closed = threading.Event()

    def packet_types(self):
        return [self.lib.refui_packet_type(i).decode() for i in range(self.lib.refui_packet_type_count())]

    def _counts(self):
        p, r, i = C.c_int(), C.c_int(), C.c_int()
        self.lib.refui_counts(self.h, C.byref(p), C.byref(r), C.byref(i))
        return p.value, r.value, i.value

    @property
    def progress(self):
        n = self._counts()[0]
        buf = np.zeros(max(n, 1), np.float32)
        n = self.lib.refui_progress(self.h, buf.ctypes.data_as(C.c_void_p), n)
        return [float(x) for x in buf[:n]]

    @property
    def rates(self):
        n = self._counts()[1]
        buf = np.zeros((max(n, 1), 2), np.float32)
        n = self.lib.refui_rates(self.h, buf.ctypes.data_as(C.c_void_p), n)
        return [(float(a), float(b)) for a, b in buf[:n]]

    def image(self, index):
        w, h = C.c_int(), C.c_int()
        if self.lib.refui_image(self.h, index, None, C.c_size_t(0), C.byref(w), C.byref(h)) != 0:
            raise IndexError(index)
        out = np.zeros((h.value, w.value, 3), np.float32)
        rc = self.lib.refui_image(self.h, index, out.ctypes.data_as(C.c_void_p), C.c_size_t(out.size), C.byref(w), C.byref(h))
        assert rc == 0
        return out

    @property
    def images(self):
        return [self.image(i) for i in range(self._counts()[2])]

    def ok(self):
        return bool(self.lib.refui_ok(self.h))

    def send_float(self, name, v):
        self.lib.refui_send_float(self.h, name.encode(), C.c_float(v))

    def send_bool(self, name, v=True):
        self.lib.refui_send_bool(self.h, name.encode(), int(bool(v)))

    def send_u32(self, name, v):
        self.lib.refui_send_u32(self.h, name.encode(), C.c_uint32(v))

    def send_string(self, name, s):
        self.lib.refui_send_string(self.h, name.encode(), s.encode())

    def wait(self, predicate, timeout=30.0):
        deadline = time.time() + timeout
        while time.time() < deadline:
            if predicate(self):
                return True
            if not self.ok():
                self.closed.set()
                return bool(predicate(self))
            time.sleep(0.02)
        return False

    def wait_closed(self, timeout=10.0):
        deadline = time.time() + timeout
        while time.time() < deadline:
            if not self.ok():
                self.closed.set()
                return True
            time.sleep(0.02)
        return False

    def close(self):
        if self.h:
            self.lib.refui_close(self.h)
            self.h = None
