"""Minimal client of the reference's remote-UI wire protocol, for the tests only.

Framing and ids follow external/packetcomms (PacketMuxer.cpp:120-138, IdManager.h:18-31): a packet is
u32 type | u32 size (both network order) | payload; type 1 is a one-byte control packet (254 Hello first,
0 heart beat, 255 good bye); application types are 2 + the index in the table of src/InterfaceServer.hpp:23-42.
Payloads are cereal binary archives: little-endian scalars, strings and vectors prefixed by a u64 count.
"""
import socket
import struct
import threading
import time

import numpy as np

PACKETS = ["stop", "detach", "progress", "sample_rate", "env_rotation", "exposure", "gamma", "fov",
           "load_nif", "render_preview", "hdr_header", "hdr_packet", "interactive_samples"]
HEARTBEAT, HELLO, GOODBYE = 0, 254, 255


class UiClient:
    def __init__(self, port, host="127.0.0.1", connect_timeout=20.0):
        deadline = time.time() + connect_timeout
        while True:
            try:
                self.sock = socket.create_connection((host, port), timeout=2.0)
                break
            except OSError:
                if time.time() > deadline:
                    raise
                time.sleep(0.05)
        self.sock.settimeout(None)
        self.lock = threading.Lock()
        self.progress = []
        self.rates = []
        self.images = []          # completed (height, width, 3) float32 RGB arrays
        self.control = []
        self.other = []
        self._hdr = None
        self.closed = threading.Event()
        self._send(1, bytes([HELLO]))
        self.thread = threading.Thread(target=self._receive, daemon=True)
        self.thread.start()

    def _send(self, ptype, payload):
        self.sock.sendall(struct.pack(">II", ptype, len(payload)) + payload)

    def send(self, name, payload):
        self._send(2 + PACKETS.index(name), payload)

    def send_float(self, name, v):
        self.send(name, struct.pack("<f", v))

    def send_bool(self, name, v=True):
        self.send(name, struct.pack("<B", 1 if v else 0))

    def send_u32(self, name, v):
        self.send(name, struct.pack("<I", v))

    def send_string(self, name, s):
        b = s.encode()
        self.send(name, struct.pack("<Q", len(b)) + b)

    def _read(self, n):
        buf = b""
        while len(buf) < n:
            chunk = self.sock.recv(n - len(buf))
            if not chunk:
                raise ConnectionError("server hung up")
            buf += chunk
        return buf

    def _receive(self):
        try:
            while True:
                ptype, size = struct.unpack(">II", self._read(8))
                payload = self._read(size) if size else b""
                with self.lock:
                    self._dispatch(ptype, payload)
        except (ConnectionError, OSError):
            pass
        finally:
            self.closed.set()

    def _dispatch(self, ptype, payload):
        if ptype == 1:
            self.control.append(payload[0])
            return
        name = PACKETS[ptype - 2]
        if name == "progress":
            self.progress.append(struct.unpack("<f", payload)[0])
        elif name == "sample_rate":
            self.rates.append(struct.unpack("<ff", payload))
        elif name == "hdr_header":
            w, h, packets = struct.unpack("<iiI", payload)
            self._hdr = {"w": w, "h": h, "packets": packets, "rows": {}}
        elif name == "hdr_packet":
            idx, count = struct.unpack("<IQ", payload[:12])
            row = np.frombuffer(payload[12:], dtype="<f4")
            assert row.size == count
            hdr = self._hdr
            hdr["rows"][idx] = row
            if len(hdr["rows"]) == hdr["packets"]:
                img = np.stack([hdr["rows"][i] for i in range(hdr["packets"])]).reshape(hdr["h"], hdr["w"], 3)
                self.images.append(img)
        else:
            self.other.append((name, payload))

    def wait(self, predicate, timeout=30.0):
        deadline = time.time() + timeout
        while time.time() < deadline:
            with self.lock:
                if predicate(self):
                    return True
            if self.closed.is_set():
                with self.lock:
                    return bool(predicate(self))
            time.sleep(0.01)
        return False

    def close(self):
        try:
            self.sock.shutdown(socket.SHUT_RDWR)
        except OSError:
            pass
        self.sock.close()
