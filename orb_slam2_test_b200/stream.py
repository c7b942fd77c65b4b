"""Python mirror of the pipelined sequence feeder of include/borb.h (borb_stream_*).

One `OrbStream` per GPU: a ring of batches whose uploads, kernels and downloads overlap on three
CUDA streams inside the library (csrc/borb_stream.inl).  It stands where the reference's drivers
loop over a sequence one frame at a time (Examples/Stereo/stereo_kitti.cc:66-108 -> Frame::Frame,
src/Frame.cc:79-151): with `stereo=True` an item is a left/right pair, both eyes are extracted and
Frame::ComputeStereoMatches (src/Frame.cc:619-834) runs on the device before anything comes back.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .extractor import KP_DTYPE, STREAM_SYMBOLS, BorbError, _check, _vp, load_library

BORB_STREAM_STEREO = 1
BORB_STREAM_PYRAMID = 2

PRODUCE_FN = C.CFUNCTYPE(C.c_int, _vp, C.c_uint64, _vp, _vp, C.c_int)
CONSUME_FN = C.CFUNCTYPE(None, _vp, _vp, C.c_uint64, C.c_int)


_bound = False


def _bind(lib) -> None:
    global _bound
    if _bound:
        return
    lib.borb_stream_open.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                     C.c_int, C.c_int, C.c_uint, C.POINTER(_vp)]
    lib.borb_stream_close.argtypes = [_vp]
    lib.borb_stream_close.restype = None
    lib.borb_stream_set_stereo.argtypes = [_vp, C.c_float, C.c_float]
    lib.borb_stream_set_camera.argtypes = [_vp, _vp, _vp, C.c_int]
    lib.borb_stream_input.argtypes = [_vp, C.c_int, C.POINTER(_vp), C.POINTER(C.c_size_t)]
    lib.borb_stream_push.argtypes = [_vp, _vp, _vp, C.c_int, C.c_size_t, C.c_size_t, C.c_uint64]
    lib.borb_stream_pop.argtypes = [_vp, C.POINTER(C.c_uint64), C.POINTER(C.c_int)]
    lib.borb_stream_release.argtypes = [_vp]
    lib.borb_stream_in_flight.argtypes = [_vp]
    lib.borb_stream_slots_free.argtypes = [_vp]
    lib.borb_stream_result.argtypes = [_vp, C.c_int, C.c_int, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(C.c_int)]
    lib.borb_stream_stereo_result.argtypes = [_vp, C.c_int, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(C.c_int),
                                              C.POINTER(C.c_int)]
    lib.borb_stream_frame_grid.argtypes = [_vp, C.c_int, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), _vp]
    lib.borb_stream_pyramid.argtypes = [_vp, C.c_int, C.c_int, C.c_int, C.POINTER(_vp), C.POINTER(C.c_int),
                                        C.POINTER(C.c_int), C.POINTER(C.c_size_t)]
    lib.borb_stream_link_bytes.argtypes = [_vp, _vp]
    lib.borb_stream_launches_per_push.argtypes = [_vp]
    lib.borb_stream_start.argtypes = [_vp, PRODUCE_FN, CONSUME_FN, _vp]
    lib.borb_stream_join.argtypes = [_vp]
    _bound = True


class OrbStream:
    """Ring of `slots` batches of `batch` items (frames, or stereo pairs) on one GPU."""

    def __init__(self, nfeatures: int, scaleFactor: float, nlevels: int, iniThFAST: int, minThFAST: int,
                 cols: int, rows: int, batch: int, slots: int = 4, device: int = -1, stereo: bool = False,
                 want_pyramid: bool = False):
        self._lib = load_library()
        _bind(self._lib)
        self._s = _vp()
        flags = (BORB_STREAM_STEREO if stereo else 0) | (BORB_STREAM_PYRAMID if want_pyramid else 0)
        _check(self._lib, self._lib.borb_stream_open(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device,
                                                     cols, rows, batch, slots, flags, C.byref(self._s)))
        self.cols, self.rows, self.batch, self.slots = cols, rows, batch, slots
        self.eyes = 2 if stereo else 1
        self.nlevels = nlevels
        self._cb = None

    def close(self) -> None:
        if getattr(self, "_s", None):
            self._lib.borb_stream_close(self._s)
            self._s = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stereo(self, mbf: float, mb: float) -> None:
        _check(self._lib, self._lib.borb_stream_set_stereo(self._s, mbf, mb))

    def set_camera(self, K=None, dist=()) -> None:
        if K is None:
            _check(self._lib, self._lib.borb_stream_set_camera(self._s, None, None, 0))
            return
        k = np.asarray(K, np.float32)
        d = np.asarray(dist, np.float32)
        _check(self._lib, self._lib.borb_stream_set_camera(self._s, k.ctypes.data, d.ctypes.data if len(d) else None, len(d)))

    # -- producer side --
    def input_buffer(self, eye: int = 0) -> np.ndarray:
        """[batch, rows, cols] uint8 view of the page-locked input buffer the next push will use."""
        p, cap = _vp(), C.c_size_t()
        _check(self._lib, self._lib.borb_stream_input(self._s, eye, C.byref(p), C.byref(cap)))
        buf = (C.c_uint8 * cap.value).from_address(p.value)
        return np.frombuffer(buf, np.uint8).reshape(self.batch, self.rows, self.cols)

    def push(self, left: np.ndarray | None, right: np.ndarray | None = None, tag: int = 0, n: int | None = None) -> None:
        """left / right: [n, rows, cols] uint8 C-contiguous; None = the frames are in input_buffer()."""
        if left is None:
            _check(self._lib, self._lib.borb_stream_push(self._s, None, None, int(n), 0, 0, tag))
            return
        assert left.dtype == np.uint8 and left.flags.c_contiguous and left.shape[1:] == (self.rows, self.cols)
        if right is not None:
            assert right.dtype == np.uint8 and right.flags.c_contiguous and right.shape == left.shape
        _check(self._lib, self._lib.borb_stream_push(self._s, left.ctypes.data, right.ctypes.data if right is not None else None,
                                                     left.shape[0], self.cols, self.rows * self.cols, tag))

    def push_ptr(self, left_ptr: int, right_ptr: int | None, n: int, tag: int = 0) -> None:
        """Densely packed frames at raw host addresses (page-locked memory is uploaded in place)."""
        _check(self._lib, self._lib.borb_stream_push(self._s, left_ptr, right_ptr, n, self.cols, self.rows * self.cols, tag))

    # -- consumer side --
    def pop(self):
        """Blocks for the oldest batch; -> (tag, n).  Its results stay readable until the next pop / release."""
        tag, n = C.c_uint64(), C.c_int()
        _check(self._lib, self._lib.borb_stream_pop(self._s, C.byref(tag), C.byref(n)))
        return tag.value, n.value

    def release(self) -> None:
        _check(self._lib, self._lib.borb_stream_release(self._s))

    def in_flight(self) -> int:
        return self._lib.borb_stream_in_flight(self._s)

    def slots_free(self) -> int:
        return self._lib.borb_stream_slots_free(self._s)

    def result(self, item: int, eye: int = 0, copy: bool = True):
        kp, de, n = _vp(), _vp(), C.c_int()
        _check(self._lib, self._lib.borb_stream_result(self._s, eye, item, C.byref(kp), C.byref(de), C.byref(n)))
        cnt = n.value
        if cnt == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        kps = np.frombuffer((C.c_uint8 * (cnt * 28)).from_address(kp.value), dtype=KP_DTYPE)
        desc = np.frombuffer((C.c_uint8 * (cnt * 32)).from_address(de.value), dtype=np.uint8).reshape(cnt, 32)
        return (kps.copy(), desc.copy()) if copy else (kps, desc)

    def count(self, item: int, eye: int = 0) -> int:
        n = C.c_int()
        _check(self._lib, self._lib.borb_stream_result(self._s, eye, item, None, None, C.byref(n)))
        return n.value

    def stereo_result(self, item: int, copy: bool = True):
        u, z, n, m = _vp(), _vp(), C.c_int(), C.c_int()
        _check(self._lib, self._lib.borb_stream_stereo_result(self._s, item, C.byref(u), C.byref(z), C.byref(n), C.byref(m)))
        cnt = n.value
        if cnt == 0:
            return np.zeros(0, np.float32), np.zeros(0, np.float32), 0
        ur = np.frombuffer((C.c_uint8 * (cnt * 4)).from_address(u.value), dtype=np.float32)
        dz = np.frombuffer((C.c_uint8 * (cnt * 4)).from_address(z.value), dtype=np.float32)
        return (ur.copy(), dz.copy(), m.value) if copy else (ur, dz, m.value)

    def frame_grid(self, item: int):
        xy, cs, ci = _vp(), _vp(), _vp()
        bounds = np.zeros(4, np.float32)
        _check(self._lib, self._lib.borb_stream_frame_grid(self._s, item, C.byref(xy), C.byref(cs), C.byref(ci), bounds.ctypes.data))
        n = self.count(item)
        start = np.ctypeslib.as_array(C.cast(cs, C.POINTER(C.c_int32)), shape=(64 * 48 + 1,)).copy()
        un = (np.ctypeslib.as_array(C.cast(xy, C.POINTER(C.c_float)), shape=(n, 2)).copy() if n else np.zeros((0, 2), np.float32))
        m = int(start[-1])
        idx = (np.ctypeslib.as_array(C.cast(ci, C.POINTER(C.c_int32)), shape=(m,)).copy() if m else np.zeros(0, np.int32))
        return un, start, idx, bounds

    def pyramid_level(self, item: int, level: int, eye: int = 0) -> np.ndarray:
        data, c, r, st = _vp(), C.c_int(), C.c_int(), C.c_size_t()
        _check(self._lib, self._lib.borb_stream_pyramid(self._s, eye, item, level, C.byref(data), C.byref(c), C.byref(r), C.byref(st)))
        buf = (C.c_uint8 * (r.value * st.value)).from_address(data.value)
        return np.lib.stride_tricks.as_strided(np.frombuffer(buf, np.uint8), shape=(r.value, c.value), strides=(st.value, 1)).copy()

    def link_bytes(self):
        """(host->device, device->host) bytes one full batch moves."""
        b = (C.c_uint64 * 2)()
        _check(self._lib, self._lib.borb_stream_link_bytes(self._s, b))
        return int(b[0]), int(b[1])

    def launches_per_push(self) -> int:
        return self._lib.borb_stream_launches_per_push(self._s)

    # -- the library's own feeder thread --
    def run(self, produce, consume) -> None:
        """produce(index, left [batch,rows,cols], right or None) -> items written (0 = end);
        consume(stream, tag, n) for every finished batch in order.  Both are called on the library's feeder
        thread; this call returns when the sequence has ended."""
        err: list = []

        def _p(_u, index, lp, rp, mx):
            try:
                left = np.frombuffer((C.c_uint8 * (mx * self.rows * self.cols)).from_address(lp), np.uint8).reshape(mx, self.rows, self.cols)
                right = (np.frombuffer((C.c_uint8 * (mx * self.rows * self.cols)).from_address(rp), np.uint8).reshape(mx, self.rows, self.cols)
                         if rp else None)
                return int(produce(int(index), left, right))
            except BaseException as e:  # noqa: BLE001 -- must not unwind through C
                err.append(e)
                return -1

        def _c(_u, _s, tag, n):
            try:
                if not err:
                    consume(self, int(tag), int(n))
            except BaseException as e:  # noqa: BLE001
                err.append(e)

        self._cb = (PRODUCE_FN(_p), CONSUME_FN(_c))
        _check(self._lib, self._lib.borb_stream_start(self._s, self._cb[0], self._cb[1], None))
        rc = self._lib.borb_stream_join(self._s)
        self._cb = None
        if err:
            raise err[0]
        _check(self._lib, rc)


__all__ = ["OrbStream", "BorbError", "STREAM_SYMBOLS"]
