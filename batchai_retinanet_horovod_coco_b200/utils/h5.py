"""Minimal h5py-like facade over the in-tree HDF5 subset (``csrc/rt/hdf5_subset.cpp``).

There is no h5py/libhdf5 on the target image, and the reference stores checkpoints as Keras
``.h5`` files (``train.py:77,113-114,406``), so the format subset Keras uses is implemented
natively and exposed here with numpy arrays.
"""

from __future__ import annotations

import ctypes as C

import numpy as np

from ..parallel import runtime as _rt

_CONFIGURED = False


def _lib():
    global _CONFIGURED
    l = _rt.lib()
    if not _CONFIGURED:
        l.b200h5_open.restype = C.c_void_p
        l.b200h5_open.argtypes = [C.c_char_p]
        l.b200h5_create.restype = C.c_void_p
        l.b200h5_create.argtypes = [C.c_char_p]
        l.b200h5_close.argtypes = [C.c_void_p]
        l.b200h5_last_error.restype = C.c_char_p
        l.b200h5_list.restype = C.c_char_p
        l.b200h5_list.argtypes = [C.c_void_p, C.c_char_p]
        l.b200h5_is_group.argtypes = [C.c_void_p, C.c_char_p]
        l.b200h5_attr_names.restype = C.c_char_p
        l.b200h5_attr_names.argtypes = [C.c_void_p, C.c_char_p]
        l.b200h5_attr_info.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                       C.POINTER(C.c_int), C.POINTER(C.c_longlong)]
        l.b200h5_attr_read.restype = C.c_longlong
        l.b200h5_attr_read.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_void_p, C.c_longlong]
        l.b200h5_dataset_info.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                          C.POINTER(C.c_int), C.POINTER(C.c_longlong)]
        l.b200h5_dataset_read.restype = C.c_longlong
        l.b200h5_dataset_read.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_longlong]
        l.b200h5_make_group.argtypes = [C.c_void_p, C.c_char_p]
        l.b200h5_write_dataset.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int, C.c_int,
                                           C.POINTER(C.c_longlong), C.c_void_p]
        l.b200h5_write_attr.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int,
                                        C.POINTER(C.c_longlong), C.c_void_p]
        _CONFIGURED = True
    return l


def _np_dtype(cls: int, size: int):
    if cls == 0:
        return {2: np.float16, 4: np.float32, 8: np.float64}[size]
    if cls == 1:
        return {1: np.int8, 2: np.int16, 4: np.int32, 8: np.int64}[size]
    if cls == 2:
        return np.dtype(f"S{size}")
    raise TypeError(f"unsupported HDF5 class {cls}")


def _encode(value):
    """python / numpy value -> (cls, elem_size, dims, contiguous numpy array)."""
    if isinstance(value, str):
        value = value.encode("utf8")
    if isinstance(value, bytes):
        arr = np.array(value if value else b"\0", dtype=f"S{max(1, len(value))}")
        return 2, arr.dtype.itemsize, [], arr
    if isinstance(value, (list, tuple)) and value and isinstance(value[0], (bytes, str)):
        items = [v.encode("utf8") if isinstance(v, str) else v for v in value]
        arr = np.array(items, dtype=f"S{max(1, max(len(i) for i in items))}")
        return 2, arr.dtype.itemsize, list(arr.shape), arr
    arr = np.ascontiguousarray(value)
    if arr.dtype.kind == "S":
        return 2, arr.dtype.itemsize, list(arr.shape), arr
    if arr.dtype.kind == "f":
        if arr.dtype.itemsize not in (4, 8):
            arr = arr.astype(np.float32)
        return 0, arr.dtype.itemsize, list(arr.shape), arr
    if arr.dtype.kind in "iub":
        arr = arr.astype(np.int64) if arr.dtype.itemsize not in (4, 8) or arr.dtype.kind != "i" else arr
        return 1, arr.dtype.itemsize, list(arr.shape), arr
    raise TypeError(f"cannot store {arr.dtype} in HDF5 subset")


class H5File:
    def __init__(self, path: str, mode: str = "r"):
        self.path, self.mode = path, mode
        l = _lib()
        self._h = l.b200h5_open(path.encode()) if mode == "r" else l.b200h5_create(path.encode())
        if not self._h:
            raise OSError(f"HDF5: {l.b200h5_last_error().decode()}")

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def close(self):
        if self._h:
            rc = _lib().b200h5_close(self._h)
            self._h = None
            if rc != 0:
                raise OSError(f"HDF5 write failed: {_lib().b200h5_last_error().decode()}")

    def _err(self):
        return OSError(f"HDF5: {_lib().b200h5_last_error().decode()}")

    # ---- reading
    def keys(self, group: str = "/"):
        s = _lib().b200h5_list(self._h, group.encode())
        if s is None:
            raise KeyError(group)
        s = s.decode()
        return s.split("\n") if s else []

    def exists(self, path: str) -> bool:
        return _lib().b200h5_is_group(self._h, path.encode()) >= 0

    def is_group(self, path: str) -> bool:
        return _lib().b200h5_is_group(self._h, path.encode()) == 1

    def attr_names(self, path: str = "/"):
        s = _lib().b200h5_attr_names(self._h, path.encode())
        if s is None:
            raise KeyError(path)
        s = s.decode()
        return s.split("\n") if s else []

    def attr(self, path: str, name: str):
        l = _lib()
        cls, es, rank = C.c_int(), C.c_int(), C.c_int()
        dims = (C.c_longlong * 8)()
        if l.b200h5_attr_info(self._h, path.encode(), name.encode(), cls, es, rank, dims) != 0:
            raise KeyError(f"{path}@{name}")
        n = l.b200h5_attr_read(self._h, path.encode(), name.encode(), None, 0)
        buf = (C.c_char * max(1, n))()
        l.b200h5_attr_read(self._h, path.encode(), name.encode(), buf, n)
        raw = bytes(buf.raw[:n])
        shape = tuple(dims[i] for i in range(rank.value))
        if cls.value == 3:
            parts = raw.split(b"\0")[:-1]
            return parts[0] if not shape else np.array(parts, dtype=object).reshape(shape)
        arr = np.frombuffer(raw, dtype=_np_dtype(cls.value, es.value)).reshape(shape)
        if cls.value == 2:
            return bytes(arr.item()).rstrip(b"\0") if not shape else np.char.rstrip(arr, b"\0")
        return arr.item() if not shape else arr.copy()

    def dataset(self, path: str) -> np.ndarray:
        l = _lib()
        cls, es, rank = C.c_int(), C.c_int(), C.c_int()
        dims = (C.c_longlong * 8)()
        if l.b200h5_dataset_info(self._h, path.encode(), cls, es, rank, dims) != 0:
            raise KeyError(path)
        shape = tuple(dims[i] for i in range(rank.value))
        out = np.empty(shape, dtype=_np_dtype(cls.value, es.value))
        if l.b200h5_dataset_read(self._h, path.encode(), out.ctypes.data_as(C.c_void_p), out.nbytes) < 0:
            raise self._err()
        return out

    # ---- writing
    def make_group(self, path: str):
        if _lib().b200h5_make_group(self._h, path.encode()) != 0:
            raise self._err()

    def write_dataset(self, path: str, value):
        cls, es, dims, arr = _encode(value)
        d = (C.c_longlong * max(1, len(dims)))(*dims)
        if _lib().b200h5_write_dataset(self._h, path.encode(), cls, es, len(dims), d, arr.ctypes.data_as(C.c_void_p)) != 0:
            raise self._err()

    def write_attr(self, path: str, name: str, value):
        cls, es, dims, arr = _encode(value)
        d = (C.c_longlong * max(1, len(dims)))(*dims)
        if _lib().b200h5_write_attr(self._h, path.encode(), name.encode(), cls, es, len(dims), d,
                                    arr.ctypes.data_as(C.c_void_p)) != 0:
            raise self._err()
