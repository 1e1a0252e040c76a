"""File set, manifest and transfer plan -- ctypes twins of the tpimd5_filter_* / tpimd5_manifest_* /
tpimd5_plan_transfer entry points (csrc/manifest.cc).

Names follow what they stand in for on the Go side: ``Filter`` is the rule list machine.Transfer
builds (/root/reference/task/common/machine/storage.go:126-138) with rclone's glob semantics,
``limit_subdir`` is LimitTransfer (storage.go:267-280), ``plan_transfer`` is rclone's
march + NeedTransfer/equal() decision for sync.CopyDir (storage.go:158).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Iterable, List, Optional, Sequence

import numpy as np

from .binding import Context, _check, load_library

MODE_MODTIME, MODE_CHECKSUM, MODE_SIZE_ONLY = 0, 1, 2
ACT_SKIP, ACT_TRANSFER_NEW, ACT_TRANSFER, ACT_FIX_MTIME = 0, 1, 2, 3


class Filter:
    def __init__(self, exclude: Optional[Iterable[str]] = None, limit_subdir: Optional[str] = None):
        self._L = load_library()
        rules = [os.fsencode(r) for r in (exclude or [])]
        arr = (C.c_char_p * max(1, len(rules)))(*rules)
        self._h = C.c_void_p()
        _check(self._L.tpimd5_filter_new(C.cast(arr, C.c_void_p), len(rules),
                                         None if limit_subdir is None else os.fsencode(limit_subdir), C.byref(self._h)))

    def include(self, rel_path: str) -> bool:
        r = self._L.tpimd5_filter_include(self._h, os.fsencode(rel_path))
        if r < 0:
            _check(r)
        return r == 1

    def include_dir(self, rel_dir: str) -> bool:
        """Would rclone descend into (and create) this directory under these rules?"""
        r = self._L.tpimd5_filter_include_dir(self._h, os.fsencode(rel_dir))
        if r < 0:
            _check(r)
        return r == 1

    def rules(self) -> List[str]:
        n = self._L.tpimd5_filter_rule_count(self._h)
        return [self._L.tpimd5_filter_rule(self._h, i).decode() for i in range(n)]

    def close(self):
        if self._h:
            self._L.tpimd5_filter_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Manifest:
    def __init__(self, handle):
        self._L = load_library()
        self._h = handle

    @classmethod
    def scan(cls, ctx: Context, root: str, flt: Optional[Filter] = None) -> "Manifest":
        """Walk `root`, keep what the filter keeps, MD5 every file on the GPU."""
        h = C.c_void_p()
        _check(load_library().tpimd5_manifest_scan(ctx._h, os.fsencode(root), flt._h if flt else None, C.byref(h)))
        return cls(h)

    @classmethod
    def from_arrays(cls, paths: Sequence[str], sizes, mtimes_ns=None, md5s=None, has_md5=None) -> "Manifest":
        n = len(paths)
        arr = (C.c_char_p * max(1, n))(*[os.fsencode(p) for p in paths])
        sz = np.ascontiguousarray(sizes, dtype=np.uint64)
        mt = None if mtimes_ns is None else np.ascontiguousarray(mtimes_ns, dtype=np.int64)
        md = None if md5s is None else np.ascontiguousarray(md5s, dtype=np.uint8).reshape(-1)
        hm = None if has_md5 is None else np.ascontiguousarray(has_md5, dtype=np.uint8)
        h = C.c_void_p()
        _check(load_library().tpimd5_manifest_from_arrays(
            C.cast(arr, C.c_void_p), sz.ctypes.data if n else None, None if mt is None else mt.ctypes.data,
            None if md is None else md.ctypes.data, None if hm is None else hm.ctypes.data, n, C.byref(h)))
        return cls(h)

    @classmethod
    def load(cls, file: str) -> "Manifest":
        h = C.c_void_p()
        _check(load_library().tpimd5_manifest_load(os.fsencode(file), C.byref(h)))
        return cls(h)

    def save(self, file: str) -> None:
        _check(self._L.tpimd5_manifest_save(self._h, os.fsencode(file)))

    def write_md5sum(self, file: str) -> None:
        _check(self._L.tpimd5_manifest_write_md5sum(self._h, os.fsencode(file)))

    def __len__(self) -> int:
        return int(self._L.tpimd5_manifest_count(self._h))

    def entry(self, i: int):
        """-> (path, size, mtime_ns, md5 bytes or None)"""
        p, s, m, hm = C.c_char_p(), C.c_uint64(), C.c_int64(), C.c_int()
        d = (C.c_uint8 * 16)()
        _check(self._L.tpimd5_manifest_entry(self._h, i, C.byref(p), C.byref(s), C.byref(m), d, C.byref(hm)))
        return os.fsdecode(p.value), int(s.value), int(m.value), (bytes(d) if hm.value else None)

    def entries(self):
        return [self.entry(i) for i in range(len(self))]

    def paths(self) -> List[str]:
        return [self.entry(i)[0] for i in range(len(self))]

    def close(self):
        if self._h:
            self._L.tpimd5_manifest_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def plan_transfer(ctx: Optional[Context], src: Manifest, dst: Manifest, mode: int = MODE_MODTIME, window_ns: int = 1):
    """-> (actions uint8[len(src)] of ACT_*, n_transfer, n_dst_only).  ctx may be None only when no pair
    needs its digests compared (there is no CPU comparison to fall back to)."""
    n = len(src)
    act = np.zeros(max(1, n), dtype=np.uint8)
    nt, nd = C.c_size_t(), C.c_size_t()
    _check(load_library().tpimd5_plan_transfer(ctx._h if ctx else None, src._h, dst._h, mode, window_ns, act.ctypes.data,
                                               C.byref(nt), C.byref(nd)))
    return act[:n], int(nt.value), int(nd.value)
