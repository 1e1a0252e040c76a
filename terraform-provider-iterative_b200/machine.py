"""`machine` -- the reference's storage-sync entry points for this path, same names and arguments.

Mirrors /root/reference/task/common/machine/storage.go: ``Transfer(ctx, source, destination,
exclude)`` (:123), ``LimitTransfer(subdir, rules)`` (:267), ``isRcloneFilter`` (:282).  ``ctx`` is a
``Context`` (the GPU) where Go has a ``context.Context``.  Errors come back as exceptions the way Go
returns ``error``.  Only local directories are moved here; cloud remotes stay with rclone.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Sequence

from .binding import Context, _check, load_library

DEFAULT_TRANSFER_EXCLUDES = ["- /main.tf", "- /terraform.tfstate*", "- /.terraform**"]   # storage.go:37-41


class _Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("files_listed", "bytes_listed", "transferred", "bytes_transferred", "skipped",
                                          "mtime_fixed", "dst_only", "bytes_hashed")]


def is_rclone_filter(rule: str) -> bool:
    return rule.startswith("+ ") or rule.startswith("- ")


def _clean_join_root(p: str) -> str:
    parts: List[str] = []
    for seg in ("/" + p).split("/"):
        if seg in ("", "."):
            continue
        if seg == "..":
            if parts:
                parts.pop()
            continue
        parts.append(seg)
    return "/" + "/".join(parts)


def LimitTransfer(subdir: str, rules: Sequence[str]) -> List[str]:
    """storage.go:267-280: keep only one subdirectory (the DirectoryOut of a Pull)."""
    d = _clean_join_root(subdir)
    if subdir in ("", ".") or d == "/":
        return list(rules)
    return list(rules) + ["+ " + d, "+ " + d + "/**", "- /**"]


def Transfer(ctx: Context, source: str, destination: str, exclude: Optional[Sequence[str]] = None, checksum: bool = False,
             size_only: bool = False) -> Dict[str, int]:
    """Sync `source` into `destination` the way machine.Transfer does (rclone CopyDir, empty
    directories included).  `checksum` is rclone's CheckSum config flag (BASELINE config 1)."""
    rules = [os.fsencode(r) for r in (exclude or [])]
    arr = (C.c_char_p * max(1, len(rules)))(*rules)
    st = _Stats()
    mode = 2 if size_only else (1 if checksum else 0)
    _check(load_library().tpimd5_transfer_local(ctx._h, os.fsencode(source), os.fsencode(destination), C.cast(arr, C.c_void_p),
                                                len(rules), None, mode, C.byref(st)))
    return {n: int(getattr(st, n)) for n, _ in _Stats._fields_}
