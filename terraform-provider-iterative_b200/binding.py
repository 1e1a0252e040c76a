"""ctypes binding of include/tpimd5.h -- one Python method per C entry point, no logic of its own.

Mirrors the cgo package sketched in INTEGRATION.md: the names follow the Go/rclone interfaces the
C-ABI replaces (``Stream`` has Go ``hash.Hash``'s Write/Sum/Reset/Size/BlockSize).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_NAME = "libtpimd5.so"
_lib: Optional[C.CDLL] = None

# every symbol include/tpimd5.h declares (tests check the built library exports each one)
EXPORTS = [
    "tpimd5_device_count", "tpimd5_init", "tpimd5_destroy", "tpimd5_ctx_devices", "tpimd5_get_progress", "tpimd5_last_error",
    "tpimd5_version", "tpimd5_hash_buffers", "tpimd5_hash_files", "tpimd5_s3_part_size", "tpimd5_s3_nparts",
    "tpimd5_s3_etag_buffer", "tpimd5_s3_etag_file", "tpimd5_s3_plan_parts", "tpimd5_units_extent", "tpimd5_ingest_probe", "tpimd5_s3_upload_plan", "tpimd5_stream_open", "tpimd5_stream_write",
    "tpimd5_stream_write_many", "tpimd5_stream_sum", "tpimd5_stream_reset", "tpimd5_stream_close", "tpimd5_diff",
    "tpimd5_dev_hash_buffers", "tpimd5_dev_s3_etag", "tpimd5_dev_diff", "tpimd5_dev_synth_fill",
    "tpimd5_dev_s3_etag_noalloc", "tpimd5_filter_new", "tpimd5_filter_free", "tpimd5_filter_include", "tpimd5_filter_include_dir",
    "tpimd5_filter_rule_count", "tpimd5_filter_rule", "tpimd5_manifest_scan", "tpimd5_manifest_from_arrays",
    "tpimd5_manifest_free", "tpimd5_manifest_count", "tpimd5_manifest_entry", "tpimd5_manifest_save",
    "tpimd5_manifest_load", "tpimd5_manifest_write_md5sum", "tpimd5_plan_transfer", "tpimd5_transfer_local",
    "tpimd5_hex", "tpimd5_base64", "tpimd5_md5_from_etag", "tpimd5_s3_etag_text",
    "tpimd5_shard_lpt", "tpimd5_dev_int_probe", "tpimd5_launch_count", "tpimd5_set_variant",
]


class TpiMd5Error(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"tpimd5 error {code}: {message}")
        self.code = code


def lib_path() -> str:
    return os.path.join(_HERE, _LIB_NAME)


def load_library() -> C.CDLL:
    """Loads the in-tree library.  No fallback of any kind: a missing build is an error."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise TpiMd5Error(-2, f"{path} is missing: run ./build.sh (or __graft_entry__.build()); "
                              "there is no CPU implementation to fall back to")
    L = C.CDLL(path)
    vp, u64, u32, sz, i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_size_t, C.c_int
    sig = {
        "tpimd5_device_count": (i32, []),
        "tpimd5_init": (i32, [vp, i32, sz, C.POINTER(vp)]),
        "tpimd5_destroy": (None, [vp]),
        "tpimd5_ctx_devices": (i32, [vp]),
        "tpimd5_get_progress": (i32, [vp, vp]),
        "tpimd5_last_error": (C.c_char_p, []),
        "tpimd5_version": (C.c_char_p, []),
        "tpimd5_hash_buffers": (i32, [vp, vp, vp, vp, sz, vp]),
        "tpimd5_hash_files": (i32, [vp, vp, vp, sz, vp, C.POINTER(sz)]),
        "tpimd5_s3_part_size": (u64, [u64, u64, u32]),
        "tpimd5_s3_nparts": (u32, [u64, u64]),
        "tpimd5_s3_etag_buffer": (i32, [vp, vp, u64, u64, vp, vp, C.POINTER(u32)]),
        "tpimd5_s3_etag_file": (i32, [vp, C.c_char_p, u64, u64, vp, vp, C.POINTER(u32)]),
        "tpimd5_s3_plan_parts": (u32, [u64, u64, u64, u32, C.POINTER(i32), C.POINTER(u64)]),
        "tpimd5_units_extent": (C.c_int, [vp, vp, vp, C.c_size_t, C.POINTER(u64)]),
        "tpimd5_ingest_probe": (i32, [vp, vp, vp, vp, sz, C.POINTER(sz)]),
        "tpimd5_s3_upload_plan": (i32, [vp, C.c_char_p, u64, u64, u64, u32, u32, vp, vp, vp]),
        "tpimd5_stream_open": (i32, [vp, C.POINTER(vp)]),
        "tpimd5_stream_write": (i32, [vp, vp, sz]),
        "tpimd5_stream_write_many": (i32, [vp, vp, vp, sz]),
        "tpimd5_stream_sum": (i32, [vp, vp]),
        "tpimd5_stream_reset": (i32, [vp]),
        "tpimd5_stream_close": (None, [vp]),
        "tpimd5_diff": (i32, [vp, vp, vp, sz, vp, C.POINTER(sz)]),
        "tpimd5_dev_hash_buffers": (i32, [vp, i32, vp, u64, vp, vp, u64, sz, vp, vp]),
        "tpimd5_dev_s3_etag": (i32, [vp, i32, vp, u64, u64, vp, vp, vp]),
        "tpimd5_dev_diff": (i32, [vp, i32, vp, vp, sz, vp, vp, vp]),
        "tpimd5_dev_synth_fill": (i32, [vp, i32, vp, u64, u64, u64, u64, vp]),
        "tpimd5_dev_s3_etag_noalloc": (i32, [i32, vp, u64, u64, vp, vp]),
        "tpimd5_filter_new": (i32, [vp, sz, C.c_char_p, C.POINTER(vp)]),
        "tpimd5_filter_free": (None, [vp]),
        "tpimd5_filter_include": (i32, [vp, C.c_char_p]),
        "tpimd5_filter_include_dir": (i32, [vp, C.c_char_p]),
        "tpimd5_filter_rule_count": (sz, [vp]),
        "tpimd5_filter_rule": (C.c_char_p, [vp, sz]),
        "tpimd5_manifest_scan": (i32, [vp, C.c_char_p, vp, C.POINTER(vp)]),
        "tpimd5_manifest_from_arrays": (i32, [vp, vp, vp, vp, vp, sz, C.POINTER(vp)]),
        "tpimd5_manifest_free": (None, [vp]),
        "tpimd5_manifest_count": (sz, [vp]),
        "tpimd5_manifest_entry": (i32, [vp, sz, C.POINTER(C.c_char_p), C.POINTER(u64), C.POINTER(C.c_int64), vp,
                                        C.POINTER(i32)]),
        "tpimd5_manifest_save": (i32, [vp, C.c_char_p]),
        "tpimd5_manifest_load": (i32, [C.c_char_p, C.POINTER(vp)]),
        "tpimd5_manifest_write_md5sum": (i32, [vp, C.c_char_p]),
        "tpimd5_plan_transfer": (i32, [vp, vp, vp, i32, C.c_int64, vp, C.POINTER(sz), C.POINTER(sz)]),
        "tpimd5_transfer_local": (i32, [vp, C.c_char_p, C.c_char_p, vp, sz, C.c_char_p, i32, vp]),
        "tpimd5_hex": (None, [vp, vp]),
        "tpimd5_base64": (None, [vp, vp]),
        "tpimd5_md5_from_etag": (i32, [C.c_char_p, vp]),
        "tpimd5_s3_etag_text": (i32, [vp, u32, vp, sz]),
        "tpimd5_shard_lpt": (i32, [vp, sz, i32, vp]),
        "tpimd5_dev_int_probe": (i32, [vp, i32, i32, i32, u32, vp, C.POINTER(u64), vp]),
        "tpimd5_launch_count": (u64, [vp]),
        "tpimd5_set_variant": (i32, [vp, i32]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def _check(rc: int) -> None:
    if rc != 0:
        raise TpiMd5Error(rc, (load_library().tpimd5_last_error() or b"").decode("utf-8", "replace"))


def device_count() -> int:
    return int(load_library().tpimd5_device_count())


def s3_part_size(size: int, chunk_size: int = 5 << 20, max_parts: int = 10000) -> int:
    return int(load_library().tpimd5_s3_part_size(size, chunk_size, max_parts))


def s3_nparts(size: int, part_size: int) -> int:
    return int(load_library().tpimd5_s3_nparts(size, part_size))


class _S3Plan(C.Structure):
    _fields_ = [("multipart", C.c_int32), ("has_md5", C.c_int32), ("part_size", C.c_uint64), ("nparts", C.c_uint32),
                ("md5", C.c_uint8 * 16), ("content_md5", C.c_char * 25), ("etag", C.c_char * 48)]


PLAN_NO_WHOLE_MD5 = 1


def s3_plan_parts(size: int, chunk_size: int = 0, upload_cutoff: int = 0, max_parts: int = 0):
    """rclone's single-vs-multipart decision for an object of `size` bytes -> (multipart, part_size, nparts)."""
    multi, part = C.c_int32(), C.c_uint64()
    n = load_library().tpimd5_s3_plan_parts(size, chunk_size, upload_cutoff, max_parts, C.byref(multi), C.byref(part))
    return bool(multi.value), int(part.value), int(n)


def s3_etag_text(etag_md5: bytes, nparts: int) -> str:
    """rclone's expected multipart ETag string: lowercase hex + "-" + part count."""
    d = (C.c_uint8 * 16).from_buffer_copy(bytes(etag_md5))
    out = C.create_string_buffer(64)
    _check(load_library().tpimd5_s3_etag_text(d, nparts, out, 64))
    return out.value.decode()


def md5_hex(md5: bytes) -> str:
    d = (C.c_uint8 * 16).from_buffer_copy(bytes(md5))
    out = C.create_string_buffer(33)
    load_library().tpimd5_hex(d, out)
    return out.value.decode()


def md5_base64(md5: bytes) -> str:
    """Content-MD5 / Md5chksum / GCS md5Hash form of a digest."""
    d = (C.c_uint8 * 16).from_buffer_copy(bytes(md5))
    out = C.create_string_buffer(25)
    load_library().tpimd5_base64(d, out)
    return out.value.decode()


def md5_from_etag(etag: str) -> Optional[bytes]:
    """rclone's rule: an S3 ETag is an MD5 only if, lowercased and without quotes, it is exactly 32 hex digits."""
    d = (C.c_uint8 * 16)()
    return bytes(d) if load_library().tpimd5_md5_from_etag(etag.encode(), d) == 1 else None


def shard_lpt(lengths, n_shards: int) -> np.ndarray:
    """Unit -> shard (GPU) assignment, byte balanced.  Pure host logic (runs without a GPU)."""
    ln = np.ascontiguousarray(lengths, dtype=np.uint64)
    out = np.zeros(ln.size, dtype=np.uint32)
    _check(load_library().tpimd5_shard_lpt(ln.ctypes.data, ln.size, n_shards, out.ctypes.data))
    return out


NOALLOC_MAX_PARTS = 10000


def dev_s3_etag_noalloc(device: int, d_data, size: int, part_size: int, d_etag, stream=None) -> None:
    """S3 ETag of device-resident data for callers that may not allocate device memory (a
    checkpoint hook inside a training process whose memory pool owns the GPU, a host program on
    a fixed memory budget).  Needs no Context; the part digests go to an array built into the
    library (at most NOALLOC_MAX_PARTS parts).  Calls that can overlap on one GPU race on it."""
    _check(load_library().tpimd5_dev_s3_etag_noalloc(device, _ptr(d_data), size, part_size, _ptr(d_etag), _ptr(stream)))


def _u8(buf) -> np.ndarray:
    a = buf if isinstance(buf, np.ndarray) else np.frombuffer(buf, dtype=np.uint8)
    if a.dtype != np.uint8:
        a = a.view(np.uint8)
    if not a.flags["C_CONTIGUOUS"]:
        a = np.ascontiguousarray(a)
    return a.reshape(-1)


def _ptr(x) -> Optional[int]:
    """Device pointers may be ints, torch tensors (data_ptr) or None."""
    if x is None:
        return None
    if hasattr(x, "data_ptr"):
        return int(x.data_ptr())
    return int(x)


class PathList:
    """A listing's paths as the C array the file entry points take (char *const *), built once and reused: a caller
    that hashes the same listing again (a checkpoint loop) does not pay for re-encoding 10^5..10^6 strings."""

    def __init__(self, paths: Sequence[str]):
        self.n = len(paths)
        self._keep = [os.fsencode(p) for p in paths]
        self.array = (C.c_char_p * max(1, self.n))(*self._keep)

    def __len__(self) -> int:
        return self.n


class Context:
    """One library context = a set of B200s plus their staging memory (tpimd5_init/destroy)."""

    def __init__(self, devices: Optional[Sequence[int]] = None, staging_bytes: int = 0):
        L = load_library()
        self._L = L
        self._h = C.c_void_p()
        if devices is None:
            devices = [0]
        ids = (C.c_int * len(devices))(*devices)
        _check(L.tpimd5_init(ids, len(devices), staging_bytes, C.byref(self._h)))
        self.devices = list(devices)

    def close(self) -> None:
        if self._h:
            self._L.tpimd5_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- host-memory entry points -----------------------------------------------------------
    def hash_buffers(self, base, offsets, lengths, out: Optional[np.ndarray] = None) -> np.ndarray:
        """MD5 of each (offset, length) unit of `base` -> uint8[n, 16] (into `out` if given, the way a Go
        caller reuses its digest slice)."""
        off = np.ascontiguousarray(offsets, dtype=np.uint64)
        ln = np.ascontiguousarray(lengths, dtype=np.uint64)
        if off.shape != ln.shape or off.ndim != 1:
            raise ValueError("offsets and lengths must be 1-D and the same length")
        if out is None:
            out = np.empty((off.size, 16), dtype=np.uint8)
        elif out.dtype != np.uint8 or out.shape != (off.size, 16) or not out.flags["C_CONTIGUOUS"]:
            raise ValueError("out must be a C-contiguous uint8[n, 16]")
        if isinstance(base, int):
            bp = base
        else:
            a = _u8(base)
            end = C.c_uint64()
            _check(self._L.tpimd5_units_extent(self._h, off.ctypes.data, ln.ctypes.data, off.size, C.byref(end)))
            if end.value > a.size:
                raise ValueError("a unit runs past the end of the buffer")
            bp = a.ctypes.data
        _check(self._L.tpimd5_hash_buffers(self._h, bp, off.ctypes.data, ln.ctypes.data, off.size, out.ctypes.data))
        return out

    def hash_files(self, paths, sizes=None) -> np.ndarray:
        """paths: a sequence of str, or a PathList built once for a listing that is hashed repeatedly."""
        pl = paths if isinstance(paths, PathList) else PathList(paths)
        n, arr = pl.n, pl.array
        sz = None if sizes is None else np.ascontiguousarray(sizes, dtype=np.uint64)
        out = np.zeros((n, 16), dtype=np.uint8)
        failed = C.c_size_t()
        rc = self._L.tpimd5_hash_files(self._h, C.cast(arr, C.c_void_p), None if sz is None else sz.ctypes.data, n,
                                       out.ctypes.data, C.byref(failed))
        if rc != 0:
            e = TpiMd5Error(rc, (self._L.tpimd5_last_error() or b"").decode("utf-8", "replace"))
            e.failed_index = None if failed.value == C.c_size_t(-1).value else int(failed.value)
            raise e
        return out

    def ingest_probe(self, paths, sizes, file_offsets=None) -> None:
        """Reads the listed file ranges into the staging buffers the way hash_files does and stops there (no copy
        to the device, no hashing): the host-ingest ceiling of the file entry points."""
        pl = paths if isinstance(paths, PathList) else PathList(paths)
        sz = np.ascontiguousarray(sizes, dtype=np.uint64)
        fo = None if file_offsets is None else np.ascontiguousarray(file_offsets, dtype=np.uint64)
        if sz.size != pl.n or (fo is not None and fo.size != pl.n):
            raise ValueError("paths, sizes and file_offsets must have the same length")
        failed = C.c_size_t()
        _check(self._L.tpimd5_ingest_probe(self._h, C.cast(pl.array, C.c_void_p), None if fo is None else fo.ctypes.data,
                                           sz.ctypes.data, pl.n, C.byref(failed)))

    def s3_etag_buffer(self, data, part_size: int):
        """-> (etag_md5 bytes, nparts, part_digests uint8[nparts, 16])."""
        a = _u8(data)
        n = s3_nparts(a.size, part_size)
        parts = np.zeros((max(n, 1), 16), dtype=np.uint8)
        etag = (C.c_uint8 * 16)()
        npo = C.c_uint32()
        _check(self._L.tpimd5_s3_etag_buffer(self._h, a.ctypes.data if a.size else None, a.size, part_size,
                                             parts.ctypes.data, etag, C.byref(npo)))
        return bytes(etag), int(npo.value), parts[: npo.value]

    def s3_etag_file(self, path: str, size: int, part_size: int):
        n = s3_nparts(size, part_size)
        parts = np.zeros((max(n, 1), 16), dtype=np.uint8)
        etag = (C.c_uint8 * 16)()
        npo = C.c_uint32()
        _check(self._L.tpimd5_s3_etag_file(self._h, os.fsencode(path), size, part_size, parts.ctypes.data, etag,
                                           C.byref(npo)))
        return bytes(etag), int(npo.value), parts[: npo.value]

    def s3_upload_plan(self, path: str, size: int, chunk_size: int = 0, upload_cutoff: int = 0, max_parts: int = 0,
                       no_whole_md5: bool = False) -> dict:
        """What rclone's S3 back-end computes before uploading `path`: mode, part size, per-part Content-MD5,
        whole-file Content-MD5 / Md5chksum and the ETag to expect back (tpimd5_s3_upload_plan)."""
        _, _, n = s3_plan_parts(size, chunk_size, upload_cutoff, max_parts)
        parts = np.zeros((max(n, 1), 16), dtype=np.uint8)
        b64 = C.create_string_buffer(25 * max(n, 1))
        plan = _S3Plan()
        _check(self._L.tpimd5_s3_upload_plan(self._h, os.fsencode(path), size, chunk_size, upload_cutoff, max_parts,
                                             PLAN_NO_WHOLE_MD5 if no_whole_md5 else 0, C.byref(plan), parts.ctypes.data, b64))
        k = int(plan.nparts)
        return {"multipart": bool(plan.multipart), "part_size": int(plan.part_size), "nparts": k,
                "md5": bytes(plan.md5) if plan.has_md5 else None,
                "content_md5": plan.content_md5.decode() if plan.has_md5 else None, "etag": plan.etag.decode(),
                "part_md5": parts[:k].copy(), "part_content_md5": [b64.raw[25 * i: 25 * i + 24].decode() for i in range(k)]}

    def diff(self, old16, new16) -> np.ndarray:
        o, n = _u8(old16), _u8(new16)
        if o.size != n.size or o.size % 16:
            raise ValueError("manifests must be the same length, 16 bytes per unit")
        rows = o.size // 16
        idx = np.zeros(max(rows, 1), dtype=np.uint32)
        cnt = C.c_size_t()
        _check(self._L.tpimd5_diff(self._h, o.ctypes.data, n.ctypes.data, rows, idx.ctypes.data, C.byref(cnt)))
        return idx[: cnt.value].copy()

    def stream(self) -> "Stream":
        return Stream(self)

    def stream_write_many(self, streams: Sequence["Stream"], chunks: Sequence) -> None:
        n = len(streams)
        keep = [_u8(c) for c in chunks]
        hs = (C.c_void_p * n)(*[s._h for s in streams])
        ps = (C.c_void_p * n)(*[(k.ctypes.data if k.size else None) for k in keep])
        ls = (C.c_size_t * n)(*[k.size for k in keep])
        _check(self._L.tpimd5_stream_write_many(C.cast(hs, C.c_void_p), C.cast(ps, C.c_void_p), C.cast(ls, C.c_void_p), n))

    # ---- device-memory entry points (pointers: ints or torch tensors) ------------------------
    def dev_hash_buffers(self, d_base, base_bytes: int, n: int, d_digests, d_offsets=None, d_lengths=None,
                         stride: int = 0, stream=None, slot: int = 0) -> None:
        _check(self._L.tpimd5_dev_hash_buffers(self._h, slot, _ptr(d_base), base_bytes, _ptr(d_offsets), _ptr(d_lengths),
                                               stride, n, _ptr(d_digests), _ptr(stream)))

    def dev_s3_etag(self, d_data, size: int, part_size: int, d_etag, d_part_digests=None, stream=None, slot: int = 0) -> None:
        _check(self._L.tpimd5_dev_s3_etag(self._h, slot, _ptr(d_data), size, part_size, _ptr(d_part_digests), _ptr(d_etag),
                                          _ptr(stream)))

    def dev_diff(self, d_old, d_new, n: int, d_idx, d_count, stream=None, slot: int = 0) -> None:
        _check(self._L.tpimd5_dev_diff(self._h, slot, _ptr(d_old), _ptr(d_new), n, _ptr(d_idx), _ptr(d_count), _ptr(stream)))

    def dev_synth_fill(self, d_out, seed: int, first_file_index: int, file_len: int, n_files: int, stream=None,
                       slot: int = 0) -> None:
        _check(self._L.tpimd5_dev_synth_fill(self._h, slot, _ptr(d_out), seed, first_file_index, file_len, n_files,
                                             _ptr(stream)))

    def dev_int_probe(self, mode: int, ctas_per_sm: int, iters: int, d_out, stream=None, slot: int = 0) -> int:
        blocks = C.c_uint64()
        _check(self._L.tpimd5_dev_int_probe(self._h, slot, mode, ctas_per_sm, iters, _ptr(d_out), C.byref(blocks),
                                            _ptr(stream)))
        return int(blocks.value)

    def progress(self) -> dict:
        """Counters of the context's host-pointer calls so far (what a 10 s stats reporter polls)."""
        buf = (C.c_uint64 * 5)()
        _check(self._L.tpimd5_get_progress(self._h, buf))
        return dict(zip(("calls_started", "calls_finished", "units_done", "bytes_staged", "bytes_done"), (int(x) for x in buf)))

    def launch_count(self) -> int:
        return int(self._L.tpimd5_launch_count(self._h))

    def set_variant(self, variant: int) -> None:
        _check(self._L.tpimd5_set_variant(self._h, variant))


class Stream:
    """Go hash.Hash over the GPU path (crypto/md5's New()): Write / Sum / Reset / Size / BlockSize."""

    Size = 16
    BlockSize = 64

    def __init__(self, ctx: Context):
        self._ctx = ctx
        self._L = ctx._L
        self._h = C.c_void_p()
        _check(self._L.tpimd5_stream_open(ctx._h, C.byref(self._h)))

    def write(self, data) -> int:
        a = _u8(data)
        if a.size:
            _check(self._L.tpimd5_stream_write(self._h, a.ctypes.data, a.size))
        return int(a.size)

    def sum(self) -> bytes:
        out = (C.c_uint8 * 16)()
        _check(self._L.tpimd5_stream_sum(self._h, out))
        return bytes(out)

    def reset(self) -> None:
        _check(self._L.tpimd5_stream_reset(self._h))

    def close(self) -> None:
        if self._h:
            self._L.tpimd5_stream_close(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
