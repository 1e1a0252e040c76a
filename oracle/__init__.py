"""CPU oracle for the TPI workdir-checksum path -- TEST INFRASTRUCTURE, never shipped.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package, and only as the checker or the timed CPU
baseline.  The product (``terraform-provider-iterative_b200``) never imports it.

It wraps ``libtpi_oracle.so`` (built from ``tpi_oracle.c`` by ``make -C oracle``), a plain-C
restatement of Go ``crypto/md5`` (RFC 1321), rclone v1.57.0's S3 multipart ETag / part-size
rule and ``equal()`` decision table (reached from
``/root/reference/task/common/machine/storage.go:158``), plus pure-Python restatements of the
small string-level rules (rclone filter globs, md5sum manifest lines).

The rclone rules restated here (``equal``, the S3 part-size growth rule and upload plan, ``md5_from_etag``,
filter globs and ``glob_to_dir_globs``) come from knowledge of rclone v1.57.0; its source is not on this machine,
so they are UNVERIFIED against that release: agreement between this package and ``csrc/`` shows the two
restatements are consistent, nothing more.

Parity status: digest parity is *unpinned by the reference's own tests* (they hold no digest);
the oracle is pinned on RFC 1321 vectors, hashlib, OpenSSL and coreutils md5sum instead.  The
file-set membership rules ARE pinned on the reference's ``storage_test.go:55-101`` cases.
"""
from __future__ import annotations

import ctypes as C
import os
import re
import subprocess
from typing import Iterable, List, Optional, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libtpi_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile the C restatement (gcc only; a second or two)."""
    src = os.path.join(_HERE, "tpi_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        u8p, u64p, u32p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)
        L.orc_md5.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        L.orc_md5.restype = None
        L.orc_md5_ctx_size.restype = C.c_size_t
        L.orc_md5_init.argtypes = [C.c_void_p]
        L.orc_md5_update.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_md5_sum.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_md5_state.argtypes = [C.c_void_p, C.c_void_p, u64p, u32p]
        L.orc_md5_batch.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int]
        L.orc_md5_batch.restype = C.c_int
        L.orc_s3_part_size.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32]
        L.orc_s3_part_size.restype = C.c_uint64
        L.orc_s3_nparts.argtypes = [C.c_uint64, C.c_uint64]
        L.orc_s3_nparts.restype = C.c_uint32
        L.orc_s3_etag.argtypes = [C.c_int, C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, u32p, C.c_int]
        L.orc_s3_etag.restype = C.c_int
        L.orc_diff.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.orc_diff.restype = C.c_size_t
        L.orc_equal.argtypes = [C.c_uint64, C.c_uint64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_int,
                                C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_int)]
        L.orc_equal.restype = C.c_int
        L.orc_synth_fill.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint64]
        L.orc_synth_fill.restype = None
        L.orc_synth_fill_files.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_int]
        L.orc_synth_fill_files.restype = None
        L.orc_has_openssl.restype = C.c_int
        L.orc_md5_files.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int]
        L.orc_md5_files.restype = C.c_int
        _lib = L
    return _lib


ENGINE_RESTATED = 0   # the audited loop-form restatement (the checker)
ENGINE_OPENSSL = 1    # OpenSSL's assembly MD5 (timing stand-in for Go's amd64 md5block)


def _as_u8(buf) -> np.ndarray:
    a = np.frombuffer(buf, dtype=np.uint8) if not isinstance(buf, np.ndarray) else buf
    if a.dtype != np.uint8:
        a = a.view(np.uint8)
    return np.ascontiguousarray(a)


def md5(data) -> bytes:
    """md5.Sum(data) -> 16 raw bytes."""
    a = _as_u8(data)
    out = (C.c_uint8 * 16)()
    lib().orc_md5(a.ctypes.data if a.size else None, a.size, out)
    return bytes(out)


def md5_hex(data) -> str:
    return md5(data).hex()


class MD5Stream:
    """Go hash.Hash semantics: Write any chunking, Sum does not disturb the stream, Reset."""

    def __init__(self):
        self._ctx = C.create_string_buffer(lib().orc_md5_ctx_size())
        lib().orc_md5_init(self._ctx)

    def write(self, data) -> int:
        a = _as_u8(data)
        if a.size:
            lib().orc_md5_update(self._ctx, a.ctypes.data, a.size)
        return a.size

    def sum(self) -> bytes:
        out = (C.c_uint8 * 16)()
        lib().orc_md5_sum(self._ctx, out)
        return bytes(out)

    def reset(self) -> None:
        lib().orc_md5_init(self._ctx)

    def state(self) -> Tuple[Tuple[int, int, int, int], int, int]:
        st = (C.c_uint32 * 4)()
        ln, nb = C.c_uint64(), C.c_uint32()
        lib().orc_md5_state(self._ctx, st, C.byref(ln), C.byref(nb))
        return tuple(st), ln.value, nb.value


def md5_batch(base, offsets, lengths, threads: int = 1, engine: int = ENGINE_RESTATED) -> np.ndarray:
    """One digest per (offset, length) unit of a packed buffer -> uint8[n,16]."""
    a = _as_u8(base)
    off = np.ascontiguousarray(offsets, dtype=np.uint64)
    ln = np.ascontiguousarray(lengths, dtype=np.uint64)
    assert off.shape == ln.shape and off.ndim == 1
    if off.size:
        assert int((off + ln).max()) <= a.size, "unit runs past the end of the buffer"
    out = np.zeros((off.size, 16), dtype=np.uint8)
    if engine == ENGINE_OPENSSL and not lib().orc_has_openssl():
        raise RuntimeError("oracle built without OpenSSL")
    lib().orc_md5_batch(engine, a.ctypes.data, off.ctypes.data, ln.ctypes.data, off.size, out.ctypes.data, threads)
    return out


def md5_files(paths, threads: int = 1, engine: int = ENGINE_RESTATED) -> np.ndarray:
    """rclone-shaped file hashing: one file per worker, 32 KiB reads -> uint8[n,16]."""
    if hasattr(paths, "array") and hasattr(paths, "n"):      # a pre-encoded listing (same shape as the binding's PathList)
        n, arr = paths.n, paths.array
    else:
        n = len(paths)
        arr = (C.c_char_p * max(1, n))(*[os.fsencode(p) for p in paths])
    out = np.zeros((n, 16), dtype=np.uint8)
    if lib().orc_md5_files(engine, C.cast(arr, C.c_void_p), n, out.ctypes.data, threads) != 0:
        raise OSError("orc_md5_files: a file could not be read")
    return out


def s3_part_size(size: int, chunk_size: int = 5 << 20, max_parts: int = 10000) -> int:
    return int(lib().orc_s3_part_size(size, chunk_size, max_parts))


def s3_nparts(size: int, part_size: int) -> int:
    return int(lib().orc_s3_nparts(size, part_size))


def s3_etag(data, part_size: int, threads: int = 1, engine: int = ENGINE_RESTATED):
    """-> (etag_md5 16 bytes, nparts, part_digests uint8[nparts,16]).  Text ETag: s3_etag_text."""
    a = _as_u8(data)
    n = s3_nparts(a.size, part_size)
    parts = np.zeros((n, 16), dtype=np.uint8)
    etag = (C.c_uint8 * 16)()
    npo = C.c_uint32()
    rc = lib().orc_s3_etag(engine, a.ctypes.data if a.size else None, a.size, part_size, parts.ctypes.data, etag,
                           C.byref(npo), threads)
    if rc != 0:
        raise ValueError(f"orc_s3_etag failed ({rc})")
    return bytes(etag), int(npo.value), parts


def s3_etag_text(etag_md5: bytes, nparts: int) -> str:
    return f"{etag_md5.hex()}-{nparts}"


def md5_from_etag(etag: str) -> Optional[bytes]:
    """rclone backend/s3: the ETag, lowercased and with its quotes trimmed, is the object's MD5 iff it matches
    ^[0-9a-f]{32}$ (so never for multipart "-N" or SSE-KMS ETags).  Restated from memory of v1.57.0."""
    e = etag.lower().strip('"')
    return bytes.fromhex(e) if re.fullmatch(r"[0-9a-f]{32}", e) else None


def s3_upload_plan(data, chunk_size: int = 5 << 20, upload_cutoff: int = 200 << 20, max_parts: int = 10000,
                   whole_md5: bool = True) -> dict:
    """What rclone v1.57.0 backend/s3 Object.Update computes for one object before uploading it (SURVEY.md
    appendix C; restated from memory, the module is not on this machine): a single PutObject below
    upload_cutoff (Content-MD5 = base64 MD5, ETag = hex MD5), else multipart: part size by the growth rule,
    a base64 Content-MD5 per part, Md5chksum metadata = base64 whole-file MD5, ETag hex(md5(md5s))-N."""
    import base64
    a = _as_u8(data)
    multi = a.size >= upload_cutoff
    whole = md5(a) if (whole_md5 or not multi) else None
    if not multi:
        return {"multipart": False, "part_size": a.size, "nparts": 1, "md5": whole,
                "content_md5": base64.b64encode(whole).decode(), "etag": whole.hex(),
                "part_md5": np.frombuffer(whole, dtype=np.uint8).reshape(1, 16).copy(),
                "part_content_md5": [base64.b64encode(whole).decode()]}
    part = s3_part_size(a.size, chunk_size, max_parts)
    etag, n, parts = s3_etag(a, part)
    return {"multipart": True, "part_size": part, "nparts": n, "md5": whole,
            "content_md5": base64.b64encode(whole).decode() if whole else None, "etag": s3_etag_text(etag, n),
            "part_md5": parts, "part_content_md5": [base64.b64encode(bytes(p)).decode() for p in parts]}


def diff(old16, new16) -> np.ndarray:
    o, n = _as_u8(old16).reshape(-1, 16), _as_u8(new16).reshape(-1, 16)
    assert o.shape == n.shape
    idx = np.zeros(o.shape[0], dtype=np.uint32)
    k = lib().orc_diff(o.ctypes.data, n.ctypes.data, o.shape[0], idx.ctypes.data)
    return idx[:k].copy()


MODE_MODTIME, MODE_CHECKSUM, MODE_SIZE_ONLY = 0, 1, 2


def equal(src_size, dst_size, src_mtime_ns, dst_mtime_ns, src_md5: Optional[bytes], dst_md5: Optional[bytes],
          mode: int = MODE_MODTIME, window_ns: int = 1) -> Tuple[bool, bool]:
    """rclone equal(): -> (equal, fix_mtime)."""
    fix = C.c_int()
    s = (C.c_uint8 * 16).from_buffer_copy(src_md5) if src_md5 else None
    d = (C.c_uint8 * 16).from_buffer_copy(dst_md5) if dst_md5 else None
    r = lib().orc_equal(src_size, dst_size, src_mtime_ns, dst_mtime_ns, window_ns, s, 1 if src_md5 else 0, d,
                        1 if dst_md5 else 0, mode, C.byref(fix))
    return bool(r), bool(fix.value)


def synth_bytes(seed: int, file_index: int, n: int, byte_offset: int = 0) -> np.ndarray:
    """Counter-PRNG workdir content (same bytes the GPU generator kernel must produce)."""
    out = np.empty(n, dtype=np.uint8)
    if n:
        lib().orc_synth_fill(seed, file_index, byte_offset, out.ctypes.data, n)
    return out


def synth_files(seed: int, first_index: int, file_len: int, n_files: int, out: Optional[np.ndarray] = None,
                threads: int = 1) -> np.ndarray:
    """n_files equal files back to back (file i at i*file_len): the packed synthetic workdir, filled by `threads` workers."""
    if out is None:
        out = np.empty(file_len * n_files, dtype=np.uint8)
    assert out.dtype == np.uint8 and out.size >= file_len * n_files and out.flags["C_CONTIGUOUS"]
    if file_len * n_files:
        lib().orc_synth_fill_files(seed, first_index, file_len, n_files, out.ctypes.data, threads)
    return out


def synth_packed(seed: int, lengths: Sequence[int], align: int = 1, first_index: int = 0):
    """A packed buffer of files `first_index..`: -> (buffer uint8[], offsets u64[], lengths u64[])."""
    ln = np.asarray(lengths, dtype=np.uint64)
    off = np.zeros(ln.size, dtype=np.uint64)
    pos = 0
    for i, l in enumerate(ln.tolist()):
        pos = (pos + align - 1) // align * align
        off[i] = pos
        pos += l
    buf = np.zeros(max(pos, 1), dtype=np.uint8)
    for i, (o, l) in enumerate(zip(off.tolist(), ln.tolist())):
        if l:
            lib().orc_synth_fill(seed, first_index + i, 0, buf[o:].ctypes.data, l)
    return buf, off, ln


# --------------------------------------------------------------------------------------------
# md5sum text manifest (/root/reference/task/common/machine/machine-script.sh.tpl:111)
# --------------------------------------------------------------------------------------------
def md5sum_line(digest: bytes, path: str) -> str:
    """coreutils md5sum text-mode line; names holding backslash/newline get the escaped form."""
    if "\\" in path or "\n" in path or "\r" in path:
        esc = path.replace("\\", "\\\\").replace("\n", "\\n").replace("\r", "\\r")
        return "\\" + digest.hex() + "  " + esc + "\n"
    return digest.hex() + "  " + path + "\n"


# --------------------------------------------------------------------------------------------
# rclone v1.57.0 fs/filter glob -> regexp (restated), and TPI's rule assembly
# (/root/reference/task/common/machine/storage.go:37-41,126-138,267-284).
# --------------------------------------------------------------------------------------------
DEFAULT_TRANSFER_EXCLUDES = ["- /main.tf", "- /terraform.tfstate*", "- /.terraform**"]


def is_rclone_filter(rule: str) -> bool:
    return rule.startswith("+ ") or rule.startswith("- ")


def _clean_join_root(p: str) -> str:
    """Go filepath.Join("/", p) == Clean("/" + p)."""
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


def transfer_rules(exclude: Optional[Iterable[str]]) -> List[str]:
    rules = list(DEFAULT_TRANSFER_EXCLUDES) + list(exclude or [])
    out = []
    for r in rules:
        if not is_rclone_filter(r):
            r = "- " + _clean_join_root(r)
        out.append(r)
    return out


def limit_transfer(subdir: str, rules: List[str]) -> List[str]:
    d = _clean_join_root(subdir)[1:] if subdir not in ("", ".") else "."
    if d in ("", "."):
        return list(rules)
    return list(rules) + ["+ " + _clean_join_root(d), "+ " + _clean_join_root(d) + "/**", "- /**"]


def glob_to_regex(glob: str) -> str:
    """rclone filter glob: leading '/' anchors at the root, otherwise matches after any '/';
    '*' = run of non-'/', '**' = anything, '?' = one non-'/', [...] class, {a,b} alternatives."""
    out = []
    if glob.startswith("/"):
        out.append("^")
        glob = glob[1:]
    else:
        out.append("(^|/)")
    i, n, depth = 0, len(glob), 0
    while i < n:
        c = glob[i]
        if c == "\\" and i + 1 < n:
            out.append(re.escape(glob[i + 1])); i += 2; continue
        if c == "*":
            j = i
            while j < n and glob[j] == "*":
                j += 1
            out.append(".*" if j - i >= 2 else "[^/]*")
            i = j; continue
        if c == "?":
            out.append("[^/]")
        elif c == "[":
            j = glob.find("]", i + 1)
            if j < 0:
                raise ValueError("mismatched '[' in glob")
            body = glob[i + 1:j]
            if body.startswith("!"):
                body = "^" + body[1:]
            out.append("[" + body + "]"); i = j + 1; continue
        elif c == "{":
            out.append("("); depth += 1
        elif c == "}":
            if depth == 0:
                raise ValueError("mismatched '}' in glob")
            out.append(")"); depth -= 1
        elif c == "," and depth > 0:
            out.append("|")
        else:
            out.append(re.escape(c))
        i += 1
    if depth:
        raise ValueError("mismatched '{' in glob")
    out.append("$")
    return "".join(out)


def glob_to_dir_globs(glob: str) -> List[str]:
    """rclone fs/filter globToDirGlobs: directories a file glob could match in (as globs for "dir/")."""
    out: List[str] = []
    while True:
        i, j = glob.rfind("/"), glob.rfind("**")
        what = ""
        if j > i:
            i, what = j, "**"
        if i < 0:
            if not out:
                out.append("/**")
            break
        glob = glob[:i]
        g = glob + what + "/"
        if not out or out[-1] != g:
            out.append(g)
    return out


class FilterSet:
    """rclone filter semantics (fs/filter Filter.Add / Include / IncludeDirectory): the first matching
    file rule decides for a file, no match => included; directories are decided by the derived
    directory rules, matched against "dir/"."""

    def __init__(self, rules: Iterable[str]):
        self.rules = []
        self.dir_rules = []
        for r in rules:
            if not is_rclone_filter(r):
                raise ValueError(f"not an rclone filter rule: {r!r}")
            include = r[0] == "+"
            g = r[2:].strip()
            is_dir = g.endswith("/")
            is_file = not is_dir
            if is_dir and not include:          # excluding "dir/" means "dir/**"
                g += "**"
            if "**" in g:
                is_dir = is_file = True
            rx = re.compile(glob_to_regex(g))
            if is_file:
                self.rules.append((include, rx))
                if include or g == "*":
                    for dg in glob_to_dir_globs(g):
                        self.dir_rules.append((include, re.compile(glob_to_regex(dg))))
            if is_dir:
                self.dir_rules.append((include, rx))

    def include_file(self, rel_path: str) -> bool:
        rel_path = rel_path.lstrip("/")
        for inc, rx in self.rules:
            if rx.search(rel_path):
                return inc
        return True

    def include_dir(self, rel_dir: str) -> bool:
        d = rel_dir.strip("/")
        if not d:
            return True
        d += "/"
        for inc, rx in self.dir_rules:
            if rx.search(d):
                return inc
        return True

    def include_path(self, rel_path: str) -> bool:
        """A file as a filtered WALK sees it: every directory on the way must be entered."""
        parts = rel_path.strip("/").split("/")
        for k in range(1, len(parts)):
            if not self.include_dir("/".join(parts[:k])):
                return False
        return self.include_file(rel_path)


def list_transfer(files: Iterable[str], exclude: Optional[Iterable[str]] = None) -> List[str]:
    """The relative paths machine.Transfer would move, given the source file listing."""
    fs = FilterSet(transfer_rules(exclude))
    return sorted(f for f in files if fs.include_path(f))
