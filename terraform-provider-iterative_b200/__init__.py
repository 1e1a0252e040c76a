"""terraform-provider-iterative_b200 -- host-side binding of the B200 workdir-checksum path.

The product is ``libtpimd5.so`` (``csrc/``, C-ABI in ``include/tpimd5.h``): hand-written sm_100a
kernels that replace the MD5 / S3-ETag arithmetic TPI reaches through rclone from
``/root/reference/task/common/machine/storage.go:123-159`` (``machine.Transfer``).  This package is
the thin ctypes layer over that C-ABI -- the Python twin of the cgo package a TPI maintainer would
add -- plus the host logic that sits either side of the hash (``manifest``,
``sharding``).  There is no CPU fallback: importing works anywhere, but every compute call needs
the library and a B200, and fails loudly otherwise.

The directory name is not a Python identifier; import it with
``importlib.import_module("terraform-provider-iterative_b200")`` or through the ``tpi_b200`` shim
at the repo root.
"""
from .binding import (  # noqa: F401
    Context,
    PathList,
    Stream,
    TpiMd5Error,
    NOALLOC_MAX_PARTS,
    dev_s3_etag_noalloc,
    device_count,
    lib_path,
    load_library,
    md5_base64,
    md5_from_etag,
    md5_hex,
    s3_etag_text,
    s3_nparts,
    s3_part_size,
    s3_plan_parts,
    shard_lpt,
)

from . import machine, manifest, sharding  # noqa: F401,E402
from .manifest import Filter, Manifest, plan_transfer  # noqa: F401,E402

__all__ = [
    "Filter", "Manifest", "plan_transfer", "machine", "manifest", "sharding",
    "Context", "PathList", "Stream", "TpiMd5Error", "NOALLOC_MAX_PARTS", "dev_s3_etag_noalloc", "device_count", "lib_path", "load_library",
    "s3_etag_text", "s3_nparts", "s3_part_size", "s3_plan_parts", "shard_lpt",
]
