//go:build cuda && linux && amd64

// Package cuda binds libtpimd5.so (include/tpimd5.h): the B200 implementation of the MD5 / S3-ETag
// arithmetic that terraform-provider-iterative reaches through rclone under machine.Transfer
// (task/common/machine/storage.go:123-159).  One Go function per C function, no logic of its own:
// everything that can be wrong is below the C-ABI, where it is tested.
//
// SOURCE ONLY: the image this was written on has no Go toolchain and no rclone module cache, so this
// file has never been compiled.  Build: CGO_ENABLED=1 go build -tags cuda ./...
package cuda

/*
#cgo CFLAGS: -I${SRCDIR}/../../../include
#cgo LDFLAGS: -L${SRCDIR}/../../../terraform-provider-iterative_b200 -ltpimd5
#include <stdlib.h>
#include "tpimd5.h"
*/
import "C"

import (
	"errors"
	"fmt"
	"runtime"
	"unsafe"
)

// Context is one tpimd5_ctx: a set of GPUs plus their staging memory.  Safe for concurrent use;
// calls that share a GPU overlap inside the library.
type Context struct{ h *C.tpimd5_ctx }

// lastErr must run on the OS thread that made the failing call (tpimd5_last_error is per thread).
func lastErr(rc C.int) error {
	if rc == 0 {
		return nil
	}
	return fmt.Errorf("tpimd5 (%d): %s", int(rc), C.GoString(C.tpimd5_last_error()))
}

// NewContext opens the given devices (nil = device 0).  staging 0 = 256 MiB per staging buffer.
func NewContext(devices []int, staging uint64) (*Context, error) {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	if len(devices) == 0 {
		devices = []int{0}
	}
	ids := make([]C.int, len(devices))
	for i, d := range devices {
		ids[i] = C.int(d)
	}
	var h *C.tpimd5_ctx
	if err := lastErr(C.tpimd5_init(&ids[0], C.int(len(ids)), C.size_t(staging), &h)); err != nil {
		return nil, err
	}
	c := &Context{h}
	runtime.SetFinalizer(c, func(c *Context) { c.Close() })
	return c, nil
}

func (c *Context) Close() {
	if c.h != nil {
		C.tpimd5_destroy(c.h)
		c.h = nil
	}
}

// HashFiles replaces one rclone Object.Hash(ctx, hash.MD5) per file with one call per listing.
// sizes come from the listing (fs.Object.Size()); a file whose size differs from its listing, in
// either direction, is an error (failed = its index), never a digest.
func (c *Context) HashFiles(paths []string, sizes []uint64) (digests [][16]byte, failed int, err error) {
	n := len(paths)
	digests = make([][16]byte, n)
	failed = -1
	if n == 0 {
		return digests, failed, nil
	}
	if len(sizes) != n {
		return nil, failed, errors.New("tpimd5: HashFiles: len(sizes) != len(paths)")
	}
	cpaths := make([]*C.char, n) // C strings: the C side must not see Go string memory
	for i, p := range paths {
		cpaths[i] = C.CString(p)
	}
	defer func() {
		for _, p := range cpaths {
			C.free(unsafe.Pointer(p))
		}
	}()
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	var bad C.size_t
	rc := C.tpimd5_hash_files(c.h, (**C.char)(unsafe.Pointer(&cpaths[0])),
		(*C.uint64_t)(unsafe.Pointer(&sizes[0])), C.size_t(n),
		(*C.uint8_t)(unsafe.Pointer(&digests[0])), &bad)
	if rc != 0 {
		if bad != ^C.size_t(0) {
			failed = int(bad)
		}
		return nil, failed, lastErr(rc)
	}
	return digests, failed, nil
}

// S3ETagFile replaces uploadMultipart's per-part md5.Sum and the final md5.Sum(md5s).
func (c *Context) S3ETagFile(path string, size, partSize uint64) (etag [16]byte, parts [][16]byte, err error) {
	np := uint32(C.tpimd5_s3_nparts(C.uint64_t(size), C.uint64_t(partSize)))
	if np == 0 {
		return etag, nil, errors.New("tpimd5: S3ETagFile: part size 0 or more than 2^32 parts")
	}
	parts = make([][16]byte, np)
	cp := C.CString(path)
	defer C.free(unsafe.Pointer(cp))
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	var n C.uint32_t
	rc := C.tpimd5_s3_etag_file(c.h, cp, C.uint64_t(size), C.uint64_t(partSize),
		(*C.uint8_t)(unsafe.Pointer(&parts[0])), (*C.uint8_t)(unsafe.Pointer(&etag[0])), &n)
	return etag, parts, lastErr(rc)
}

// S3Plan is what backend/s3 Object.Update needs before it uploads one local file.
type S3Plan struct {
	Multipart      bool
	PartSize       uint64
	Parts          int
	MD5            *[16]byte // nil when the whole-file MD5 was switched off (multipart only)
	ContentMD5     string    // base64: Content-MD5 (single PutObject) / X-Amz-Meta-Md5chksum (multipart)
	ETag           string    // the ETag to expect back
	PartMD5        [][16]byte
	PartContentMD5 []string
}

// S3UploadPlan: zero chunkSize / uploadCutoff / maxParts select rclone's defaults (5 MiB, 200 MiB, 10,000).
// noWholeMD5 is rclone's --s3-disable-checksum.
func (c *Context) S3UploadPlan(path string, size, chunkSize, uploadCutoff uint64, maxParts uint32, noWholeMD5 bool) (*S3Plan, error) {
	np := int(C.tpimd5_s3_plan_parts(C.uint64_t(size), C.uint64_t(chunkSize), C.uint64_t(uploadCutoff), C.uint32_t(maxParts), nil, nil))
	if np == 0 {
		return nil, errors.New("tpimd5: S3UploadPlan: too many parts")
	}
	partMD5 := make([][16]byte, np)
	b64 := make([]byte, 25*np)
	cp := C.CString(path)
	defer C.free(unsafe.Pointer(cp))
	var flags C.uint32_t
	if noWholeMD5 {
		flags = C.TPIMD5_PLAN_NO_WHOLE_MD5
	}
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	var plan C.tpimd5_s3_plan
	rc := C.tpimd5_s3_upload_plan(c.h, cp, C.uint64_t(size), C.uint64_t(chunkSize), C.uint64_t(uploadCutoff), C.uint32_t(maxParts),
		flags, &plan, (*C.uint8_t)(unsafe.Pointer(&partMD5[0])), (*C.char)(unsafe.Pointer(&b64[0])))
	if err := lastErr(rc); err != nil {
		return nil, err
	}
	out := &S3Plan{Multipart: plan.multipart != 0, PartSize: uint64(plan.part_size), Parts: int(plan.nparts),
		ETag: C.GoString(&plan.etag[0]), PartMD5: partMD5, PartContentMD5: make([]string, np)}
	if plan.has_md5 != 0 {
		var d [16]byte
		for i := range d {
			d[i] = byte(plan.md5[i])
		}
		out.MD5 = &d
		out.ContentMD5 = C.GoString(&plan.content_md5[0])
	}
	for i := 0; i < np; i++ {
		out.PartContentMD5[i] = string(b64[25*i : 25*i+24])
	}
	return out, nil
}

// Diff replaces the per-pair srcHash != dstHash of operations.checkHashes for two manifests:
// ascending indices of the rows that differ.
func (c *Context) Diff(old, new [][16]byte) ([]uint32, error) {
	if len(old) != len(new) {
		return nil, errors.New("tpimd5: Diff: manifests differ in length")
	}
	if len(old) == 0 {
		return nil, nil
	}
	idx := make([]uint32, len(old))
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	var k C.size_t
	rc := C.tpimd5_diff(c.h, (*C.uint8_t)(unsafe.Pointer(&old[0])), (*C.uint8_t)(unsafe.Pointer(&new[0])),
		C.size_t(len(old)), (*C.uint32_t)(unsafe.Pointer(&idx[0])), &k)
	return idx[:k], lastErr(rc)
}

// MD5FromETag is rclone's test whether an S3 ETag is an MD5 at all (lowercased, quotes trimmed, 32 hex).
func MD5FromETag(etag string) (md5 [16]byte, ok bool) {
	ce := C.CString(etag)
	defer C.free(unsafe.Pointer(ce))
	ok = C.tpimd5_md5_from_etag(ce, (*C.uint8_t)(unsafe.Pointer(&md5[0]))) == 1
	return md5, ok
}

// Hash is a hash.Hash (what crypto/md5's New() returns) over the GPU path.  One Hash is one MD5 chain,
// i.e. one GPU lane: it exists for interface completeness (io.Copy into a hasher); whole listings should
// go through HashFiles.
type Hash struct{ s *C.tpimd5_stream }

func (c *Context) NewHash() (*Hash, error) {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	var s *C.tpimd5_stream
	if err := lastErr(C.tpimd5_stream_open(c.h, &s)); err != nil {
		return nil, err
	}
	h := &Hash{s}
	runtime.SetFinalizer(h, func(h *Hash) { h.Close() })
	return h, nil
}

// Write never fails for lack of data (hash.Hash contract); a CUDA failure panics, as there is no error
// path in the interface and no CPU fallback by design.
func (h *Hash) Write(p []byte) (int, error) {
	if len(p) == 0 {
		return 0, nil
	}
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	// the C side copies or hashes p before returning, so passing Go memory is within the cgo rules
	if err := lastErr(C.tpimd5_stream_write(h.s, unsafe.Pointer(&p[0]), C.size_t(len(p)))); err != nil {
		return 0, err
	}
	return len(p), nil
}

// Sum appends the digest of everything written so far; the stream continues (Go semantics).
func (h *Hash) Sum(b []byte) []byte {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	var d [16]byte
	if err := lastErr(C.tpimd5_stream_sum(h.s, (*C.uint8_t)(unsafe.Pointer(&d[0])))); err != nil {
		panic(err)
	}
	return append(b, d[:]...)
}
func (h *Hash) Reset()         { C.tpimd5_stream_reset(h.s) }
func (h *Hash) Size() int      { return 16 }
func (h *Hash) BlockSize() int { return 64 }
func (h *Hash) Close() {
	if h.s != nil {
		C.tpimd5_stream_close(h.s)
		h.s = nil
	}
}
