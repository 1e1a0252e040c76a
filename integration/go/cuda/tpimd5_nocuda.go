//go:build !(cuda && linux && amd64)

// Package cuda, default build (CGO_ENABLED=0, Makefile:22,29 of the reference): the GPU path is not
// compiled in.  There is no CPU implementation behind these names on purpose -- selecting the GPU path
// in a build that lacks it fails loudly at Transfer time instead of silently hashing on the CPU.
package cuda

import "errors"

var ErrNotBuilt = errors.New("tpimd5: built without -tags cuda (the GPU checksum path needs CGO_ENABLED=1 on linux/amd64)")

type Context struct{}
type Hash struct{}
type S3Plan struct{}

func NewContext(devices []int, staging uint64) (*Context, error) { return nil, ErrNotBuilt }
func (c *Context) Close()                                         {}
func (c *Context) HashFiles(paths []string, sizes []uint64) ([][16]byte, int, error) {
	return nil, -1, ErrNotBuilt
}
func (c *Context) S3ETagFile(path string, size, partSize uint64) ([16]byte, [][16]byte, error) {
	return [16]byte{}, nil, ErrNotBuilt
}
func (c *Context) S3UploadPlan(path string, size, chunkSize, uploadCutoff uint64, maxParts uint32, noWholeMD5 bool) (*S3Plan, error) {
	return nil, ErrNotBuilt
}
func (c *Context) Diff(old, new [][16]byte) ([]uint32, error) { return nil, ErrNotBuilt }
func (c *Context) NewHash() (*Hash, error)                    { return nil, ErrNotBuilt }
func MD5FromETag(etag string) ([16]byte, bool)                { return [16]byte{}, false }
