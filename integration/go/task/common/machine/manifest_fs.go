//go:build cuda && linux && amd64

// manifest_fs.go -- puts a GPU-built MD5 manifest under rclone's Object.Hash for LOCAL file systems.
//
// machine.Transfer (storage.go:123-159) creates two fs.Fs (storage.go:140-148) and hands them to
// sync.CopyDir (storage.go:158).  With --checksum, rclone then calls Object.Hash(MD5) once per object and
// side, one file at a time, on 8 checker goroutines.  wrapLocal lists a local side once, hashes the whole
// listing in one tpimd5_hash_files call, and answers every later Hash(MD5) from that manifest -- the
// pattern of rclone's own hasher overlay backend.  Both sides of a local->local sync are wrapped
// (task/k8s/task.go:294-299, and the reference's own unit test).
//
// SOURCE ONLY: never compiled (no Go toolchain, no rclone module on the build image).
package machine

import (
	"context"
	"encoding/hex"
	"fmt"
	"io"
	"os"
	"path/filepath"
	"time"

	"github.com/rclone/rclone/fs"
	"github.com/rclone/rclone/fs/hash"
	"github.com/rclone/rclone/fs/walk"

	"terraform-provider-iterative/cuda"
)

// gpuChecksumEnabled: the GPU path is opt-in per process.
func gpuChecksumEnabled() bool { return os.Getenv("TPI_GPU_CHECKSUM") == "1" }

// gpuContext is shared by every Transfer of the plugin process; the library overlaps concurrent calls.
var gpuContext *cuda.Context

func sharedGPU() (*cuda.Context, error) {
	if gpuContext != nil {
		return gpuContext, nil
	}
	c, err := cuda.NewContext(nil, 0)
	if err != nil {
		return nil, err // no CPU fallback: a GPU build that cannot reach its GPU fails the Transfer
	}
	gpuContext = c
	return c, nil
}

type manifestEntry struct {
	md5   string // lowercase hex, rclone's hash string form
	size  int64
	mtime time.Time
}

// manifestFs wraps a local fs.Fs; every object it hands out answers Hash(MD5) from the manifest.
type manifestFs struct {
	fs.Fs
	features *fs.Features
	root     string
	entries  map[string]manifestEntry // remote (slash-separated relative path) -> digest and what was listed
}

type manifestObject struct {
	fs.Object
	parent *manifestFs
}

// wrapLocal returns f unchanged unless the GPU path is switched on and f is a local file system.
func wrapLocal(ctx context.Context, f fs.Fs, root string) (fs.Fs, error) {
	if !gpuChecksumEnabled() || !f.Features().IsLocal {
		return f, nil
	}
	gpu, err := sharedGPU()
	if err != nil {
		return nil, err
	}
	var remotes, paths []string
	var sizes []uint64
	var mtimes []time.Time
	// walk.ListR applies the filter that Transfer installed into ctx (storage.go:124-138), so this is
	// exactly the file set sync.CopyDir will see; the local backend skips symlinks by default.
	err = walk.ListR(ctx, f, "", true, -1, walk.ListObjects, func(entries fs.DirEntries) error {
		for _, e := range entries {
			if o, ok := e.(fs.Object); ok {
				remotes = append(remotes, o.Remote())
				paths = append(paths, filepath.Join(root, filepath.FromSlash(o.Remote())))
				sizes = append(sizes, uint64(o.Size()))
				mtimes = append(mtimes, o.ModTime(ctx))
			}
		}
		return nil
	})
	if err == fs.ErrorDirNotFound {
		err = nil // an empty destination: nothing to hash
	}
	if err != nil {
		return nil, err
	}
	digests, failed, err := gpu.HashFiles(paths, sizes) // tpimd5_hash_files: one GPU pass over the listing
	if err != nil {
		if failed >= 0 {
			return nil, fmt.Errorf("%s: %w", remotes[failed], err)
		}
		return nil, err
	}
	m := &manifestFs{Fs: f, root: root, entries: make(map[string]manifestEntry, len(remotes))}
	for i, r := range remotes {
		m.entries[r] = manifestEntry{md5: hex.EncodeToString(digests[i][:]), size: int64(sizes[i]), mtime: mtimes[i]}
	}
	m.features = (&fs.Features{}).Fill(ctx, m).Mask(ctx, f).WrapsFs(m, f)
	return m, nil
}

func (f *manifestFs) Features() *fs.Features { return f.features }
func (f *manifestFs) Hashes() hash.Set       { return hash.Set(hash.MD5) }

func (f *manifestFs) wrap(o fs.Object) fs.Object {
	if o == nil {
		return nil
	}
	return &manifestObject{Object: o, parent: f}
}

func (f *manifestFs) NewObject(ctx context.Context, remote string) (fs.Object, error) {
	o, err := f.Fs.NewObject(ctx, remote)
	if err != nil {
		return nil, err
	}
	return f.wrap(o), nil
}

func (f *manifestFs) List(ctx context.Context, dir string) (fs.DirEntries, error) {
	entries, err := f.Fs.List(ctx, dir)
	for i, e := range entries {
		if o, ok := e.(fs.Object); ok {
			entries[i] = f.wrap(o)
		}
	}
	return entries, err
}

// Put and friends write through the wrapped Fs; what they return is NOT in the manifest (it was written after
// the listing), so its Hash goes to the GPU on demand (see Hash below).
func (f *manifestFs) Put(ctx context.Context, in io.Reader, src fs.ObjectInfo, options ...fs.OpenOption) (fs.Object, error) {
	o, err := f.Fs.Put(ctx, in, src, options...)
	if err != nil {
		return nil, err
	}
	delete(f.entries, src.Remote())
	return f.wrap(o), nil
}

func (o *manifestObject) Fs() fs.Info { return o.parent }

// Hash: lowercase hex for MD5; other types are unsupported, as on any backend that lacks them.
//
// The local backend re-lstats before it trusts a cached hash and drops it when size or mtime moved.  The same
// here: an object that is in the manifest and still has the listed size and mtime is answered from it.  One
// that changed since the listing is REFUSED (an error, not a CPU hash: the digest on file belongs to other
// bytes, and this build has no CPU path).  One that is not in the manifest (copied in after the listing:
// rclone's post-transfer verify asks for it) is hashed on the GPU there and then.
func (o *manifestObject) Hash(ctx context.Context, t hash.Type) (string, error) {
	if t != hash.MD5 {
		return "", hash.ErrUnsupported
	}
	path := filepath.Join(o.parent.root, filepath.FromSlash(o.Remote()))
	st, err := os.Lstat(path)
	if err != nil {
		return "", err
	}
	if e, ok := o.parent.entries[o.Remote()]; ok {
		if st.Size() != e.size || !st.ModTime().Equal(e.mtime) {
			return "", fmt.Errorf("%s: changed since it was listed (size %d -> %d, mtime %v -> %v); refusing a stale MD5",
				o.Remote(), e.size, st.Size(), e.mtime, st.ModTime())
		}
		return e.md5, nil
	}
	gpu, err := sharedGPU()
	if err != nil {
		return "", err
	}
	d, _, err := gpu.HashFiles([]string{path}, []uint64{uint64(st.Size())})
	if err != nil {
		return "", err
	}
	return hex.EncodeToString(d[0][:]), nil
}
