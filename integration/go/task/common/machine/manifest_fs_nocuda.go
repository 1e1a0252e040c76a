//go:build !(cuda && linux && amd64)

package machine

import (
	"context"
	"errors"
	"os"

	"github.com/rclone/rclone/fs"
)

// wrapLocal in the default build (CGO_ENABLED=0): the reference's behaviour, unless the GPU path was asked
// for -- then the Transfer fails instead of quietly hashing on the CPU.
func wrapLocal(ctx context.Context, f fs.Fs, root string) (fs.Fs, error) {
	if os.Getenv("TPI_GPU_CHECKSUM") == "1" {
		return nil, errors.New("TPI_GPU_CHECKSUM=1, but this binary was built without -tags cuda")
	}
	return f, nil
}
