"""The files a TPI maintainer would take from integration/: they must apply to the reference as it is.

Runs only where the reference tree is present (this container; the GPU box has none, and nothing in the
-m gpu tests reads it).  The renderer is first pinned on the reference's own goldens (unpatched template ->
byte-identical to testdata/machine_script_{minimal,full}.golden, script_test.go:14-38); then the patched
template is rendered and must differ from those goldens in exactly the two change detectors."""
import difflib
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
MACHINE = os.path.join(REF, "task", "common", "machine")
pytestmark = pytest.mark.skipif(not os.path.isdir(MACHINE), reason="the reference tree is not on this machine")

sys.path.insert(0, os.path.join(ROOT, "integration"))


def test_renderer_reproduces_the_reference_goldens():
    import render_machine_script as r
    tpl = open(os.path.join(MACHINE, "machine-script.sh.tpl")).read()
    for name, kw in r.CASES.items():
        want = open(os.path.join(MACHINE, "testdata", name + ".golden")).read()
        assert r.render(tpl, **kw) == want, name


def test_patched_template_changes_exactly_the_two_detectors(tmp_path):
    import render_machine_script as r
    assert r.main.__module__  # imported
    sys.argv = ["render", REF, str(tmp_path)]
    assert r.main() == 0
    for name in r.CASES:
        old = open(os.path.join(MACHINE, "testdata", name + ".golden")).read().splitlines()
        new = open(tmp_path / (name + ".golden")).read().splitlines()
        removed = [l[1:] for l in difflib.unified_diff(old, new, lineterm="", n=0) if l.startswith("-") and not l.startswith("---")]
        added = [l[1:] for l in difflib.unified_diff(old, new, lineterm="", n=0) if l.startswith("+") and not l.startswith("+++")]
        assert len(removed) == 4 and len(added) == 2, (removed, added)
        assert "md5sum" in removed[0] and "find " in removed[1] and "EPOCH" in removed[2] and "EPOCH" in removed[3]
        assert added[0] == removed[0].replace("md5sum", "tpimd5sum")
        assert "tpimd5sum --changed" in added[1] and added[1].rstrip().endswith("then")
        # everything else (rclone sync lines, loop cadence, service unit) is untouched
        assert sum(1 for l in new if "rclone sync" in l) == sum(1 for l in old if "rclone sync" in l) == 2
    # bash accepts the result
    subprocess.check_call(["bash", "-n", str(tmp_path / "machine_script_full.golden")])


def test_storage_go_patch_applies_to_the_reference(tmp_path):
    out = tmp_path / "storage.go"
    subprocess.check_call(["patch", "--quiet", "-o", str(out), os.path.join(MACHINE, "storage.go"),
                           os.path.join(ROOT, "integration", "go", "task", "common", "machine", "storage.go.patch")])
    text = out.read_text()
    assert text.count("wrapLocal(ctx, ") == 2 and "sync.CopyDir(ctx, destinationFileSystem, sourceFileSystem, true)" in text


def test_go_files_are_laid_out_where_the_reference_would_take_them():
    for rel in ("go/cuda/tpimd5.go", "go/cuda/tpimd5_nocuda.go", "go/task/common/machine/manifest_fs.go",
                "go/task/common/machine/manifest_fs_nocuda.go"):
        text = open(os.path.join(ROOT, "integration", rel)).read()
        assert text.startswith("//go:build "), rel
    cgo = open(os.path.join(ROOT, "integration", "go", "cuda", "tpimd5.go")).read()
    header = open(os.path.join(ROOT, "include", "tpimd5.h")).read()
    import re
    for sym in set(re.findall(r"C\.(tpimd5_[a-z0-9_]+)\(", cgo)):
        assert re.search(r"\b" + sym + r"\(", header), f"{sym} is bound in Go but not declared in include/tpimd5.h"
