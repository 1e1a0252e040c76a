#!/usr/bin/env python3
"""Applies integration/machine-script.sh.tpl.patch to the reference's VM script template and renders it the
way the reference's Script() does (task/common/machine/script.go:22-57): base64 of the task script, of the
`NAME="value"` environment lines and of the `export NAME=value` credential lines, and the timeout.

    python integration/render_machine_script.py REFERENCE_ROOT OUT_DIR

writes OUT_DIR/machine-script.sh.tpl (patched) and OUT_DIR/machine_script_{minimal,full}.golden, the two
goldens script_test.go:14-38 pins, rendered from the patched template.  The reference's files are read where
they lie; nothing of them is copied into this repository (only the patch and this renderer are tracked).
"""
import base64
import os
import shlex
import subprocess
import sys


def render(template: str, script: str, credentials=None, variables=None, timeout=None) -> str:
    """script.go:22-57.  shellescape.Quote == shlex.quote for the characters these tests use."""
    environment = "".join(f'{k}="{v.replace(chr(34), chr(92) + chr(34))}"\n' for k, v in (variables or {}).items())
    creds = "".join("export " + shlex.quote(f"{k}={v}") + "\n" for k, v in (credentials or {}).items())
    b64 = lambda s: base64.b64encode(s.encode()).decode()
    out = template
    for key, val in (("TaskScript", b64(script)), ("Environment", b64(environment)), ("Credentials", b64(creds)),
                     ("Timeout", "infinity" if timeout is None else str(int(timeout)))):
        out = out.replace("{{." + key + "}}", val)
    return out


# the two cases of script_test.go:14-38
TEST_SCRIPT = "\n"          # `\n#!/bin/sh\necho "done"\n`[:1]
CASES = {
    "machine_script_minimal": dict(script=TEST_SCRIPT),
    "machine_script_full": dict(script=TEST_SCRIPT, credentials={"SECRET": "VALUE"}, variables={"KEY": "VALUE"}, timeout=1659919333),
}


def patched_template(reference_root: str, out_dir: str) -> str:
    here = os.path.dirname(os.path.abspath(__file__))
    src = os.path.join(reference_root, "task", "common", "machine", "machine-script.sh.tpl")
    dst = os.path.join(out_dir, "machine-script.sh.tpl")
    os.makedirs(out_dir, exist_ok=True)
    subprocess.check_call(["patch", "--quiet", "-o", dst, src, os.path.join(here, "machine-script.sh.tpl.patch")])
    return open(dst).read()


def main() -> int:
    reference_root, out_dir = sys.argv[1], sys.argv[2]
    tpl = patched_template(reference_root, out_dir)
    for name, kw in CASES.items():
        with open(os.path.join(out_dir, name + ".golden"), "w") as f:
            f.write(render(tpl, **kw))
    return 0


if __name__ == "__main__":
    sys.exit(main())
