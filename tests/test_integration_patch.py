"""The core-side change of SURVEY §8f.1 is delivered as a patch against the reference tree
(integration/patches/0001-stores-parallel.diff).  Where the reference is at hand (the build container), check
that it still applies cleanly -- on copies: the reference tree is never written to."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATCH = os.path.join(ROOT, "integration", "patches", "0001-stores-parallel.diff")
REF = "/root/reference"
FILES = ["core/backend/stores.go", "pkg/grpc/client.go", "pkg/model/loader.go"]


def test_patch_names_the_three_reference_files():
    text = open(PATCH).read()
    for f in FILES:
        assert f"--- a/{f}" in text and f"+++ b/{f}" in text
    assert "model.EnableParallelRequests" in text and "model.WithModelID(storeName)" in text


@pytest.mark.skipif(not os.path.isdir(REF) or shutil.which("patch") is None, reason="needs the reference tree and patch(1)")
def test_patch_applies_to_the_reference(tmp_path):
    for f in FILES:
        dst = tmp_path / f
        dst.parent.mkdir(parents=True, exist_ok=True)
        shutil.copy(os.path.join(REF, f), dst)
    res = subprocess.run(["patch", "-p1", "--no-backup-if-mismatch", "-i", PATCH], cwd=tmp_path, capture_output=True, text=True)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "model.EnableParallelRequests" in (tmp_path / FILES[0]).read_text()
    client = (tmp_path / FILES[1]).read_text()
    assert client.count("c.stores()") == 4 and "storesOnce" in client          # the four store calls share one connection
    assert "lastAlive" in (tmp_path / FILES[2]).read_text()
