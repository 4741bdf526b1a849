"""The N-API shim (imagestitching_b200/ts/napi/stitch_napi.cc) at least type-checks (no GPU, no Node.js).

The image has no Node.js and no node_api.h, so the shim cannot be built or run here.  This compiles it
with `g++ -fsyntax-only` against include/stitch.h and a hand-written, declaration-only subset of Node-API
(tests/napi_stub/node_api.h): every libstitch call and every napi_* call in the shim matches a prototype."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("g++") is None, reason="no g++")
def test_napi_shim_type_checks():
    src = os.path.join(ROOT, "imagestitching_b200", "ts", "napi", "stitch_napi.cc")
    r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-Wextra", "-Werror=implicit-function-declaration",
                        "-I", os.path.join(ROOT, "tests", "napi_stub"), "-I", os.path.join(ROOT, "include"), src],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-4000:]
