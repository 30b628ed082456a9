"""The C-ABI from plain C: include/gputopo.h must compile as C99 (what cgo feeds it to) and examples/extender_demo.c
must link against libgputopo.so.  Without a GPU the demo has to fail loudly at gt_create (no CPU fallback); with one it
answers a `sort` and a `bind` body."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build_demo(gt, tmp_path):
    exe = str(tmp_path / "extender_demo")
    libdir = os.path.dirname(gt.lib_path())
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I" + os.path.join(ROOT, "include"),
           os.path.join(ROOT, "examples", "extender_demo.c"), "-L" + libdir, "-lgputopo", "-Wl,-rpath," + libdir, "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_header_is_c99_and_demo_fails_loudly_without_a_gpu(gt, tmp_path):
    import torch
    exe = build_demo(gt, tmp_path)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([exe, os.path.join(ROOT, "tests", "golden", "f1_ring.topo.txt")], capture_output=True, text=True)
    assert r.returncode == 1 and "no usable CUDA device" in r.stderr


@pytest.mark.gpu
def test_demo_sort_bind_delete(gt, tmp_path):
    exe = build_demo(gt, tmp_path)
    r = subprocess.run([exe, os.path.join(ROOT, "tests", "golden", "f1_ring.topo.txt")], capture_output=True, text=True,
                       timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.splitlines()
    sort = json.loads(lines[0].split("-> ", 1)[1])
    # F1 (NV3 ring, PHB elsewhere): both nodes still have an NV3 pair free -> cost 3; the score is normalised by the
    # largest cost among the table's link kinds (SYS = 11 in the default table; user-defined codes 12-15 stay out),
    # so it is floor(10 * (11 - 3) / 11) = 7
    assert sort == [{"Host": "node-a", "Score": 7}, {"Host": "node-b", "Score": 7}]
    assert json.loads(lines[1].split("-> ", 1)[1]) == {"Error": ""}
    ann = dict(l.split(": ", 1) for l in lines[2:5])
    # node-b has GPUs 2,3 taken: the lowest-mask NV3 pair left is {0,1}
    assert ann["ALIYUN_COM_GPU_GROUP"] == "0,1" and ann["ALIYUN_COM_GPU_ASSIGNED"] == "false"
    assert lines[-1] == "pod deleted, GPUs released"
