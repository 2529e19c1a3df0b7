"""The bounds-checked build of the library (libsrt_b200_checked.so, -DSRT_CHECKED) over every kernel, against the oracle.

compute-sanitizer is closed on the GPU pool this was developed on, so the kernels carry their own checks
(csrc/srt_device.cuh SRT_CHECK): every push on a shared-memory traversal stack against the stack's capacity, every index
into a ray queue, the overflow / replay lists, the slot arrays, the candidate buffers and the sample buffer against the
array's size.  tests/kernel_driver.py renders small frames through every schedule and option (finder with its outside-ray
walk, overflow rays, tail kernel, primary cull, regenerating slots, depth 8, soup, spheres, ray entry points, BVH builder)
and compares each with the oracle; here it runs in a subprocess with SRT_LIB pointing at the checked build, and the
check-flag word must stay 0 while the self-test proves the checks are compiled in (bit 31).
"""
import os
import re
import subprocess
import sys

import pytest

import scenes

pytestmark = pytest.mark.gpu


def run_driver(lib_path):
    env = dict(os.environ)
    if lib_path:
        env["SRT_LIB"] = lib_path
    else:
        env.pop("SRT_LIB", None)
    r = subprocess.run([sys.executable, os.path.join(scenes.ROOT, "tests", "kernel_driver.py")], capture_output=True, text=True,
                       env=env, timeout=900)
    return r.returncode, r.stdout + r.stderr


def test_checked_build_sees_no_out_of_bounds_access(pkg):
    if not os.path.exists(pkg.CHECKED_LIB_PATH):
        subprocess.run(["make", "-s", "-C", os.path.dirname(pkg.CHECKED_LIB_PATH), "libsrt_b200_checked.so"], check=True)
    rc, text = run_driver(pkg.CHECKED_LIB_PATH)
    assert rc == 0 and "ALL OK" in text, text
    assert "DIFFERS" not in text
    m = re.search(r"library: (.*); check_flags over all cases 0x([0-9a-f]+); self-test raises 0x([0-9a-f]+)", text)
    assert m, text
    assert m.group(1).endswith("checked")
    assert int(m.group(2), 16) == 0, "bounds checks fired: 0x%s\n%s" % (m.group(2), text)
    assert int(m.group(3), 16) == 1 << 31  # the checks are live in this build
    assert len(re.findall(r"check_flags 0x0\b", text)) >= 14  # every rendered case reported its flag word


def test_normal_build_compiles_the_checks_away(pkg):
    rc, text = run_driver(None)
    assert rc == 0 and "ALL OK" in text, text
    m = re.search(r"library: (.*); check_flags over all cases 0x([0-9a-f]+); self-test raises 0x([0-9a-f]+)", text)
    assert m and not m.group(1).endswith("checked")
    assert int(m.group(2), 16) == 0 and int(m.group(3), 16) == 0
