"""The launch plans count on a number of resident CTAs per SM; the register allocation of a kernel
can silently take that away — in round 2 a waiting loop that only lane 0 ran moved the wide beam kernel from 48 to 80
registers (6 CTAs per SM instead of 8, -12 % throughput) with every test still green.  This reads the built library's
resource usage (cuobjdump, no GPU needed) and holds every hot kernel to the registers its plan allows:
registers x threads per CTA x planned CTAs <= 65,536."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "fast-ctc-decode_b200", "libfctc_b200.so")

# mangled-name fragment -> (threads per CTA, CTAs per SM the plan counts on)
PLANS = {
    "18beam_search_kernelILb0ELi5ELb0E": (128, 8),   # wide beam, five labels, f32: 8 CTAs of 4 warps (shared memory plan)
    "18beam_search_kernelILb0ELi5ELb1E": (128, 8),   # the same, 16-bit posteriors
    "18beam_search_kernelILb1ELi5ELb0E": (128, 8),   # crf
    "17beam_small_kernelILb0ELi5ELb0E": (128, 8),    # narrow beam: 32 warps per SM
    "21viterbi_stream_kernelIhLi4ELb1ELb1ELb1ELb0ELb0E": (128, 5),  # headline: qstring, collapse, five labels, f32 (5 CTAs by plan)
    "21viterbi_stream_kernelIhLi4ELb1ELb0ELb1ELb0ELb0E": (128, 3),  # without quality strings: 3 CTAs by plan
    "20duplex_packed_kernelILi0E": (32, 16),         # exact exp: 16 single-warp CTAs
    "20duplex_packed_kernelILi1E": (32, 24),
    "20duplex_packed_kernelILi2E": (32, 24),
}


def resource_usage():
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe):
        pytest.skip("cuobjdump not found")
    if not os.path.exists(LIB):
        pytest.skip("library not built")
    out = subprocess.run([exe, "-res-usage", LIB], capture_output=True, text=True, timeout=300).stdout
    usage = {}
    name = None
    for line in out.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            name = m.group(1)
            continue
        m = re.search(r"REG:(\d+)", line)
        if m and name:
            usage[name] = int(m.group(1))
            name = None
    return usage


def test_registers_leave_room_for_the_planned_ctas():
    usage = resource_usage()
    assert usage, "no kernels found in the library"
    for frag, (threads, ctas) in PLANS.items():
        hits = {k: v for k, v in usage.items() if frag in k}
        assert hits, f"kernel {frag} not found (renamed? update PLANS)"
        for k, regs in hits.items():
            alloc = (regs + 7) // 8 * 8  # registers are granted in units of 8 per thread
            assert alloc * threads * ctas <= 65536, f"{k}: {regs} registers x {threads} threads x {ctas} CTAs exceed the register file"
