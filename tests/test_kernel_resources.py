"""Resource budget of the kernels whose speed depends on it (no GPU needed: ptxas reports).

lk_track runs 16 blocks of 64 threads per SM with the L1 carved down to almost nothing in
favour of shared memory: it needs exactly 64 registers and NO spills - two spilled values in the
level prologue cost it 17 % on B200."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "stereo_vo_b200", "csrc")


def _ptxas_report(src, tmp_path):
    from stereo_vo_b200 import build
    nvcc = build._nvcc()
    if shutil.which(nvcc) is None and not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    flags = [f for f in build.NVCC_FLAGS if f not in ("--threads", "0")]
    out = subprocess.run([nvcc, *flags, "-Xptxas", "-v", "-c", os.path.join(CSRC, src),
                          "-o", str(tmp_path / "k.o")], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    return out.stderr + out.stdout


def test_lk_track_has_no_spills_and_64_registers(tmp_path):
    rep = _ptxas_report("lk_track.cu", tmp_path)
    blocks = re.findall(r"Compiling entry function '([^']*lk_track_kernel[^']*)'.*?"
                        r"(\d+) bytes spill stores, (\d+) bytes spill loads.*?Used (\d+) registers",
                        rep, re.S)
    assert blocks, rep
    for name, st, ld, regs in blocks:
        assert (int(st), int(ld)) == (0, 0), f"{name}: {st} B spill stores / {ld} B spill loads"
        assert int(regs) <= 64, f"{name}: {regs} registers (16 blocks per SM need <= 64)"
