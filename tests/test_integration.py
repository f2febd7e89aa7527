"""CPU: the drop-in of INTEGRATION.md section 1, performed on a scratch copy of the reference tree
(only where /root/reference exists -- it does not travel to the GPU box).  The UNMODIFIED reference
CLI must import the replacement `_ldscore`, build LDScoreParams with its own keyword set and reach
nldsc_ld_calculate; without a GPU that call fails loudly (no CPU fallback), which is what is
asserted here.  With a GPU the same command completes and writes the TSV."""
import os
import shutil
import subprocess
import sys

import pytest

from nldsc_b200 import build, capi, synth

REF = "/root/reference/nldsc"


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present")
def test_reference_cli_runs_on_the_replacement_module(tmp_path):
    build.build_module()
    tree = tmp_path / "nldsc"
    shutil.copytree(REF, tree, ignore=shutil.ignore_patterns("_ldscore", "__pycache__"))
    shutil.copy(capi.LIB_PATH, tree / "ldscore" / "libnldsc_ld.so")
    shutil.copy(build.ext_path(), tree / "ldscore" / "_ldscore.so")          # the slot of reference Makefile:8
    spec = synth.SynthSpec(n_org=200, n_snp=300, missing=0.02, seed=31)
    synth.write_bfile(str(tmp_path / "g"), spec)
    out = tmp_path / "g.l2"
    r = subprocess.run([sys.executable, "nldsc", "ld", "--bfile", str(tmp_path / "g"), "--ld-wind-cm", "1",
                        "-maf", "0.0001", "--std-thr", "1e-5", "--out", str(out), "--extra"],
                       cwd=tmp_path, capture_output=True, text=True, timeout=300)
    log = r.stdout + r.stderr
    assert "Can't load `ldscore` sub-package" not in log, log       # routines.py:5-10 import guard not tripped
    if os.path.exists("/dev/nvidiactl"):
        assert r.returncode == 0, log
        head = out.read_text().splitlines()[0].split("\t")
        assert head == ["CHR", "SNP", "BP", "L2", "L2D", "MAF", "WSA", "WSD", "WSDE", "RSTD"]
    else:
        assert "no usable CUDA device" in log, log                  # reached the C ABI; failed loudly
        assert not out.exists()
