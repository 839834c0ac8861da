"""The measurement drivers under tools/ only ever run on a GPU box, where a typo costs GPU minutes: compile every
Python driver and syntax-check every shell script here; the one-call session scripts may only name files that exist."""
import glob
import os
import py_compile
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_python_drivers_compile(tmp_path):
    files = sorted(glob.glob(os.path.join(ROOT, "tools", "*.py"))) + [os.path.join(ROOT, "bench.py"),
                                                                      os.path.join(ROOT, "__graft_entry__.py")]
    assert len(files) > 10
    for f in files:
        py_compile.compile(f, cfile=str(tmp_path / (os.path.basename(f) + "c")), doraise=True)


def test_session_scripts_parse_and_reference_existing_files():
    for sh in sorted(glob.glob(os.path.join(ROOT, "tools", "*.sh"))):
        assert subprocess.run(["bash", "-n", sh]).returncode == 0, sh
        for rel in re.findall(r"\b((?:tools|tests)/[A-Za-z0-9_./]+\.(?:py|sh))\b", open(sh).read()):
            assert os.path.exists(os.path.join(ROOT, rel)), (sh, rel)
