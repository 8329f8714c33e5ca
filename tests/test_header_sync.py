"""``include/sxb200.h`` declares every ``SXB_API`` entry point of ``csrc/*.cu`` and compiles as C and as C++."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_declares_every_entry_point():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_header.py"), "--check"], capture_output=True,
                       text=True)
    assert r.returncode == 0, r.stdout


def test_header_compiles_as_c_and_cxx(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "sxb200.h"\nint main(void) { return sxb_abi_version == 0; }\n')
    inc = os.path.join(ROOT, "include")
    for cc, extra in (("gcc", []), ("g++", ["-x", "c++"])):
        r = subprocess.run([cc, *extra, "-fsyntax-only", "-I", inc, str(src)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
