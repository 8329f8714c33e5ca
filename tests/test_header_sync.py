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


def _header_params():
    """name -> list of parameter type classes ('p' pointer, 'i' 32-bit integer, 'l' 64-bit integer, 'f' float, 'd')."""
    import re
    text = open(os.path.join(ROOT, "include", "sxb200.h")).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    out = {}
    for m in re.finditer(r"\b(?:int|long long|unsigned long long)\s+(sxb_\w+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
        name, args = m.group(1), " ".join(m.group(2).split())
        kinds = []
        if args not in ("", "void"):
            for a in args.split(","):
                a = a.strip()
                if "*" in a:
                    kinds.append("p")
                elif "long long" in a:
                    kinds.append("l")
                elif "double" in a:
                    kinds.append("d")
                elif "float" in a:
                    kinds.append("f")
                else:
                    kinds.append("i")
        out[name] = kinds
    return out


def _registered_signatures():
    """Every ``{"sxb_name": "signature"}`` entry of the package's signature tables (adjacent string literals are
    already joined by the parser)."""
    import ast
    sigs = {}
    pkg = os.path.join(ROOT, "shuffle_exchange_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if not f.endswith(".py"):
                continue
            tree = ast.parse(open(os.path.join(dirpath, f)).read())
            for node in ast.walk(tree):
                if isinstance(node, ast.Dict):
                    for k, v in zip(node.keys, node.values):
                        if isinstance(k, ast.Constant) and isinstance(k.value, str) and k.value.startswith("sxb_") \
                                and isinstance(v, ast.Constant) and isinstance(v.value, str):
                            sigs.setdefault(k.value, set()).add(v.value)
    return sigs


def test_ctypes_signatures_match_the_header():
    """A C entry point that gained or lost a parameter while its ctypes signature did not would corrupt the call's
    arguments on the GPU box only; the header (kept in step with csrc/ by tools/gen_header.py) is the referee."""
    header = _header_params()
    bad = []
    for name, variants in sorted(_registered_signatures().items()):
        if name not in header:
            continue                                   # host-library or removed names are covered elsewhere
        for sig in variants:
            want = "".join(header[name])
            # 'i' in a ctypes table may stand for any 32-bit integer; pointers, 64-bit integers and floats must agree
            if len(sig) != len(want) or any(a != b for a, b in zip(sig, want)):
                bad.append((name, sig, want))
    assert not bad, bad


def test_header_parameter_types_match_the_sources():
    """Hand-written declarations are compared with the definitions in csrc/ by parameter TYPE (names may differ)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("gen_header", os.path.join(ROOT, "tools", "gen_header.py"))
    gh = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gh)

    def kinds(args):
        out = []
        args = " ".join(args.split())
        if args in ("", "void"):
            return out
        for a in args.split(","):
            a = a.split("=")[0].strip()
            out.append("p" if "*" in a else "l" if "long long" in a else "d" if "double" in a
                       else "f" if "float" in a else "i")
        return out
    header = _header_params()
    bad = []
    for fn, name, sig, _ in gh.entry_points():
        args = sig[sig.index("(") + 1: sig.rindex(")")]
        if name in header and header[name] != kinds(args):
            bad.append((fn, name, "".join(header[name]), "".join(kinds(args))))
    assert not bad, bad


def test_host_library_signatures_match_the_header():
    """Same check for libsxb200_host.so: its ctypes tables list argument types as ``[c_i, c_p, ...]``."""
    import ast
    import re
    text = re.sub(r"/\*.*?\*/", " ", open(os.path.join(ROOT, "include", "sxb200.h")).read(), flags=re.S)
    header = {}
    for m in re.finditer(r"\b(?:int|long long)\s+(sxbh_\w+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
        args = " ".join(m.group(2).split())
        kinds = []
        if args not in ("", "void"):
            for a in args.split(","):
                kinds.append("p" if "*" in a else "l" if "long long" in a else "d" if "double" in a
                             else "f" if "float" in a else "i")
        header[m.group(1)] = "".join(kinds)
    letter = {"c_i": "i", "c_p": "p", "c_l": "l", "c_f": "f", "c_d": "d", "c_int": "i", "c_void_p": "p",
              "c_longlong": "l", "c_float": "f", "c_double": "d", "c_char_p": "p", "_P": "p"}
    bad, seen = [], 0
    pkg = os.path.join(ROOT, "shuffle_exchange_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if not f.endswith(".py"):
                continue
            for node in ast.walk(ast.parse(open(os.path.join(dirpath, f)).read())):
                if not isinstance(node, ast.Dict):
                    continue
                for k, v in zip(node.keys, node.values):
                    if not (isinstance(k, ast.Constant) and isinstance(k.value, str) and k.value.startswith("sxbh_")
                            and isinstance(v, ast.List)):
                        continue
                    names = [e.id if isinstance(e, ast.Name) else getattr(e, "attr", "?") for e in v.elts]
                    sig = "".join(letter.get(n, "?") for n in names)
                    seen += 1
                    if k.value in header and sig != header[k.value]:
                        bad.append((f, k.value, sig, header[k.value]))
    assert seen >= 20 and not bad, (seen, bad)
