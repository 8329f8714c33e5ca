#!/usr/bin/env python
"""Keep ``include/sxb200.h`` complete: every ``SXB_API`` entry point of ``csrc/*.cu`` must be declared there.

    python tools/gen_header.py            # append declarations that are missing (grouped per source file)
    python tools/gen_header.py --check    # exit 1 and list what is missing (used by tests/test_header_sync.py)

A declaration is the function's own signature as written in the source, preceded by the comment block that stands
directly above it there (that comment is the documentation of the entry point)."""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "sxb200.h")
MARK = "/* ==== declarations below are kept in step with csrc/ by tools/gen_header.py ==== */"
SIG = re.compile(r"^SXBH?_API\s+([^;{()]*?\b(sxbh?_\w+)\s*\([^;{]*?\))\s*\{", re.S | re.M)


def entry_points():
    """[(source file, name, signature text, comment text)] in source order."""
    out = []
    csrc = os.path.join(ROOT, "csrc")
    sources = [f for f in sorted(os.listdir(csrc)) if f.endswith(".cu")]
    sources += [os.path.join("host", f) for f in sorted(os.listdir(os.path.join(csrc, "host"))) if f.endswith(".cpp")]
    for fn in sources:
        text = open(os.path.join(csrc, fn)).read()
        for m in SIG.finditer(text):
            sig, name = " ".join(m.group(1).split()), m.group(2)
            # comment block immediately above
            lines = text[: m.start()].rstrip("\n").split("\n")
            block = []
            while lines and lines[-1].lstrip().startswith("//"):
                block.append(lines.pop().lstrip()[2:].rstrip())
            out.append((fn, name, sig, [b[1:] if b.startswith(" ") else b for b in reversed(block)]))
    return out


def declared(header_text):
    return set(re.findall(r"\b(sxbh?_\w+)\s*\(", header_text))


def wrap(sig, width=116):
    """Break a one-line signature at commas, continuation lines aligned behind the opening parenthesis."""
    head, args = sig.split("(", 1)
    pad = " " * (len(head) + 1)
    lines, cur = [], head + "("
    for i, piece in enumerate(args.split(", ")):
        piece = piece + ("," if i < args.count(", ") else "")
        if len(cur) + len(piece) + 1 > width and cur.strip() != head.strip() + "(":
            lines.append(cur.rstrip())
            cur = pad + piece + " "
        else:
            cur += piece + " "
    lines.append(cur.rstrip())
    return "\n".join(lines) + ";"


def _norm(sig):
    return " ".join(sig.replace("SXBH_API ", "").replace("SXB_API ", "").replace("(", " ( ").replace(")", " ) ").replace(",", " , ").split())


def stale(text, entries):
    """Entry points whose declaration in the header no longer matches the source: full text inside the generated
    section, number of parameters in the hand-written part (parameter names may differ there)."""
    cut = text.find(MARK)
    out = []
    for fn, name, sig, comment in entries:
        m = re.search(r"\bint\s+" + name + r"\s*\([^;]*\)\s*;", text)
        if not m:
            continue
        decl = m.group(0).rstrip(";").strip()
        generated = cut >= 0 and m.start() > cut
        same = _norm(decl) == _norm(sig) if generated else decl.count(",") == sig.count(",")
        if not same:
            out.append((fn, name, sig, comment, m, generated))
    return out


def main():
    text = open(HEADER).read()
    entries = entry_points()
    old = stale(text, entries)
    if "--check" in sys.argv and old:
        for fn, name, *_ in old:
            print(f"{fn}: declaration of {name} in include/sxb200.h differs from the source")
        sys.exit(1)
    for fn, name, sig, comment, m, generated in sorted(old, key=lambda e: -e[4].start()):
        if not generated:
            raise SystemExit(f"{name}: hand-written declaration is stale; edit include/sxb200.h")
        text = text[: m.start()] + wrap(sig.replace("SXBH_API ", "").replace("SXB_API ", "")) + text[m.end():]
    if old:
        open(HEADER, "w").write(text)
        print(f"refreshed {len(old)} declarations")
    have = declared(text)
    missing = [e for e in entries if e[1] not in have]
    if "--check" in sys.argv:
        for fn, name, _, _ in missing:
            print(f"{fn}: {name} is not declared in include/sxb200.h")
        sys.exit(1 if missing else 0)
    if not missing:
        print("include/sxb200.h is complete")
        return
    tail = "\n#ifdef __cplusplus\n}\n#endif\n#endif /* SXB200_H */\n"
    assert text.endswith(tail), "unexpected header epilogue"
    body = text[: -len(tail)]
    if MARK not in body:
        body += "\n" + MARK + "\n"
    cur_file = None
    for fn, name, sig, comment in missing:
        if fn != cur_file:
            body += f"\n/* ---- csrc/{fn} */\n"
            cur_file = fn
        if comment:
            body += "/* " + "\n * ".join(c.replace("*/", "* /") for c in comment) + " */\n"
        body += wrap(sig.replace("SXBH_API ", "").replace("SXB_API ", "")) + "\n"
    open(HEADER, "w").write(body + tail)
    print(f"declared {len(missing)} entry points")


if __name__ == "__main__":
    main()
