"""SASS evidence for every kernel of the library, read from the built objects (no GPU needed).

    python tools/sass_listing.py            # writes profiles/sass_mnemonics.txt and profiles/sass/<object>.sass.txt

``sass_mnemonics.txt``: per object and per kernel, how often each Blackwell-relevant instruction occurs (tcgen05 MMA =
``UTC*MMA``, TMEM load/store = ``LDTM``/``STTM``, TMA = ``UTMALDG``/``UTMASTG``/``UBLKCP``, NVLS = ``LDGMC``/``*.MC``,
legacy tensor path = ``HMMA``).  ``sass/<object>.sass.txt``: the instructions themselves (address + text) for those
mnemonics, per kernel, so each claim can be checked against an address in ``cuobjdump -sass``."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "build", "obj")
OUT = os.path.join(ROOT, "profiles")
KEY = re.compile(r"\b(UTC[A-Z]*MMA[\w.]*|UTCBAR[\w.]*|UTCCP[\w.]*|LDTM[\w.]*|STTM[\w.]*|UTMALDG[\w.]*|UTMASTG[\w.]*|UBLKCP[\w.]*|"
                 r"UTMAPF[\w.]*|LDGMC[\w.]*|REDG?MC[\w.]*|STGMC[\w.]*|HMMA[\w.]*|SYNCS[\w.]*|UCGABAR[\w.]*|ELECT[\w.]*|"
                 r"F2FP[\w.]*E4M3[\w.]*|LDGSTS[\w.]*|MEMBAR[\w.]*SYS|FENCE[\w.]*)")
INSTR = re.compile(r"/\*([0-9a-f]{4,})\*/\s+(.*?);")


def demangle(name):
    try:
        return subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip() or name
    except Exception:
        return name


def main():
    os.makedirs(os.path.join(OUT, "sass"), exist_ok=True)
    summary = []
    for fn in sorted(os.listdir(OBJ)):
        if not fn.endswith(".o") or fn.startswith("host_"):
            continue
        txt = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, fn)], capture_output=True, text=True).stdout
        kernels = collections.OrderedDict()
        cur = None
        for line in txt.splitlines():
            m = re.search(r"Function : (\S+)", line)
            if m:
                cur = demangle(m.group(1))
                cur = re.sub(r"\(.*", "", cur)[:120]
                kernels[cur] = {"n": 0, "hits": [], "hist": collections.Counter()}
                continue
            if cur is None:
                continue
            mi = INSTR.search(line)
            if not mi:
                continue
            kernels[cur]["n"] += 1
            mk = KEY.search(mi.group(2))
            if mk:
                kernels[cur]["hist"][mk.group(1)] += 1
                kernels[cur]["hits"].append(f"    /*{mi.group(1)}*/ {mi.group(2).strip()}")
        summary.append(f"== {fn}")
        with open(os.path.join(OUT, "sass", fn[:-2] + ".sass.txt"), "w") as f:
            f.write(f"# key instructions of {fn} (cuobjdump -sass, sm_100a); full listing: cuobjdump -sass build/obj/{fn}\n")
            for k, d in kernels.items():
                hist = ", ".join(f"{a} x{b}" for a, b in sorted(d["hist"].items()))
                summary.append(f"  {k}: {d['n']} instr; {hist or 'no tensor/TMA/NVLS instructions (SIMT kernel)'}")
                f.write(f"\n{k}  [{d['n']} instructions]\n")
                f.write("\n".join(d["hits"][:400]) + ("\n" if d["hits"] else ""))
                if len(d["hits"]) > 400:
                    f.write(f"    ... {len(d['hits']) - 400} more\n")
    with open(os.path.join(OUT, "sass_mnemonics.txt"), "w") as f:
        f.write("Per-kernel SASS mnemonic histogram of build/obj/*.o (tools/sass_listing.py; instruction text per address in "
                "profiles/sass/)\n" + "\n".join(summary) + "\n")
    print("\n".join(summary[:60]))


if __name__ == "__main__":
    sys.exit(main())
