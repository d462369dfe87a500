"""Regenerates the per-curve symbol-name block of include/b200prover.h: in the BLS12-377 build every entry point
b200_x is exported as b200_bls12377_x, so that both libraries can be linked into ONE binary (gnark's groth16.go links
every curve; ELF global symbols of two shared objects with the same name would all resolve to the first one loaded).
Run after adding an entry point:  python scripts/gen_symbol_renames.py"""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATH = os.path.join(ROOT, "include", "b200prover.h")
BEGIN, END = "/* BEGIN per-curve symbol names (generated: scripts/gen_symbol_renames.py) */", "/* END per-curve symbol names */"


def functions(src):
    body = src[src.index(END):] if END in src else src
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", body)))


def main():
    src = open(PATH).read()
    names = functions(src)
    block = [BEGIN, "#if defined(B200_CURVE_BLS12_377) || defined(B2_CURVE_BLS12_377)"]
    block += [f"#define {n} b200_bls12377_{n[5:]}" for n in names]
    block += ["#endif", END]
    text = "\n".join(block)
    if BEGIN in src:
        src = src[:src.index(BEGIN)] + text + src[src.index(END) + len(END):]
    else:
        anchor = "typedef struct b200_ctx b200_ctx;"
        src = src.replace(anchor, text + "\n\n" + anchor, 1)
    open(PATH, "w").write(src)
    print(len(names), "entry points")


if __name__ == "__main__":
    main()
