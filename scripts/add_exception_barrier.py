"""One-off source transformation (kept for the record): turn the body of every extern "C" entry point in csrc/ into a
function-try-block ending in B2_ABI_CATCH*, so that no C++ exception (std::bad_alloc from new / std::vector,
std::system_error from std::async, ...) crosses the C ABI.  Idempotent."""
import re
import sys

HEAD = re.compile(r'^(?:extern "C" )?(int|void|const char\*|void\*)\s+(b200_[a-z0-9_]+)\s*\(', re.M)


def transform(src: str) -> str:
    out, pos = [], 0
    for m in HEAD.finditer(src):
        if m.start() < pos:
            continue
        # end of the parameter list
        i, depth = m.end(), 1
        while depth:
            depth += {"(": 1, ")": -1}.get(src[i], 0)
            i += 1
        j = i
        while src[j] in " \n\t":
            j += 1
        if src[j] != "{":        # a declaration, or already a function-try-block
            continue
        params = src[m.end():i - 1]
        # matching closing brace of the body
        k, depth = j + 1, 1
        in_str = in_chr = False
        while depth:
            c = src[k]
            if c == '"' and src[k - 1] != "\\" and not in_chr:
                in_str = not in_str
            elif c == "'" and src[k - 1] != "\\" and not in_str:
                in_chr = not in_chr
            elif not in_str and not in_chr:
                if src.startswith("//", k):
                    k = src.index("\n", k)
                    continue
                depth += {"{": 1, "}": -1}.get(c, 0)
            k += 1
        ret = m.group(1)
        has_ctx = re.search(r"\bb200_ctx\s*\*\s*ctx\b", params) and "const b200_ctx" not in params
        if ret == "int":
            tail = " B2_ABI_CATCH_CTX(ctx)" if has_ctx else " B2_ABI_CATCH_INT"
        elif ret == "void":
            tail = " B2_ABI_CATCH_VOID"
        else:
            tail = " B2_ABI_CATCH_NULL"
        out.append(src[pos:j] + "try {" + src[j + 1:k] + tail)
        pos = k
    out.append(src[pos:])
    return "".join(out)


for path in sys.argv[1:]:
    s = open(path).read()
    t = transform(s)
    if t != s:
        open(path, "w").write(t)
        print(path, "entry points wrapped:", t.count("B2_ABI_CATCH") - s.count("B2_ABI_CATCH"))
