"""Static SASS instruction histograms of the hot kernels (cuobjdump -sass on the built objects): the evidence behind the
IMAD.WIDE / IMAD.MOV / LDL / STL counts quoted in DESIGN.md and profiles/.  python scripts/sass_histogram.py > profiles/rNN_sass_histograms.txt"""
import collections, re, subprocess, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJS = {"bn254": os.path.join(ROOT, "gnark-fork_b200", "csrc", "build"), "bls12_377": os.path.join(ROOT, "gnark-fork_b200", "csrc", "build377")}
WANT = ["msm_accumulate_kernel", "msm_accumulate_smem_kernel", "ba_prep_kernel", "ba_apply_kernel", "ba_inverse_kernel", "ntt_pass_kernel", "ntt_pass_tma_kernel",
        "fp2_mul_call", "fp2_sqr_call", "mul_sub_kernel", "msm_digits_kernel", "msm_scatter_kernel"]
KEYS = ["IMAD.WIDE", "IMAD.MOV", "IMAD.X", "IMAD.HI", "IMAD", "IADD3", "LOP3", "SEL", "MOV", "LDG", "STG", "LDL", "STL", "LDS", "STS", "CALL", "UTMALDG", "UTMASTG", "SYNCS", "CCTL", "BRA"]
def demangle(names):
    out = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))
for curve, d in OBJS.items():
    if not os.path.isdir(d):
        continue
    for obj in sorted(os.listdir(d)):
        if not obj.endswith(".o") or not obj.startswith(("msm_", "ntt")):
            continue
        sass = subprocess.run(["cuobjdump", "-sass", os.path.join(d, obj)], capture_output=True, text=True).stdout
        funcs, cur = collections.OrderedDict(), None
        for line in sass.splitlines():
            m = re.search(r"Function : (\S+)", line)
            if m:
                cur = m.group(1); funcs[cur] = collections.Counter(); continue
            m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_.]*)", line)
            if m and cur:
                funcs[cur][m.group(1)] += 1
        names = demangle(list(funcs))
        for f, c in funcs.items():
            nm = names.get(f, f)
            if not any(w in nm for w in WANT):
                continue
            agg = collections.OrderedDict((k, 0) for k in KEYS)
            for op, n in c.items():
                for k in KEYS:
                    if op == k or op.startswith(k + "."):
                        agg[k] += n; break
            short = re.sub(r"\(.*", "", nm).replace("b200::", "").replace("void ", "")
            print(f"{curve:9s} {obj:10s} {short[:70]:70s} total {sum(c.values()):6d}  " + " ".join(f"{k}={v}" for k, v in agg.items() if v))
