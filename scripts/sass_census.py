"""SASS evidence for the built library (north_star: "each kernel's design choices are evidenced by committed SASS").
Runs `cuobjdump -sass` on liblg_b200.so and writes
  profiles/<tag>_sass_census.txt        per kernel: instruction count and the counts of the mnemonics that prove the
                                        design (UTCHMMA = tcgen05.mma, .2CTA = cta_group::2, LDTM / STTM = tcgen05.ld / st,
                                        UBLKCP = cp.async.bulk (TMA engine), UTMALDG / UTMASTG = tensor-map TMA, UTCBAR =
                                        tcgen05.commit, SYNCS = mbarrier, MUFU.EX2, FFMA2 / FADD2 packed fp32, REDUX, ATOM/RED)
  profiles/<tag>_sass_hot_<kernel>.txt  the SASS around the tensor-core issue of attention, linear, fused FFN and the
                                        assignment passes, and the whole softmax sweep of the attention kernel
usage: sass_census.py [tag] [lib.so]"""
import collections, os, re, subprocess, sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
tag = sys.argv[1] if len(sys.argv) > 1 else "r2"
lib = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "cs566-project-lightglue_b200", "liblg_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout.split("\n")
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip() or n

kernels = collections.OrderedDict()
cur = None
for l in txt:
    m = re.search(r"Function : (\S+)", l)
    if m:
        cur = kernels.setdefault(m.group(1), [])
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m and cur is not None:
        cur.append((int(m.group(1), 16), m.group(2).strip()))

KEYS = [("UTCHMMA", r"\bUTCHMMA"), (".2CTA", r"UTCHMMA\.2CTA"), ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"), ("UBLKCP", r"\bUBLKCP"),
        ("UTMALDG", r"\bUTMALDG"), ("UTMASTG", r"\bUTMASTG"), ("UTCBAR", r"\bUTCBAR"), ("SYNCS", r"\bSYNCS"),
        ("MUFU.EX2", r"MUFU\.EX2"), ("FFMA2", r"\bFFMA2"), ("FADD2", r"\bFADD2"), ("REDUX", r"\bREDUX"),
        ("ATOM/RED", r"\b(ATOM|ATOMS|ATOMG|RED)\b"), ("STG", r"\bSTG"), ("LDG", r"\bLDG"), ("BAR", r"\bBAR\.")]
out = [f"SASS census of {os.path.basename(lib)} ({len(kernels)} kernels, sm_100a), `cuobjdump -sass`", ""]
out.append(f"{'kernel':72s} {'instr':>6s} " + " ".join(f"{k:>8s}" for k, _ in KEYS))
tot = collections.Counter()
for name, ins in kernels.items():
    c = {k: sum(1 for _, s in ins if re.search(rx, s)) for k, rx in KEYS}
    tot.update(c)
    short = re.sub(r"^void ", "", demangle(name))
    short = re.sub(r"\(.*", "", short)[:72]
    out.append(f"{short:72s} {len(ins):6d} " + " ".join(f"{c[k]:8d}" for k, _ in KEYS))
out.append(f"{'TOTAL':72s} {sum(len(i) for i in kernels.values()):6d} " + " ".join(f"{tot[k]:8d}" for k, _ in KEYS))
out += ["", "No UTMALDG / UTMASTG by design: every operand tile is stored in HBM as the byte image of its SWIZZLE_128B shared-memory",
        "tile, so ONE flat cp.async.bulk (UBLKCP, the same TMA engine) of 16-64 KB stages it and the UMMA descriptor reads it in",
        "place; a tensor map would re-derive per box what the producer kernel already laid out."]
os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
open(os.path.join(ROOT, "profiles", f"{tag}_sass_census.txt"), "w").write("\n".join(out) + "\n")
print("\n".join(out[:6]))


def hot(pattern, fname, ctx=14, max_regions=4, whole=None):
    for name, ins in kernels.items():
        if not re.search(pattern, demangle(name)):
            continue
        lines = [f"{demangle(name)}", ""]
        if whole:
            a0, a1 = whole(ins)
            lines.append(f"-- {a0:#x} .. {a1:#x}")
            lines += [f"  /*{a:04x}*/  {s}" for a, s in ins if a0 <= a <= a1]
        idx = [k for k, (_, s) in enumerate(ins) if "UTCHMMA" in s]
        regions, last = [], -10 ** 9
        for k in idx:
            if k - last > 2 * ctx:
                regions.append([k, k])
            else:
                regions[-1][1] = k
            last = k
        for r0, r1 in regions[:max_regions]:
            lines.append(f"-- tensor-core issue region {ins[r0][0]:#x} .. {ins[r1][0]:#x}")
            lines += [f"  /*{a:04x}*/  {s}" for a, s in ins[max(0, r0 - ctx): r1 + ctx + 1]]
        open(os.path.join(ROOT, "profiles", f"{tag}_sass_hot_{fname}.txt"), "w").write("\n".join(lines) + "\n")
        return


def attn_sweep(ins):
    mu = [a for a, s in ins if "MUFU.EX2" in s]
    return (mu[1] - 0x200, mu[-1] + 0x300) if len(mu) > 2 else (0, 0)


hot(r"attention_kernel", "attention", whole=attn_sweep)
hot(r"linear_kernel<256, 0, 2>", "linear_split_pair")
hot(r"linear_kernel<192, 2, 2>", "linear_qkv_self_pair")
hot(r"ffn_ln_gelu_kernel", "ffn_ln_gelu")
for pat, fn in ((r"assign_\w*kernel<1>", "assign_pass1"), (r"assign_\w*kernel<2>", "assign_pass2")):
    hot(pat, fn)
