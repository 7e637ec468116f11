"""Static issue-cycle estimate of the attention softmax sweep from the built library's SASS: decodes the control
codes (stall counts), walks the non-partial path from the first S load to the P store and sums the stall counts,
and reports how the MUFU.EX2 instructions are spaced (a MUFU that directly follows a MUFU holds the warp 8 cycles).
usage: sass_sweep_cost.py [lib.so]"""
import re, subprocess, sys, os
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(__file__), "..", "cs566-project-lightglue_b200", "liblg_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout.split("\n")
start = next(i for i, l in enumerate(txt) if "Function :" in l and "attention_kernel" in l)
end = next((i for i in range(start + 1, len(txt)) if "Function :" in txt[i]), len(txt))
lines = txt[start:end]
ins = []
i = 0
while i < len(lines) - 1:
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\* (0x[0-9a-f]+) \*/", lines[i])
    m2 = re.match(r"\s*/\* (0x[0-9a-f]+) \*/", lines[i + 1]) if m else None
    if m and m2:
        hi = int(m2.group(1), 16)
        ins.append((int(m.group(1), 16), m.group(2).strip(), (hi >> 41) & 0xF))
        i += 2
    else:
        i += 1
addr = {a: k for k, (a, _, _) in enumerate(ins)}
# the sweep: the LAST pair of adjacent "LDTM.x32 .. tmem[UR+0]" / "+0x20" loads before the STTM pair at column 0
stt = [k for k, (a, s, _) in enumerate(ins) if s.startswith("STTM.x32") and "+0x" not in s.split("tmem")[1].split("]")[0]]
k_end = stt[-1] + 1
cands = [k for k, (a, s, _) in enumerate(ins) if k < k_end and s.startswith("LDTM.x32") and "+0x" not in s.split("tmem")[1].split("]")[0]]


def walk(k_start):
    k, cyc, n, mufu, clump, clumps = k_start, 0, 0, 0, 0, []
    while k <= k_end:
        a, s, st = ins[k]
        cyc += max(st, 1); n += 1
        if "MUFU.EX2" in s:
            mufu += 1; clump += 1
        else:
            if clump: clumps.append(clump)
            clump = 0
        m = re.match(r"@(!?)P\d+ BRA (0x[0-9a-f]+)", s)
        if m and int(m.group(2), 16) in addr and k < addr[int(m.group(2), 16)] <= k_end:
            k = addr[int(m.group(2), 16)]  # forward branch over the partial-block mask code: taken on full blocks
            continue
        k += 1
    return n, mufu, cyc, clumps


for k_start in cands:
    n, mufu, cyc, clumps = walk(k_start)
    if mufu >= 64:
        print(f"sweep {ins[k_start][0]:#x}..{ins[k_end][0]:#x}: {n} instructions on the full-block path, {mufu} MUFU.EX2, "
              f"static issue cycles (sum of stall counts) = {cyc}")
        print("MUFU run lengths:", clumps)
